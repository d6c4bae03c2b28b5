// find_b200/csrc/step.cu — one host call that enqueues FIND's whole 3-D training step (no host sync, graph-capturable):
//   field forward on the shared template + registration transform     model.py:1009 -> 455-504
//   bidirectional Chamfer of the deformed vertices vs the target cloud losses.py:87-88 (plain branch)
//   colour MSE against per-vertex targets (stand-in for TextureLossGTSpace, losses.py:22-57; optional)
//   loss = w_chamf * chamfer + w_tex * mse                             model.py:1157-1158, opts.py:97-99
//   backward of all of it: d/d(MLP weights), d/d(latent rows), d/d(reg rows)   trainer.py:121
// It only composes the entry points of chamfer.cu / field.cu; the kernels are theirs.
#include "common.cuh"
#include "../../include/find_b200.h"
#include "field.cuh"
#include "chamfer_hooks.cuh"

namespace fb {
namespace step {

static inline size_t align1k(size_t x) { return (x + 1023) & ~(size_t)1023; }

struct Carve {
    size_t field, knn, idx_x, idx_y, grad_verts, gv_scale, grad_col, sums, total;
    size_t field_bytes, knn_bytes, mse_part_bytes;
    // mesh terms (findb200_train_step_mesh): surface samples of the deformed template and the smoothness workspace
    size_t samples, fidx, bary, grad_samples, sample_ws, smooth_ws, smooth2;
    size_t sample_ws_bytes, smooth_ws_bytes;
};
static Carve carve(int B, int V, int P2, int S = 0, int nnz = 0, int F = 0) {
    Carve c{};
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align1k(off + bytes); return o; };
    const int P1 = S > 0 ? S : V;
    c.field_bytes = findb200_field_workspace_bytes(B, V, 1, 1);
    c.knn_bytes = findb200_knn_workspace_bytes(B, P1, P2);
    c.field = take(c.field_bytes);
    c.knn = take(c.knn_bytes);
    c.idx_x = take((size_t)B * P1 * 8);
    c.idx_y = take((size_t)B * P2 * 8);
    c.grad_verts = take((size_t)B * V * 24);       // float [B][V][3], or int64 fixed point on the fused plain path
    c.gv_scale = take((size_t)B * 4);
    c.grad_col = take((size_t)B * V * 12);
    c.mse_part_bytes = (size_t)8 * B * (((P1 > P2 ? P1 : P2) + 511) / 512);
    c.sums = take(1024 + c.mse_part_bytes);     // mse sum | chamfer scalar | block partials of the colour term (fused tail)
    if (S > 0) {
        c.samples = take((size_t)B * S * 12);
        c.fidx = take((size_t)B * S * 8);
        c.bary = take((size_t)B * S * 12);
        c.grad_samples = take((size_t)B * S * 12);
        c.sample_ws_bytes = findb200_sample_workspace_bytes(B, F);
        c.sample_ws = take(c.sample_ws_bytes);
    }
    if (nnz > 0) {
        c.smooth_ws_bytes = findb200_mesh_smooth_workspace_bytes(B, V, nnz);
        c.smooth_ws = take(c.smooth_ws_bytes);
        c.smooth2 = take(64);
    }
    c.total = off;
    return c;
}

// grad_col = 2 w (col - target) / n ; per-block partial sums of (col - target)^2 -> one double atomic per block
__global__ void __launch_bounds__(256) col_mse_kernel(const float* __restrict__ col, const float* __restrict__ tgt, size_t n,
                                                      float w_over_n, float* __restrict__ grad_col, double* sum) {
    fb::pdl_wait_trigger();
    double acc = 0.0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float d = col[i] - tgt[i];
        grad_col[i] = 2.0f * w_over_n * d;
        acc += (double)d * (double)d;
    }
    __shared__ double sh[256];
    sh[threadIdx.x] = acc;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) atomicAdd(sum, sh[0]);
}

__global__ void combine_loss_kernel(const float* chamf, const double* mse_sum, double n, float w_chamf, float w_tex,
                                    const float* smooth2, float w_smooth, float* loss3, float* loss_extra) {
    fb::pdl_wait_trigger();
    if (threadIdx.x || blockIdx.x) return;
    const float c = *chamf;
    const float t = mse_sum ? (float)(*mse_sum / n) : 0.f;
    float sm = 0.f;
    if (smooth2) {                         // MeshSmoothnessLoss, losses.py:97: 0.1 * laplacian + 10 * edge
        sm = 0.1f * smooth2[1] + 10.f * smooth2[0];
        if (loss_extra) { loss_extra[0] = smooth2[0]; loss_extra[1] = smooth2[1]; }
    }
    loss3[0] = w_chamf * c + w_tex * t + w_smooth * sm;
    loss3[1] = c;
    loss3[2] = t;
}

__global__ void bump_counter_kernel(unsigned long long* c) { fb::pdl_wait_trigger(); *c += 1ull; }

}  // namespace step
}  // namespace fb

using namespace fb;
using namespace fb::step;

extern "C" {

size_t findb200_train_step_workspace_bytes(int nfeet, int V, int P2) {
    if (nfeet < 1 || V < 1 || P2 < 0) return 0;
    return carve(nfeet, V, P2).total;
}

size_t findb200_train_step_mesh_workspace_bytes(int nfeet, int V, int P2, int num_samples, int nnz, int F) {
    if (nfeet < 1 || V < 1 || P2 < 0 || num_samples < 0 || nnz < 0 || F < 0) return 0;
    return carve(nfeet, V, P2, num_samples, nnz, F).total;
}

}  // extern "C"

static int run_step(const findb200_field_params* p, const float* template_pos, const float* z_shape, const float* z_pose,
                    const float* z_tex, const float* reg, const float* target, const int64_t* target_len,
                    const float* col_target, int nfeet, int V, int P2, float w_chamf, float w_tex, float* loss3, float* disp,
                    float* col, float* verts, const findb200_field_grads* gw, float* gz_shape, float* gz_pose, float* gz_tex,
                    float* greg, const findb200_mesh_terms* mt, void* workspace, size_t workspace_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(p && template_pos && target && loss3 && disp && col && verts, FINDB200_ERR_ARG,
               "train_step: a required pointer is null");
    FB_REQUIRE(nfeet >= 1 && V >= 1 && P2 >= 1, FINDB200_ERR_SHAPE, "train_step: bad sizes nfeet=%d V=%d P2=%d", nfeet, V, P2);
    const int S = mt ? mt->num_samples : 0;
    const bool smooth = mt && mt->w_smooth != 0.f;
    if (mt) {
        FB_REQUIRE(S >= 0, FINDB200_ERR_SHAPE, "train_step_mesh: num_samples < 0");
        FB_REQUIRE(!(S > 0) || (mt->faces && mt->F >= 1), FINDB200_ERR_ARG, "train_step_mesh: sampling needs the template faces");
        FB_REQUIRE(!smooth || (mt->faces && mt->edges && mt->vf_ptr && mt->vf_face && mt->vf_corner && mt->F >= 1 && mt->E >= 1 &&
                               mt->nnz >= 1), FINDB200_ERR_ARG, "train_step_mesh: smoothness needs faces, edges and the incidence list");
    }
    const Carve c = carve(nfeet, V, P2, S, smooth ? mt->nnz : 0, S > 0 ? mt->F : 0);
    FB_REQUIRE(workspace && workspace_bytes >= c.total, FINDB200_ERR_WORKSPACE,
               "train_step: workspace too small (%zu < %zu bytes)", workspace_bytes, c.total);
    uint8_t* ws = (uint8_t*)workspace;
    int64_t* idx_x = (int64_t*)(ws + c.idx_x);
    int64_t* idx_y = (int64_t*)(ws + c.idx_y);
    float* grad_verts = (float*)(ws + c.grad_verts);
    float* grad_col = (float*)(ws + c.grad_col);
    double* mse_sum = (double*)(ws + c.sums);
    float* chamf = (float*)(ws + c.sums + 64);

    // plain branch: the part of the Chamfer that does not depend on the forward pass (sort of the target clouds, curve order of
    // the template, clearing of the accumulators) runs on a side stream beside the forward kernels (chamfer_hooks.cuh)
    const bool presort = S == 0 && !smooth && fb::chamfer::presort_supported(nfeet, V, P2);
    cudaEvent_t ev_join = nullptr;
    int rc = 0;
    if (presort) {
        cudaStream_t side = nullptr;
        cudaEvent_t ev_fork = nullptr;
        rc = fb::side_stream(0, &side, &ev_fork, &ev_join);
        if (rc) return rc;
        size_t zoff = 0, zbytes = 0;
        fb::field::field_bwd_zero_region(nfeet, V, 1, &zoff, &zbytes);
        FB_CUDA(cudaEventRecord(ev_fork, st));
        FB_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
        rc = fb::chamfer::train_presort(template_pos, target, target_len, nfeet, V, P2, ws + c.grad_verts, (size_t)nfeet * V * 24,
                                        ws + c.field + zoff, zbytes, ws + c.knn, c.knn_bytes, side);
        if (rc) return rc;
        FB_CUDA(cudaEventRecord(ev_join, side));
    }
    rc = findb200_field_forward(p, template_pos, 1, z_shape, z_pose, z_tex, reg, nfeet, V, 1, disp, col, verts,
                                ws + c.field, c.field_bytes, stream);
    if (rc) return rc;
    if (S > 0) {
        // losses.py:67: Chamfer is taken on S area-weighted surface samples of the deformed template
        float* samples = (float*)(ws + c.samples);
        int64_t* fidx = (int64_t*)(ws + c.fidx);
        float* bary = (float*)(ws + c.bary);
        float* grad_samples = (float*)(ws + c.grad_samples);
        rc = findb200_sample_points_forward(verts, 0, mt->faces, 1, nullptr, nfeet, V, mt->F, S, mt->seed, mt->draw_counter, nullptr,
                                            nullptr, samples, fidx, bary, nullptr, 0, 0, nullptr, ws + c.sample_ws,
                                            c.sample_ws_bytes, stream);
        if (rc) return rc;
        if (mt->draw_counter) {
            FB_CUDA(fb::launch_pdl(bump_counter_kernel, 1, 1, 0, st, mt->draw_counter));
            FB_CUDA(cudaGetLastError());
        }
        rc = findb200_chamfer_forward(samples, target, nullptr, target_len, nfeet, S, P2, chamf, nullptr, idx_x, nullptr, idx_y,
                                      ws + c.knn, c.knn_bytes, stream);
        if (rc) return rc;
        rc = findb200_chamfer_backward(samples, target, nullptr, target_len, nfeet, S, P2, idx_x, idx_y, nullptr, w_chamf,
                                       grad_samples, nullptr, stream);
        if (rc) return rc;
        rc = findb200_sample_points_backward(grad_samples, mt->faces, 1, fidx, bary, nfeet, V, mt->F, S, 1.0f, 0, grad_verts, stream);
        if (rc) return rc;
    } else if (!smooth) {
        // plain branch (losses.py:87-88), everything after the kNN in ONE launch: indices, loss3, d/d(verts), colour MSE and its
        // gradient; the sort pass of the kNN also clears grad_verts and the accumulators of the field backward
        const size_t n = (size_t)nfeet * V * 3;
        const bool tex = col_target != nullptr && w_tex != 0.f;
        size_t zoff = 0, zbytes = 0;
        fb::field::field_bwd_zero_region(nfeet, V, 1, &zoff, &zbytes);
        long long* gv_fx = (long long*)(ws + c.grad_verts);
        float* gv_scale = (float*)(ws + c.gv_scale);
        if (presort) FB_CUDA(cudaStreamWaitEvent(st, ev_join, 0));
        fb::chamfer::g_train_presorted = presort;
        rc = findb200_chamfer_train(verts, target, nullptr, target_len, nfeet, V, P2, w_chamf, tex ? col : nullptr, col_target, n,
                                    tex ? w_tex : 0.f, gv_fx, gv_scale, grad_col, idx_x, idx_y, loss3, (double*)(ws + c.sums + 1024),
                                    c.mse_part_bytes, ws + c.field + zoff, zbytes, ws + c.knn, c.knn_bytes, stream);
        fb::chamfer::g_train_presorted = false;
        if (rc) return rc;
        fb::field::g_field_bwd_prezeroed = true;
        fb::field::g_grad_verts_fx = gv_fx;
        fb::field::g_grad_verts_fx_scale = gv_scale;
        rc = findb200_field_backward(p, template_pos, 1, z_shape, z_pose, z_tex, reg, nfeet, V, disp, col, nullptr,
                                     tex ? grad_col : nullptr, nullptr, gw, gz_shape, gz_pose, gz_tex, greg,
                                     ws + c.field, c.field_bytes, stream);
        fb::field::g_field_bwd_prezeroed = false;
        fb::field::g_grad_verts_fx = nullptr;
        fb::field::g_grad_verts_fx_scale = nullptr;
        return rc;
    } else {
        rc = findb200_chamfer_forward(verts, target, nullptr, target_len, nfeet, V, P2, chamf, nullptr, idx_x, nullptr, idx_y,
                                      ws + c.knn, c.knn_bytes, stream);
        if (rc) return rc;
        rc = findb200_chamfer_backward(verts, target, nullptr, target_len, nfeet, V, P2, idx_x, idx_y, nullptr, w_chamf,
                                       grad_verts, nullptr, stream);
        if (rc) return rc;
    }
    float* smooth2 = nullptr;
    if (smooth) {                          // model.py:1032-1033 on the deformed template; weights losses.py:97
        smooth2 = (float*)(ws + c.smooth2);
        rc = findb200_mesh_smooth(verts, nfeet, V, mt->faces, mt->F, mt->edges, mt->E, mt->vf_ptr, mt->vf_face, mt->vf_corner,
                                  mt->nnz, 0.f, 10.f * mt->w_smooth, 0.1f * mt->w_smooth, nullptr, smooth2, grad_verts, 1,
                                  ws + c.smooth_ws, c.smooth_ws_bytes, stream);
        if (rc) return rc;
    }
    const size_t n = (size_t)nfeet * V * 3;
    const bool tex = col_target != nullptr && w_tex != 0.f;
    if (tex) {
        FB_MARK("col_mse", st);
        FB_CUDA(cudaMemsetAsync(mse_sum, 0, 8, st));
        int blocks = (int)((n + 255) / 256);
        if (blocks > 1184) blocks = 1184;
        FB_CUDA(fb::launch_pdl(col_mse_kernel, blocks, 256, 0, st, col, col_target, n, w_tex / (float)n, grad_col, mse_sum));
        FB_CUDA(cudaGetLastError());
    }
    FB_MARK("combine_loss", st);
    FB_CUDA(fb::launch_pdl(combine_loss_kernel, 1, 32, 0, st, chamf, tex ? mse_sum : nullptr, (double)n, w_chamf, tex ? w_tex : 0.f, smooth2,
                                          smooth ? mt->w_smooth : 0.f, loss3, mt ? mt->loss_extra : nullptr));
    FB_CUDA(cudaGetLastError());
    return findb200_field_backward(p, template_pos, 1, z_shape, z_pose, z_tex, reg, nfeet, V, disp, col, nullptr,
                                   tex ? grad_col : nullptr, grad_verts, gw, gz_shape, gz_pose, gz_tex, greg,
                                   ws + c.field, c.field_bytes, stream);
}

extern "C" {

int findb200_train_step(const findb200_field_params* p, const float* template_pos, const float* z_shape,
                        const float* z_pose, const float* z_tex, const float* reg, const float* target,
                        const int64_t* target_len, const float* col_target, int nfeet, int V, int P2, float w_chamf,
                        float w_tex, float* loss3, float* disp, float* col, float* verts,
                        const findb200_field_grads* gw, float* gz_shape, float* gz_pose, float* gz_tex, float* greg,
                        void* workspace, size_t workspace_bytes, void* stream) {
    return run_step(p, template_pos, z_shape, z_pose, z_tex, reg, target, target_len, col_target, nfeet, V, P2, w_chamf, w_tex,
                    loss3, disp, col, verts, gw, gz_shape, gz_pose, gz_tex, greg, nullptr, workspace, workspace_bytes, stream);
}

int findb200_train_step_mesh(const findb200_field_params* p, const float* template_pos, const float* z_shape,
                             const float* z_pose, const float* z_tex, const float* reg, const float* target,
                             const int64_t* target_len, const float* col_target, int nfeet, int V, int P2, float w_chamf,
                             float w_tex, float* loss3, float* disp, float* col, float* verts,
                             const findb200_field_grads* gw, float* gz_shape, float* gz_pose, float* gz_tex, float* greg,
                             const findb200_mesh_terms* mesh_terms, void* workspace, size_t workspace_bytes, void* stream) {
    FB_REQUIRE(mesh_terms, FINDB200_ERR_ARG, "train_step_mesh: mesh_terms is null");
    return run_step(p, template_pos, z_shape, z_pose, z_tex, reg, target, target_len, col_target, nfeet, V, P2, w_chamf, w_tex,
                    loss3, disp, col, verts, gw, gz_shape, gz_pose, gz_tex, greg, mesh_terms, workspace, workspace_bytes, stream);
}

int findb200_train_step_knn_stats(const void* workspace, int nfeet, int V, int P2, unsigned long long* out4_host,
                                  void* stream) {
    FB_REQUIRE(workspace && out4_host && nfeet >= 1 && V >= 1 && P2 >= 1, FINDB200_ERR_ARG, "train_step_knn_stats: bad argument");
    return findb200_knn_stats((const uint8_t*)workspace + carve(nfeet, V, P2).knn, out4_host, stream);
}

}  // extern "C"

// ---- latent-table row gather / gradient scatter (LatentVector.__getitem__, model.py:130-152, and its autograd) ----
namespace fb {
namespace step {
// torch's `self.data[idx]` raises a device-side assert for an index outside the table; so do these kernels
__device__ __forceinline__ long long checked_row(const int64_t* idx, int b, int n_rows, const char* what) {
    const long long r = idx[b];
    if (r < 0 || r >= (long long)n_rows) {
        if (threadIdx.x == 0) printf("find_b200: %s: index %lld of batch entry %d is outside the table [0, %d)\n", what, r, b, n_rows);
        __trap();
    }
    return r;
}
__global__ void gather_rows_kernel(const float* __restrict__ table, int n_rows, int width, const int64_t* __restrict__ idx, int B,
                                   float* __restrict__ out) {
    fb::pdl_wait_trigger();
    const int b = blockIdx.x;
    const float* src = table + (size_t)checked_row(idx, b, n_rows, "gather_rows") * width;
    for (int j = threadIdx.x; j < width; j += blockDim.x) out[(size_t)b * width + j] = src[j];
}
__global__ void scatter_add_rows_kernel(float* __restrict__ grad_table, int n_rows, int width, const int64_t* __restrict__ idx, int B,
                                        const float* __restrict__ g) {
    fb::pdl_wait_trigger();
    const int b = blockIdx.x;
    float* dst = grad_table + (size_t)checked_row(idx, b, n_rows, "scatter_add_rows") * width;
    for (int j = threadIdx.x; j < width; j += blockDim.x) atomicAdd(dst + j, g[(size_t)b * width + j]);   // duplicates accumulate
}
struct GatherJobs { const float* table[8]; float* out[8]; int width[8]; int n_rows[8]; const int64_t* idx[8]; };
__global__ void gather_rows_multi_kernel(const GatherJobs J) {
    fb::pdl_wait_trigger();
    const int b = blockIdx.x, k = blockIdx.y;
    const int width = J.width[k];
    const float* src = J.table[k] + (size_t)checked_row(J.idx[k], b, J.n_rows[k], "gather_rows_multi") * width;
    float* dst = J.out[k] + (size_t)b * width;
    for (int j = threadIdx.x; j < width; j += blockDim.x) dst[j] = src[j];
}
// Dense table gradient of one step, WITHOUT atomics: block (r, k) owns row r of table k and adds, in batch order, the
// gradient rows of every batch entry that indexed it (feet sharing a label, dataset.py:191-201) — what autograd's
// index_put_(accumulate=True) gives after zero_grad, but bit-reproducible.  accumulate != 0 adds to what is there.
struct ScatterJobs { const float* g[8]; float* grad_table[8]; int width[8]; int n_rows[8]; const int64_t* idx[8]; };
__global__ void scatter_rows_multi_kernel(const ScatterJobs J, int B, int accumulate) {
    fb::pdl_wait_trigger();
    const int r = blockIdx.x, k = blockIdx.y;
    if (r >= J.n_rows[k]) return;
    const int width = J.width[k];
    const int64_t* idx = J.idx[k];
    float* dst = J.grad_table[k] + (size_t)r * width;
    for (int j = threadIdx.x; j < width; j += blockDim.x) {
        float acc = accumulate ? dst[j] : 0.f;
        for (int b = 0; b < B; ++b)
            if (idx[b] == (int64_t)r) acc += J.g[k][(size_t)b * width + j];
        dst[j] = acc;
    }
}
}  // namespace step
}  // namespace fb

extern "C" {
int findb200_gather_rows_multi(const findb200_gather_job* jobs, int n_jobs, const int64_t* idx, int B, void* stream) {
    FB_REQUIRE(jobs && n_jobs >= 1 && n_jobs <= 8 && B >= 1, FINDB200_ERR_ARG, "gather_rows_multi: bad argument (1..8 jobs)");
    fb::step::GatherJobs J{};
    for (int k = 0; k < n_jobs; ++k) {
        FB_REQUIRE(jobs[k].table && jobs[k].out && jobs[k].width >= 1 && jobs[k].n_rows >= 1 && (jobs[k].idx || idx), FINDB200_ERR_ARG,
                   "gather_rows_multi: job %d is incomplete", k);
        J.table[k] = jobs[k].table; J.out[k] = jobs[k].out; J.width[k] = jobs[k].width; J.n_rows[k] = jobs[k].n_rows;
        J.idx[k] = jobs[k].idx ? jobs[k].idx : idx;
    }
    FB_MARK("gather_rows", (cudaStream_t)stream);
    FB_CUDA(fb::launch_pdl(fb::step::gather_rows_multi_kernel, dim3(B, n_jobs), 128, 0, (cudaStream_t)stream, J));
    FB_CUDA(cudaGetLastError());
    FB_MARK_END((cudaStream_t)stream);
    return 0;
}
int findb200_scatter_rows_multi(const findb200_scatter_job* jobs, int n_jobs, const int64_t* idx, int B, int accumulate,
                                void* stream) {
    FB_REQUIRE(jobs && n_jobs >= 1 && n_jobs <= 8 && B >= 1, FINDB200_ERR_ARG, "scatter_rows_multi: bad argument (1..8 jobs)");
    fb::step::ScatterJobs J{};
    int max_rows = 1;
    for (int k = 0; k < n_jobs; ++k) {
        FB_REQUIRE(jobs[k].g && jobs[k].grad_table && jobs[k].width >= 1 && jobs[k].n_rows >= 1 && (jobs[k].idx || idx), FINDB200_ERR_ARG,
                   "scatter_rows_multi: job %d is incomplete", k);
        J.g[k] = jobs[k].g; J.grad_table[k] = jobs[k].grad_table; J.width[k] = jobs[k].width; J.n_rows[k] = jobs[k].n_rows;
        J.idx[k] = jobs[k].idx ? jobs[k].idx : idx;
        if (jobs[k].n_rows > max_rows) max_rows = jobs[k].n_rows;
    }
    FB_MARK("scatter_rows", (cudaStream_t)stream);
    FB_CUDA(fb::launch_pdl(fb::step::scatter_rows_multi_kernel, dim3(max_rows, n_jobs), 128, 0, (cudaStream_t)stream, J, B, accumulate));
    FB_CUDA(cudaGetLastError());
    FB_MARK_END((cudaStream_t)stream);
    return 0;
}
int findb200_gather_rows(const float* table, int n_rows, int width, const int64_t* idx, int B, float* out, void* stream) {
    FB_REQUIRE(table && idx && out && n_rows >= 1 && width >= 1 && B >= 1, FINDB200_ERR_ARG, "gather_rows: bad argument");
    FB_MARK("gather_rows", (cudaStream_t)stream);
    FB_CUDA(fb::launch_pdl(fb::step::gather_rows_kernel, B, 128, 0, (cudaStream_t)stream, table, n_rows, width, idx, B, out));
    FB_CUDA(cudaGetLastError());
    FB_MARK_END((cudaStream_t)stream);
    return 0;
}
int findb200_scatter_add_rows(float* grad_table, int n_rows, int width, const int64_t* idx, int B, const float* g,
                              void* stream) {
    FB_REQUIRE(grad_table && idx && g && n_rows >= 1 && width >= 1 && B >= 1, FINDB200_ERR_ARG, "scatter_add_rows: bad argument");
    FB_MARK("scatter_add_rows", (cudaStream_t)stream);
    FB_CUDA(fb::launch_pdl(fb::step::scatter_add_rows_kernel, B, 128, 0, (cudaStream_t)stream, grad_table, n_rows, width, idx, B, g));
    FB_CUDA(cudaGetLastError());
    FB_MARK_END((cudaStream_t)stream);
    return 0;
}
}
