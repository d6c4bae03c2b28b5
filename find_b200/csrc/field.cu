// find_b200/csrc/field.cu — host orchestration + small CUDA-core kernels of the displacement field (C ABI in
// include/find_b200.h).  The tensor-core work is in field_chain.cu (forward / dX chains) and field_dw.cu (dW GEMMs).
//
// Step structure (training, template shared by all feet — the get_meshes call site, model.py:473-477):
//   forward : pack_weights | latent_bias | TRUNK_FWD (once per template tile) | HEAD_FWD (per head, foot, tile)
//   backward: grad_prep (tanh', registration-transform backward, max |g| for the fp16 gradient scale)
//             | HEAD_BWD | TRUNK_BWD | dw_kernel (all weight-gradient GEMMs in one launch) | reduce_dw | finish_small
// With per-foot distinct points (TextureLossGTSpace call site, losses.py:50) the trunk runs per foot as well.
#include "field.cuh"
#include "chamfer_hooks.cuh"
#include <cstdio>
#include <cstdlib>
#include "field_dw.cuh"
#include <math.h>
#include <string.h>

#include "field_chain_params.cuh"

namespace fb {
namespace field {

static inline size_t align1k(size_t x) { return (x + 1023) & ~(size_t)1023; }
static inline int cdiv(int a, int b) { return (a + b - 1) / b; }

bool chain_pairs_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("FINDB200_CHAIN"); on = (e && e[0] == '1') ? 0 : 1; }
    return on != 0;
}

static int g_sms = 0;
static int num_sms() {
    if (!g_sms) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
        if (g_sms <= 0) g_sms = 148;
    }
    return g_sms;
}

// ---- packed weight layout ---------------------------------------------------------------------------------
static PackLayout make_layout() {
    PackLayout L;
    memset(&L, 0, sizeof(L));
    uint32_t off = 0;
    auto put = [&](GemmW& g, int n_chunks, uint32_t chunk_bytes, int N, int k16_last) {
        g.off = off; g.chunk_bytes = chunk_bytes; g.n_chunks = (uint16_t)n_chunks; g.N = (uint16_t)N;
        g.k16_last = (uint16_t)k16_last; g.group = 1;
        off += (uint32_t)n_chunks * chunk_bytes;
    };
    put(L.fwd[P_T0], kEncChunks, kWChunkBytes, 256, 1);          // K = 515 -> 8 chunks + 16 columns of the 9th
    for (int l = P_T1; l <= P_T4; ++l) put(L.fwd[l], 4, kWChunkBytes, 256, 4);
    for (int h = 0; h < 2; ++h) {
        const int b = h ? P_C1 : P_D1;
        for (int l = 0; l < 3; ++l) put(L.fwd[b + l], 4, kWChunkBytes, 256, 4);
        put(L.fwd[b + 3], 4, kWSmallChunkBytes, 16, 4);          // 256 -> 3, N padded to 16
        L.fwd[b + 3].group = 4;                                  // 4 x 2 KB in one slot: one ring round trip, not four
    }
    for (int l = P_T1; l <= P_T4; ++l) put(L.bwd[l], 4, kWChunkBytes, 256, 4);
    for (int h = 0; h < 2; ++h) {
        const int b = h ? P_C1 : P_D1;
        for (int l = 0; l < 3; ++l) put(L.bwd[b + l], 4, kWChunkBytes, 256, 4);
        put(L.bwd[b + 3], 1, kWChunkBytes, 256, 1);              // 3 -> 256: one K = 16 step
    }
    L.total_bytes = off;
    return L;
}

// One thread per 16-B chunk of a weight image.  Image row i, internal k:
//   forward  images: value = W[i][col0 + k]            (i < rows, k < kcols), K-permuted for base.0
//   backward images: value = W[k][col0 + i]            (i < kcols, k < rows)
struct PackJob {
    const float* src;
    int ld, rows, kcols, col0, transpose, perm, img_rows, n_chunks;
    uint32_t dst_off;
};
struct PackParams { PackJob j[2 * P_COUNT]; int n; uint8_t* dst; };

__device__ __forceinline__ void pack_weights_body(const PackParams& P, int job) {
    const PackJob& J = P.j[job];
    const int per_chunk = J.img_rows * 8;
    const int total = per_chunk * J.n_chunks;
    for (int id = blockIdx.x * blockDim.x + threadIdx.x; id < total; id += gridDim.x * blockDim.x) {
        const int c = id / per_chunk, r = (id % per_chunk) / 8, ch = id % 8;
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
            const int k = c * 64 + ch * 8 + e;
            float x = 0.f;
            if (!J.transpose) {
                int kc = J.perm ? enc_ref_col(k) : k;
                if (r < J.rows && kc >= 0 && kc < J.kcols) x = J.src[(size_t)r * J.ld + J.col0 + kc];
            } else {
                if (r < J.kcols && k < J.rows) x = J.src[(size_t)k * J.ld + J.col0 + r];
            }
            v[e] = x;
        }
        uint4 pk;
        __half2 h;
        h = __floats2half2_rn(v[0], v[1]); pk.x = *reinterpret_cast<uint32_t*>(&h);
        h = __floats2half2_rn(v[2], v[3]); pk.y = *reinterpret_cast<uint32_t*>(&h);
        h = __floats2half2_rn(v[4], v[5]); pk.z = *reinterpret_cast<uint32_t*>(&h);
        h = __floats2half2_rn(v[6], v[7]); pk.w = *reinterpret_cast<uint32_t*>(&h);
        uint8_t* d = P.dst + J.dst_off + (size_t)c * (J.img_rows * 128) + (size_t)r * 128 + ((ch ^ (r & 7)) << 4);
        *reinterpret_cast<uint4*>(d) = pk;
    }
}

// biasfoot[head][b][n] = b1[n] + sum_j W1[n][256 + j] * z_b[j]   (the latent columns of mlp_*.0, model.py:428-437)
struct LatentBiasParams {
    const float *Wd, *bd, *Wc, *bc;      // mlp_disp.0 / mlp_col.0
    const float *z_shape, *z_pose, *z_tex;
    int S, Pz, T, B;
    float* out;
};
// one warp per output row n: lanes stride over the latent columns (coalesced weight reads), shuffle-reduce
__device__ __forceinline__ void latent_bias_body(const LatentBiasParams& P, int b, int head, int zslot) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    for (int n = zslot * 8 + wid; n < kW; n += 8 * 8) {
        float acc = 0.f;
        if (head == 0) {
            const int L = P.S + P.Pz;
            const float* w = P.Wd + (size_t)n * (kW + L) + kW;
            for (int j = lane; j < L; j += 32)
                acc = fmaf(w[j], j < P.S ? P.z_shape[(size_t)b * P.S + j] : P.z_pose[(size_t)b * P.Pz + (j - P.S)], acc);
        } else {
            const float* w = P.Wc + (size_t)n * (kW + P.T) + kW;
            for (int j = lane; j < P.T; j += 32) acc = fmaf(w[j], P.z_tex[(size_t)b * P.T + j], acc);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (lane == 0) P.out[((size_t)head * P.B + b) * kW + n] = acc + (head ? P.bc[n] : P.bd[n]);
    }
}

// pack_weights + latent_bias in one launch (they are independent): blockIdx.y < P.n is a pack job; the remaining rows of the
// grid are the (foot, head, row-slot) blocks of latent_bias, numbered row-major
__global__ void __launch_bounds__(256) field_prologue_kernel(const __grid_constant__ PackParams P, const LatentBiasParams L) {
    fb::pdl_wait_trigger();
    if ((int)blockIdx.y < P.n) { pack_weights_body(P, blockIdx.y); return; }
    const int lb = ((int)blockIdx.y - P.n) * gridDim.x + blockIdx.x;
    if (lb >= L.B * 2 * 8) return;
    latent_bias_body(L, lb % L.B, (lb / L.B) & 1, lb / (2 * L.B));
}

__device__ __forceinline__ void euler_xyz_d(const float* e, float (&R)[3][3]) {
    float sa, ca, sb, cb, sc, cc;
    sincosf(e[0], &sa, &ca); sincosf(e[1], &sb, &cb); sincosf(e[2], &sc, &cc);
    R[0][0] = cb * cc;                 R[0][1] = -cb * sc;                R[0][2] = sb;
    R[1][0] = sa * sb * cc + ca * sc;  R[1][1] = -sa * sb * sc + ca * cc; R[1][2] = -sa * cb;
    R[2][0] = -ca * sb * cc + sa * sc; R[2][1] = ca * sb * sc + sa * cc;  R[2][2] = ca * cb;
}

// Backward of the output non-linearities and of the registration transform, per (foot, vertex):
//   g_u = (R g_X) * S  (+ grad_disp),  d(pre-tanh)_disp = g_u * 0.1 * (1 - t^2), t = disp / 0.1
//   d(pre-tanh)_col = grad_col * 0.5 * (1 - t^2), t = 2 (col - avg) - 1
// plus per-foot sums for d(reg) and the global max |d(pre-tanh)| that fixes the fp16 gradient scale.
struct GradPrepParams {
    const float *pos, *disp, *col, *reg, *avg_col, *grad_disp, *grad_col, *grad_verts;
    int B, V, shared;
    float* dovals;       // [2][B][V][4]
    uint32_t* maxabs;
    long long* regsums;  // [B][12] fixed point (kRegFx)
    const long long* grad_verts_fx;    // optional fixed-point form of grad_verts (replaces it when non-null)
    const float* grad_verts_fx_scale;  // [B]
};
__global__ void __launch_bounds__(256) grad_prep_kernel(const GradPrepParams P) {
    fb::pdl_wait_trigger();
    const int b = blockIdx.y;
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    float sums[12];
#pragma unroll
    for (int i = 0; i < 12; ++i) sums[i] = 0.f;
    float m = 0.f;
    if (v < P.V) {
        const size_t o = ((size_t)b * P.V + v) * 3;
        float g0 = 0.f, g1 = 0.f, g2 = 0.f;
        if (P.grad_disp) { g0 = P.grad_disp[o]; g1 = P.grad_disp[o + 1]; g2 = P.grad_disp[o + 2]; }
        if (P.grad_verts || P.grad_verts_fx) {
            float x0, x1, x2;
            if (P.grad_verts_fx) {
                const double sc = (double)P.grad_verts_fx_scale[b];
                x0 = (float)((double)P.grad_verts_fx[o] * sc); x1 = (float)((double)P.grad_verts_fx[o + 1] * sc);
                x2 = (float)((double)P.grad_verts_fx[o + 2] * sc);
            } else { x0 = P.grad_verts[o]; x1 = P.grad_verts[o + 1]; x2 = P.grad_verts[o + 2]; }
            if (P.reg) {
                const float* rg = P.reg + (size_t)b * 9;
                float R[3][3];
                euler_xyz_d(rg + 3, R);
                const float* pp = P.pos + ((size_t)(P.shared ? 0 : b) * P.V + v) * 3;
                const float u[3] = {pp[0] + P.disp[o], pp[1] + P.disp[o + 1], pp[2] + P.disp[o + 2]};
                const float gx[3] = {x0, x1, x2};
                sums[0] = x0; sums[1] = x1; sums[2] = x2;
#pragma unroll
                for (int i = 0; i < 3; ++i)
#pragma unroll
                    for (int j = 0; j < 3; ++j) sums[3 + i * 3 + j] = u[i] * gx[j];
                g0 += rg[6] * (R[0][0] * x0 + R[0][1] * x1 + R[0][2] * x2);
                g1 += rg[7] * (R[1][0] * x0 + R[1][1] * x1 + R[1][2] * x2);
                g2 += rg[8] * (R[2][0] * x0 + R[2][1] * x1 + R[2][2] * x2);
            } else {
                g0 += x0; g1 += x1; g2 += x2;
            }
        }
        const float t0 = 10.f * P.disp[o], t1 = 10.f * P.disp[o + 1], t2 = 10.f * P.disp[o + 2];
        float4 dd = make_float4(g0 * 0.1f * (1.f - t0 * t0), g1 * 0.1f * (1.f - t1 * t1), g2 * 0.1f * (1.f - t2 * t2), 0.f);
        *reinterpret_cast<float4*>(P.dovals + ((size_t)b * P.V + v) * 4) = dd;
        m = fmaxf(fmaxf(fabsf(dd.x), fabsf(dd.y)), fabsf(dd.z));
        float4 dc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (P.grad_col) {
            const float a0 = P.avg_col ? P.avg_col[0] : 0.f, a1 = P.avg_col ? P.avg_col[1] : 0.f, a2 = P.avg_col ? P.avg_col[2] : 0.f;
            const float c0 = 2.f * (P.col[o] - a0) - 1.f, c1 = 2.f * (P.col[o + 1] - a1) - 1.f, c2 = 2.f * (P.col[o + 2] - a2) - 1.f;
            dc = make_float4(P.grad_col[o] * 0.5f * (1.f - c0 * c0), P.grad_col[o + 1] * 0.5f * (1.f - c1 * c1),
                             P.grad_col[o + 2] * 0.5f * (1.f - c2 * c2), 0.f);
            m = fmaxf(m, fmaxf(fmaxf(fabsf(dc.x), fabsf(dc.y)), fabsf(dc.z)));
        }
        *reinterpret_cast<float4*>(P.dovals + (((size_t)P.B + b) * P.V + v) * 4) = dc;
    }
    // block reductions
    __shared__ float sh[12][8];
    __shared__ float shm[8];
    const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    const bool have_gv = P.grad_verts || P.grad_verts_fx;
    if (P.reg && have_gv) {
#pragma unroll
        for (int i = 0; i < 12; ++i) {
            float s = sums[i];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if (lane == 0) sh[i][wp] = s;
        }
    }
    if (lane == 0) shm[wp] = m;
    __syncthreads();
    if (threadIdx.x < 12 && P.reg && have_gv) {
        float s = 0.f;
        for (int w = 0; w < 8; ++w) s += sh[threadIdx.x][w];
        fx_add(P.regsums + (size_t)b * 12 + threadIdx.x, s, kRegFx);
    }
    if (threadIdx.x == 0) {
        float mm = 0.f;
        for (int w = 0; w < 8; ++w) mm = fmaxf(mm, shm[w]);
        if (mm == mm && mm < 3.0e38f) atomicMax(P.maxabs, __float_as_uint(mm));   // non-negative floats order like their bits
    }
}

__device__ __forceinline__ float inv_grad_scale(uint32_t maxabs_bits) {
    const float m = __uint_as_float(maxabs_bits);
    if (!(m > 0.f) || !(m < 3.0e38f)) return 1.0f;
    int e;
    frexpf(m, &e);
    return ldexpf(1.0f, e - 4);
}

// dst[r][dst_col(c)] = inv_scale * sum_s partial[s][r][src_col0 + c]      (fixed order => deterministic)
struct ReduceJob {
    float* dst;
    const float* partial;
    int dst_ld, rows, cols, dst_col0, perm_kbase, ncta, src_col0;   // perm_kbase >= 0: dst col = enc_ref_col(kbase + c)
};
constexpr int kMaxReduceJobs = 20;
struct ReduceParams { ReduceJob j[kMaxReduceJobs]; int n; const uint32_t* maxabs; };

__device__ __forceinline__ void reduce_dw_body(const ReduceParams& P, int job) {
    const ReduceJob& J = P.j[job];
    if (!J.dst) return;
    const float inv = inv_grad_scale(*P.maxabs);
    if ((J.cols & 3) == 0 && (J.src_col0 & 3) == 0) {
        // four columns per thread: one 16-byte load per partial, all of them independent (the launch is a pure HBM read)
        const int c4n = J.cols >> 2, total4 = J.rows * c4n;
        for (int id = blockIdx.x * blockDim.x + threadIdx.x; id < total4; id += gridDim.x * blockDim.x) {
            const int r = id / c4n, c = (id % c4n) << 2;
            const float4* p = reinterpret_cast<const float4*>(J.partial + (size_t)r * kW + J.src_col0 + c);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 8
            for (int s = 0; s < J.ncta; ++s) {
                const float4 v = __ldg(p + (size_t)s * (kW * kW / 4));
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            const float a[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                int dc = J.dst_col0 + c + e;
                if (J.perm_kbase >= 0) {
                    dc = enc_ref_col(J.perm_kbase + c + e);
                    if (dc < 0) continue;
                }
                J.dst[(size_t)r * J.dst_ld + dc] = a[e] * inv;
            }
        }
        return;
    }
    const int total = J.rows * J.cols;
    for (int id = blockIdx.x * blockDim.x + threadIdx.x; id < total; id += gridDim.x * blockDim.x) {
        const int r = id / J.cols, c = id % J.cols;
        float acc = 0.f;
        const float* p = J.partial + (size_t)r * kW + J.src_col0 + c;
        for (int s = 0; s < J.ncta; ++s) acc += p[(size_t)s * kW * kW];
        int dc = J.dst_col0 + c;
        if (J.perm_kbase >= 0) {
            dc = enc_ref_col(J.perm_kbase + c);
            if (dc < 0) continue;
        }
        J.dst[(size_t)r * J.dst_ld + dc] = acc * inv;
    }
}

// Everything that hangs off the column sums: bias gradients, latent columns of mlp_*.0.weight, latent-code
// gradients, registration gradients.  grid.x = blocks of 256 threads, grid.y = section.
struct FinishParams {
    const uint32_t* maxabs;
    const long long* colsum;      // see Workspace::colsum (fixed point, kColsumFx)
    const long long* regsums;     // [B][12] (fixed point, kRegFx)
    const float *Wd, *Wc;         // mlp_disp.0.weight / mlp_col.0.weight (for the latent-code gradients)
    const float *z_shape, *z_pose, *z_tex, *reg;
    int B, S, Pz, T, heads_mask, frozen, have_reg;
    float* gW[P_COUNT];
    float* gb[P_COUNT];
    float *gz_shape, *gz_pose, *gz_tex, *greg;
};
// colsum layout (floats)
__host__ __device__ inline size_t cs_foot(int head, int B) { return (size_t)head * B * kW; }                 // [2][B][256]
__host__ __device__ inline size_t cs_head(int head, int l, int B) { return (size_t)2 * B * kW + ((size_t)head * 2 + l) * kW; }   // l = 0: layer 2, 1: layer 3
__host__ __device__ inline size_t cs_out(int head, int B) { return (size_t)2 * B * kW + 4 * kW + (size_t)head * 16; }
__host__ __device__ inline size_t cs_trunk(int l, int B) { return (size_t)2 * B * kW + 4 * kW + 32 + (size_t)l * kW; }
__host__ __device__ inline size_t cs_total(int B) { return (size_t)2 * B * kW + 4 * kW + 32 + 5 * kW; }

__device__ __forceinline__ void finish_small_body(const FinishParams& P, int sec) {
    const float inv = inv_grad_scale(*P.maxabs);
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (sec == 0) {
        // bias gradients (13 vectors) : id = param * 256 + n
        if (P.frozen) return;
        const int prm = id / kW, n = id % kW;
        if (prm >= P_COUNT || !P.gb[prm]) return;
        float v = 0.f;
        if (prm <= P_T4) v = fx_get(P.colsum + cs_trunk(prm, P.B) + n, kColsumFx);
        else {
            const int head = prm >= P_C1, l = prm - (head ? P_C1 : P_D1);
            if (l == 3 && n >= 3) return;
            if (!(P.heads_mask & (1 << head))) v = 0.f;
            else if (l == 0) { for (int b = 0; b < P.B; ++b) v += fx_get(P.colsum + cs_foot(head, P.B) + (size_t)b * kW + n, kColsumFx); }
            else if (l == 3) v = fx_get(P.colsum + cs_out(head, P.B) + n, kColsumFx);
            else v = fx_get(P.colsum + cs_head(head, l - 1, P.B) + n, kColsumFx);
        }
        P.gb[prm][n] = v * inv;
    } else if (sec == 1 || sec == 2) {
        // latent columns of mlp_disp.0.weight (sec 1) / mlp_col.0.weight (sec 2): dW[n][256 + j] = sum_b s_b[n] z_b[j]
        if (P.frozen) return;
        const int head = sec - 1;
        const int L = head ? P.T : P.S + P.Pz;
        float* g = P.gW[head ? P_C1 : P_D1];
        if (!g || id >= kW * L) return;
        const int n = id / L, j = id % L;
        float v = 0.f;
        if (P.heads_mask & (1 << head)) {
            for (int b = 0; b < P.B; ++b) {
                float z;
                if (head) z = P.z_tex[(size_t)b * P.T + j];
                else z = j < P.S ? P.z_shape[(size_t)b * P.S + j] : P.z_pose[(size_t)b * P.Pz + (j - P.S)];
                v = fmaf(fx_get(P.colsum + cs_foot(head, P.B) + (size_t)b * kW + n, kColsumFx), z, v);
            }
        }
        g[(size_t)n * (kW + L) + kW + j] = v * inv;
    } else if (sec == 3) {
        // latent-code gradients: gz[b][j] = sum_n W1[n][256 + j] s_b[n].  A block owns 32 consecutive (b, j) pairs; its
        // 8 warps split the 256 rows, so every thread issues 32 independent coalesced loads, then a shared-memory fold.
        __shared__ float part[8][33];
        const int Ld = P.S + P.Pz, Lt = P.T;
        const int per = Ld + Lt;
        const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
        const int pid = blockIdx.x * 32 + lane;
        const bool live = per > 0 && pid < P.B * per;
        const int b = live ? pid / per : 0, j = live ? pid % per : 0;
        const int head = j >= Ld;
        const int jj = head ? j - Ld : j;
        float v = 0.f;
        if (live && (P.heads_mask & (1 << head))) {
            const float* W = head ? P.Wc : P.Wd;
            const int ld = kW + (head ? Lt : Ld);
            const long long* s = P.colsum + cs_foot(head, P.B) + (size_t)b * kW + wq * 32;
            const float* Wj = W + (size_t)(wq * 32) * ld + kW + jj;
            float v0 = 0.f, v1 = 0.f, v2 = 0.f, v3 = 0.f;
#pragma unroll
            for (int n = 0; n < 32; n += 4) {
                v0 = fmaf(Wj[(size_t)n * ld], fx_get(s + n, kColsumFx), v0);
                v1 = fmaf(Wj[(size_t)(n + 1) * ld], fx_get(s + n + 1, kColsumFx), v1);
                v2 = fmaf(Wj[(size_t)(n + 2) * ld], fx_get(s + n + 2, kColsumFx), v2);
                v3 = fmaf(Wj[(size_t)(n + 3) * ld], fx_get(s + n + 3, kColsumFx), v3);
            }
            v = (v0 + v1) + (v2 + v3);
        }
        part[wq][lane] = v;
        __syncthreads();
        if (wq == 0 && live) {
            float t = 0.f;
#pragma unroll
            for (int q = 0; q < 8; ++q) t += part[q][lane];
            t *= inv;
            if (head) { if (P.gz_tex) P.gz_tex[(size_t)b * P.T + jj] = t; }
            else if (jj < P.S) { if (P.gz_shape) P.gz_shape[(size_t)b * P.S + jj] = t; }
            else { if (P.gz_pose) P.gz_pose[(size_t)b * P.Pz + (jj - P.S)] = t; }
        }
    } else {
        // registration gradients (reg = [t(3), euler(3), S(3)], model.py:481-488); not scaled (fp32 path)
        if (!P.greg || id >= P.B) return;
        const int b = id;
        float* g = P.greg + (size_t)b * 9;
        if (!P.have_reg) { for (int i = 0; i < 9; ++i) g[i] = 0.f; return; }
        float s[12];
        for (int i = 0; i < 12; ++i) s[i] = fx_get(P.regsums + (size_t)b * 12 + i, kRegFx);
        const float* rg = P.reg + (size_t)b * 9;
        float sa, ca, sb, cb, sc, cc;
        sincosf(rg[3], &sa, &ca); sincosf(rg[4], &sb, &cb); sincosf(rg[5], &sc, &cc);
        float R[3][3];
        euler_xyz_d(rg + 3, R);
        const float dRa[3][3] = {{0.f, 0.f, 0.f}, {-R[2][0], -R[2][1], -R[2][2]}, {R[1][0], R[1][1], R[1][2]}};
        const float dRb[3][3] = {{-sb * cc, sb * sc, cb}, {sa * cb * cc, -sa * cb * sc, sa * sb}, {-ca * cb * cc, ca * cb * sc, -ca * sb}};
        const float dRc[3][3] = {{R[0][1], -R[0][0], 0.f}, {R[1][1], -R[1][0], 0.f}, {R[2][1], -R[2][0], 0.f}};
        float ga = 0.f, gbb = 0.f, gc = 0.f, gS[3];
        for (int i = 0; i < 3; ++i) {
            float acc = 0.f;
            for (int j = 0; j < 3; ++j) {
                const float Q = s[3 + i * 3 + j];
                acc += R[i][j] * Q;
                const float M = rg[6 + i] * Q;
                ga += M * dRa[i][j]; gbb += M * dRb[i][j]; gc += M * dRc[i][j];
            }
            gS[i] = acc;
        }
        g[0] = s[0]; g[1] = s[1]; g[2] = s[2];
        g[3] = ga; g[4] = gbb; g[5] = gc;
        g[6] = gS[0]; g[7] = gS[1]; g[8] = gS[2];
    }
}

// Shared template: the trunk sees the SUM over feet of the heads' layer-1 gradients (the template rows are the same
// vertices for every foot).  Elementwise over tile images: out[hd][tile] = sum_f dZ1[hd][f][tile], fp32 accumulate.
// grid (trunk_tiles * 16, heads), block 256; one uint4 (8 halves) per thread.
__global__ void __launch_bounds__(256) sum_feet_kernel(const uint8_t* __restrict__ hdz, uint8_t* __restrict__ hsum, int B,
                                                       int tpf2, int heads_mask) {
    fb::pdl_wait_trigger();
    const int hd = (heads_mask == 2) ? 1 : blockIdx.y;
    const size_t j = (size_t)blockIdx.x * 256 + threadIdx.x;            // uint4 index inside [tpf2 tiles]
    const size_t per_foot = (size_t)tpf2 * (kTileBytes / 16);
    const uint4* src = reinterpret_cast<const uint4*>(hdz + (size_t)hd * 3 * B * tpf2 * kTileBytes) + j;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int f0 = 0; f0 < B; f0 += 8) {
        uint4 v[8];
#pragma unroll
        for (int f = 0; f < 8; ++f)
            v[f] = (f0 + f < B) ? __ldg(src + (size_t)(f0 + f) * per_foot) : make_uint4(0, 0, 0, 0);
#pragma unroll
        for (int f = 0; f < 8; ++f) {
            const __half2* h = reinterpret_cast<const __half2*>(&v[f]);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 x = __half22float2(h[i]);
                acc[2 * i] += x.x;
                acc[2 * i + 1] += x.y;
            }
        }
    }
    uint4 o;
    __half2* oh = reinterpret_cast<__half2*>(&o);
#pragma unroll
    for (int i = 0; i < 4; ++i) oh[i] = __floats2half2_rn(acc[2 * i], acc[2 * i + 1]);
    reinterpret_cast<uint4*>(hsum + (size_t)hd * tpf2 * kTileBytes)[j] = o;
}

// reduce_dw + finish_small in one launch: blockIdx.y < R.n is a reduce job, the rest are the sections of finish_small
__global__ void __launch_bounds__(256) reduce_finish_kernel(const __grid_constant__ ReduceParams R, const __grid_constant__ FinishParams F) {
    fb::pdl_wait_trigger();
    if ((int)blockIdx.y < R.n) reduce_dw_body(R, blockIdx.y);
    else finish_small_body(F, (int)blockIdx.y - R.n);
}

// ---- workspace --------------------------------------------------------------------------------------------
static Workspace carve(int B, int V, int shared, int for_backward) {
    Workspace w;
    memset(&w, 0, sizeof(w));
    const PackLayout L = make_layout();
    w.tpf = cdiv(V > 0 ? V : 1, kTile);
    w.tpf2 = (w.tpf + 1) & ~1;
    w.trunk_feet = shared ? 1 : B;
    w.trunk_tiles = w.trunk_feet * w.tpf2;
    const size_t head_tiles = (size_t)B * w.tpf2;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t o = off; off = align1k(off + bytes); return o; };
    w.wpack = take(L.total_bytes);
    w.biasfoot = take((size_t)2 * B * kW * 4);
    w.tact = take((size_t)5 * w.trunk_tiles * kTileBytes);
    w.scal = take(1024);
    if (for_backward) {
        // scal | regsums | colsum are adjacent: the backward zeroes them with ONE memset node
        w.regsums = take((size_t)B * 12 * 8);
        w.colsum = take(cs_total(B) * 8);
        w.enc = take((size_t)w.trunk_tiles * kEncTileBytes);
        w.hact = take((size_t)2 * 3 * head_tiles * kTileBytes);
        w.hdz = take((size_t)2 * 3 * head_tiles * kTileBytes);
        w.hsum = take(shared ? (size_t)2 * w.trunk_tiles * kTileBytes : 0);
        w.hmask = take((size_t)2 * 3 * head_tiles * kMaskTileBytes);
        w.tmask = take((size_t)5 * w.trunk_tiles * kMaskTileBytes);
        w.doimg = take((size_t)2 * head_tiles * kImgBytes);
        w.tdz = take((size_t)5 * w.trunk_tiles * kTileBytes);
        w.dovals = take((size_t)2 * B * V * 16);
        w.n_partials = kMaxDwCtas;
        w.partial = take((size_t)w.n_partials * kW * kW * 4);
    }
    w.total = off;
    return w;
}

struct Sizes { int S, Pz, T; };

static int check_common(const findb200_field_params* p, int B, int V) {
    FB_REQUIRE(p != nullptr, FINDB200_ERR_ARG, "field: params is null");
    FB_REQUIRE(B >= 1 && V >= 1, FINDB200_ERR_SHAPE, "field: need nfeet >= 1 and V >= 1 (got %d, %d)", B, V);
    FB_REQUIRE((long long)B * V < (1ll << 31) / 16, FINDB200_ERR_SHAPE, "field: nfeet * V too large");
    for (int i = 0; i < P_COUNT; ++i)
        FB_REQUIRE(p->weight[i] && p->bias[i], FINDB200_ERR_ARG, "field: weight/bias %d is null", i);
    FB_REQUIRE(p->fourier_B != nullptr, FINDB200_ERR_ARG, "field: fourier_B is null");
    FB_REQUIRE(p->shapevec_size >= 0 && p->posevec_size >= 0 && p->texvec_size >= 0, FINDB200_ERR_SHAPE,
               "field: negative latent size");
    return 0;
}

static void fill_chain_common(ChainParams& C, const findb200_field_params* p, const Workspace& w, uint8_t* ws, int B, int V,
                              int shared) {
    memset(&C, 0, sizeof(C));
    C.wpack = ws + w.wpack;
    C.lay = make_layout();
    C.B = B; C.V = V; C.tpf = w.tpf; C.tpf2 = w.tpf2; C.trunk_feet = w.trunk_feet; C.trunk_tiles = w.trunk_tiles;
    C.shared = shared;
    C.Bmat = p->fourier_B;
    for (int i = 0; i < P_COUNT; ++i) C.bias[i] = p->bias[i];
    C.biasfoot = (const float*)(ws + w.biasfoot);
    C.enc = ws + w.enc; C.tact = ws + w.tact; C.hact = ws + w.hact; C.hdz = ws + w.hdz; C.doimg = ws + w.doimg;
    C.tdz = ws + w.tdz;
    C.hmask = (uint32_t*)(ws + w.hmask); C.tmask = (uint32_t*)(ws + w.tmask);
    C.avg_col = p->avg_col;
    C.dovals = (const float*)(ws + w.dovals);
    C.maxabs = (const uint32_t*)(ws + w.scal);
    C.dbg = g_chain_dbg;
    C.dbg_flags = g_chain_dbg_flags;
}

}  // namespace field
}  // namespace fb
namespace fb {
namespace field {
float g_dw_early_weight = -1.f;
thread_local bool g_field_bwd_prezeroed = false;
thread_local const long long* g_grad_verts_fx = nullptr;
thread_local const float* g_grad_verts_fx_scale = nullptr;
void field_bwd_zero_region(int nfeet, int V, int pos_shared, size_t* offset, size_t* bytes) {
    const Workspace w = carve(nfeet, V, pos_shared, 1);
    *offset = w.scal;
    *bytes = (w.colsum - w.scal) + cs_total(nfeet) * 8;
}

static int run_pack(const findb200_field_params* p, const Workspace& w, uint8_t* ws, const LatentBiasParams& LB, cudaStream_t st) {
    const PackLayout L = make_layout();
    PackParams PP;
    memset(&PP, 0, sizeof(PP));
    PP.dst = ws + w.wpack;
    const int in_dims[P_COUNT] = {kEnc, kW, kW, kW, kW, kW + p->shapevec_size + p->posevec_size, kW, kW, kW,
                                  kW + p->texvec_size, kW, kW, kW};
    const int out_dims[P_COUNT] = {kW, kW, kW, kW, kW, kW, kW, kW, 3, kW, kW, kW, 3};
    int n = 0;
    for (int i = 0; i < P_COUNT; ++i) {
        PackJob& J = PP.j[n++];
        J.src = p->weight[i]; J.ld = in_dims[i]; J.rows = out_dims[i]; J.kcols = (i == P_T0) ? kEnc : kW; J.col0 = 0;
        J.transpose = 0; J.perm = (i == P_T0); J.img_rows = L.fwd[i].N; J.n_chunks = L.fwd[i].n_chunks; J.dst_off = L.fwd[i].off;
    }
    for (int i = 0; i < P_COUNT; ++i) {
        if (i == P_T0) continue;
        PackJob& J = PP.j[n++];
        J.src = p->weight[i]; J.ld = in_dims[i]; J.rows = out_dims[i]; J.kcols = kW; J.col0 = 0;
        J.transpose = 1; J.perm = 0; J.img_rows = 256; J.n_chunks = L.bwd[i].n_chunks; J.dst_off = L.bwd[i].off;
    }
    PP.n = n;
    FB_MARK("field_prologue", st);
    const int lb_rows = cdiv(LB.B * 2 * 8, 36);
    FB_CUDA(fb::launch_pdl(field_prologue_kernel, dim3(36, n + lb_rows), 256, 0, st, PP, LB));
    FB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace field
}  // namespace fb

using namespace fb;
using namespace fb::field;

extern "C" {

size_t findb200_field_workspace_bytes(int nfeet, int V, int pos_shared, int for_backward) {
    if (nfeet < 1 || V < 1) return 0;
    return carve(nfeet, V, pos_shared, for_backward).total;
}

int findb200_field_forward(const findb200_field_params* p, const float* pos, int pos_shared, const float* z_shape,
                           const float* z_pose, const float* z_tex, const float* reg, int nfeet, int V,
                           int save_for_backward, float* disp, float* col, float* verts, void* ws_, size_t ws_bytes,
                           void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int B = nfeet;
    int rc = check_common(p, B, V);
    if (rc) return rc;
    FB_REQUIRE(pos && disp, FINDB200_ERR_ARG, "field_forward: pos/disp is null");
    FB_REQUIRE((p->shapevec_size == 0 || z_shape) && (p->posevec_size == 0 || z_pose) && (p->texvec_size == 0 || z_tex),
               FINDB200_ERR_ARG, "field_forward: a latent block has non-zero size but a null pointer");
    const Workspace w = carve(B, V, pos_shared, save_for_backward);
    FB_REQUIRE(ws_ && ws_bytes >= w.total, FINDB200_ERR_WORKSPACE, "field_forward: workspace too small (%zu < %zu bytes)",
               ws_bytes, w.total);
    uint8_t* ws = (uint8_t*)ws_;
    const int sms = num_sms();

    LatentBiasParams LB{p->weight[P_D1], p->bias[P_D1], p->weight[P_C1], p->bias[P_C1], z_shape, z_pose, z_tex,
                        p->shapevec_size, p->posevec_size, p->texvec_size, B, (float*)(ws + w.biasfoot)};
    rc = run_pack(p, w, ws, LB, st);
    if (rc) return rc;

    ChainParams C;
    fill_chain_common(C, p, w, ws, B, V, pos_shared);
    C.save = save_for_backward;
    C.pos = pos;
    C.disp = disp; C.col = col; C.verts = verts; C.reg = reg;
    // trunk: one tile per CTA while that still fills the machine poorly (latency), two when there are many tiles
    C.enc_cached = (save_for_backward && (p->flags & FINDB200_FIELD_ENC_CACHED)) ? 1 : 0;
    if (!(save_for_backward && (p->flags & FINDB200_FIELD_TRUNK_CACHED))) {
        const int nt = (w.trunk_tiles <= 2 * sms) ? 1 : 2;
        C.n_items = w.trunk_tiles / nt;
        rc = chain_pairs_enabled() ? launch_chain2(0, C, sms, st) : launch_chain(0, nt, C, sms, st);
        if (rc) return rc;
    }
    {
        C.heads_mask = col ? 3 : 1;
        C.n_heads = col ? 2 : 1;
        const int nt = 2;
        C.n_items = C.n_heads * B * (w.tpf2 / nt);
        rc = chain_pairs_enabled() ? launch_chain2(1, C, sms, st) : launch_chain(1, nt, C, sms, st);
        if (rc) return rc;
    }
    FB_MARK_END(st);
    return 0;
}

int findb200_field_backward(const findb200_field_params* p, const float* pos, int pos_shared, const float* z_shape,
                            const float* z_pose, const float* z_tex, const float* reg, int nfeet, int V,
                            const float* disp, const float* col, const float* grad_disp, const float* grad_col,
                            const float* grad_verts, const findb200_field_grads* gw, float* gz_shape, float* gz_pose,
                            float* gz_tex, float* greg, void* ws_, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const int B = nfeet;
    int rc = check_common(p, B, V);
    if (rc) return rc;
    FB_REQUIRE(pos && disp, FINDB200_ERR_ARG, "field_backward: pos/disp is null");
    FB_REQUIRE(!grad_col || col, FINDB200_ERR_ARG, "field_backward: grad_col given without col");
    const Workspace w = carve(B, V, pos_shared, 1);
    FB_REQUIRE(ws_ && ws_bytes >= w.total, FINDB200_ERR_WORKSPACE, "field_backward: workspace too small (%zu < %zu bytes)",
               ws_bytes, w.total);
    uint8_t* ws = (uint8_t*)ws_;
    const int sms = num_sms();
    const int frozen = gw == nullptr;
    const int heads_mask = 1 | (grad_col ? 2 : 0);
    const int n_heads = grad_col ? 2 : 1;
    const size_t head_tiles = (size_t)B * w.tpf2;

    // scalars, per-foot sums, column sums, d(pre-tanh) images start from zero
    // one memset: global max | per-foot reg sums | column sums (adjacent in the workspace).  The d(pre-tanh) images need no
    // clearing: head_bwd rewrites columns 0..15 of every row each step, and whatever sits in columns 16..63 only reaches
    // rows >= 16 of the 256 -> 3 layers' dW partial, which reduce_dw never reads (J.rows = 3).
    if (!g_field_bwd_prezeroed) {
        FB_MARK("bwd_memsets", st);
        FB_CUDA(cudaMemsetAsync(ws + w.scal, 0, (w.colsum - w.scal) + cs_total(B) * 8, st));
    }

    GradPrepParams GP{pos, disp, col, reg, p->avg_col, grad_disp, grad_col, grad_verts, B, V, pos_shared,
                      (float*)(ws + w.dovals), (uint32_t*)(ws + w.scal), (long long*)(ws + w.regsums), g_grad_verts_fx,
                      g_grad_verts_fx_scale};
    FB_MARK("grad_prep", st);
    FB_CUDA(fb::launch_pdl(grad_prep_kernel, dim3(cdiv(V, 256), B), 256, 0, st, GP));
    FB_CUDA(cudaGetLastError());

    ChainParams C;
    fill_chain_common(C, p, w, ws, B, V, pos_shared);
    C.pos = pos; C.reg = reg;
    C.heads_mask = heads_mask; C.n_heads = n_heads;
    {
        const int nt = 2;
        C.n_items = n_heads * B * (w.tpf2 / nt);
        rc = chain_pairs_enabled() ? launch_chain2(2, C, sms, st) : launch_chain(2, nt, C, sms, st);
        if (rc) return rc;
    }
    long long* colsum = (long long*)(ws + w.colsum);
    float* partial = (float*)(ws + w.partial);

    DwParams D;
    memset(&D, 0, sizeof(D));
    int ng = 0;
    int gemm_param[kMaxDwGemms];     // which parameter the GEMM feeds, and which K block of base.0
    int gemm_kblock[kMaxDwGemms];
    auto add = [&](const uint8_t* A, uint32_t atb, int a_imgs, const uint8_t* Bm, uint32_t btb, uint32_t boff, int b_imgs,
                   int mh, int N, int feet, int afs, int bfs, long long* cs, int csmode, int cscols, int prm, int kblock) {
        DwGemm& g = D.g[ng];
        g.A = A; g.Bm = Bm; g.a_tile_bytes = atb; g.b_tile_bytes = btb; g.b_img_off = boff; g.a_imgs = a_imgs; g.b_imgs = b_imgs;
        g.mhalves = mh; g.N = N; g.feet = feet; g.tpf = w.tpf; g.a_foot_stride = afs; g.b_foot_stride = bfs;
        g.colsum = cs; g.colsum_mode = csmode; g.colsum_cols = cscols;
        gemm_param[ng] = prm; gemm_kblock[ng] = kblock;
        ++ng;
    };
    for (int head = 0; head < 2; ++head) {
        if (!(heads_mask & (1 << head))) continue;
        const int base = head ? P_C1 : P_D1;
        uint8_t* hact = ws + w.hact + (size_t)head * 3 * head_tiles * kTileBytes;
        uint8_t* hdz = ws + w.hdz + (size_t)head * 3 * head_tiles * kTileBytes;
        uint8_t* doimg = ws + w.doimg + (size_t)head * head_tiles * kImgBytes;
        const uint8_t* h5 = ws + w.tact + (size_t)4 * w.trunk_tiles * kTileBytes;
        if (frozen) {
            // only the per-foot column sums of dZ1 are needed (latent-code gradients): stream A alone
            add(hdz, kTileBytes, 4, h5, kTileBytes, 0, 0, 0, 256, B, w.tpf2, pos_shared ? 0 : w.tpf2,
                colsum + cs_foot(head, B), 2, 256, -1, 0);
            continue;
        }
        // layer 1: dZ1^T h5 (+ per-foot column sums)
        add(hdz, kTileBytes, 4, h5, kTileBytes, 0, 4, 2, 256, B, w.tpf2, pos_shared ? 0 : w.tpf2, colsum + cs_foot(head, B), 2, 256,
            base, 0);
        // layers 2, 3: dZ_l^T a_{l-1}
        for (int l = 1; l <= 2; ++l)
            add(hdz + (size_t)l * head_tiles * kTileBytes, kTileBytes, 4, hact + (size_t)(l - 1) * head_tiles * kTileBytes,
                kTileBytes, 0, 4, 2, 256, B, w.tpf2, w.tpf2, colsum + cs_head(head, l - 1, B), 1, 256, base + l, 0);
        // output layer: do^T a3 (A = the 16-column d(pre-tanh) image, M' = 128 with rows 3.. ignored)
        add(doimg, kImgBytes, 1, hact + (size_t)2 * head_tiles * kTileBytes, kTileBytes, 0, 4, 1, 256, B, w.tpf2, w.tpf2,
            colsum + cs_out(head, B), 1, 16, base + 3, 0);
    }
    const int n_head_gemms = ng;
    if (!frozen) {
        uint8_t* tact = ws + w.tact;
        uint8_t* tdz = ws + w.tdz;
        uint8_t* enc = ws + w.enc;
        for (int l = 4; l >= 1; --l)
            add(tdz + (size_t)l * w.trunk_tiles * kTileBytes, kTileBytes, 4, tact + (size_t)(l - 1) * w.trunk_tiles * kTileBytes,
                kTileBytes, 0, 4, 2, 256, w.trunk_feet, w.tpf2, w.tpf2, colsum + cs_trunk(l, B), 1, 256, P_T0 + l, 0);
        for (int kb = 0; kb < 3; ++kb)
            add(tdz, kTileBytes, 4, enc, kEncTileBytes, (uint32_t)kb * 4 * kImgBytes, kb < 2 ? 4 : 1, 2, kb < 2 ? 256 : 64,
                w.trunk_feet, w.tpf2, w.tpf2, kb == 0 ? colsum + cs_trunk(0, B) : nullptr, kb == 0 ? 1 : 0, 256, P_T0, kb);
    }
    D.n_gemms = ng;
    D.dbg = g_chain_dbg;
    // Work split.  Every GEMM streams its tiles through a 3-stage ring; under load a stage costs a fixed part (ring round trip)
    // plus its bytes, and bytes that were written by the kernels just before this one (the trunk's activations / dZ, 5 MB each)
    // or that every foot re-reads (the shared trunk output h5) come from the 126 MB L2 (measured per stage on cfg2, all CTAs
    // streaming: 1.25 us for 64 KB from HBM, 0.97 us for 40 KB, 1.14 us for 32 KB HBM + 32 KB L2, 0.75 us for 40 KB from L2:
    // profiles/r2_dw_cta_lifetimes.txt).  The cost-weighted unit lists of all GEMMs are laid end to end and cut into one equal
    // piece per CTA; a piece that crosses a GEMM border becomes two segments, each with its own partial (deterministic: the
    // cut depends only on the shapes).
    static float c_fix = -1.f, c_hbm = 0.f, c_l2 = 0.f, c_seg = 0.f;
    if (c_fix < 0.f) {
        const char* e = getenv("FINDB200_DW_COST");        // "fixed_us,us_per_64KB_hbm,us_per_64KB_l2,us_per_segment" (tuning)
        float a = 0.70f, b = 0.55f, c = 0.45f, d = 6.0f;
        if (e) sscanf(e, "%f,%f,%f,%f", &a, &b, &c, &d);
        c_fix = a; c_hbm = b; c_l2 = c; c_seg = d;
    }
    // every segment ends with the drain of its MMAs and a 256 KB partial written from TMEM: a CTA whose piece crosses a GEMM
    // border pays that twice (measured: such CTAs ended 6..8 us after their neighbours, profiles/r2_dw_split.md)
    const long long seg_cost = (long long)(c_seg * 1000.0f);
    long long unit_cost[kMaxDwGemms], total_cost = 0;      // in ns per unit (one tile = two stages)
    for (int i = 0; i < ng; ++i) {
        const DwGemm& g = D.g[i];
        const bool trunk = g.feet == w.trunk_feet && g.A >= ws + w.tdz && g.A < ws + w.tdz + (size_t)5 * w.trunk_tiles * kTileBytes;
        const bool b_l2 = trunk || (g.b_foot_stride == 0 && g.feet > 1);
        const double a_kb = g.a_imgs * 8.0, b_kb = g.b_imgs * 8.0;                      // per stage
        const double hbm = (trunk ? 0.0 : a_kb) + (b_l2 ? 0.0 : b_kb), l2 = (trunk ? a_kb : 0.0) + (b_l2 ? b_kb : 0.0);
        double us = c_fix + hbm / 64.0 * c_hbm + l2 / 64.0 * c_l2;
        const double mma_us = g.mhalves * 4 * 0.09;                                     // 172 cycles per K = 16 MMA
        if (us < mma_us) us = mma_us;
        unit_cost[i] = (long long)(2.0 * us * 1000.0);
        if (unit_cost[i] < 1) unit_cost[i] = 1;
        total_cost += unit_cost[i] * g.feet * g.tpf;
    }
    int n_ctas = sms < kMaxDwCtas ? sms : kMaxDwCtas;
    {
        long long total_units = 0;
        for (int i = 0; i < ng; ++i) total_units += (long long)D.g[i].feet * D.g[i].tpf;
        if (total_units < n_ctas) n_ctas = (int)total_units;
        if (n_ctas < 1) n_ctas = 1;
    }
    total_cost += (long long)(n_ctas + ng - 1) * seg_cost;
    FB_REQUIRE(n_ctas + ng <= kMaxDwSegs, FINDB200_ERR_UNSUPPORTED, "field_backward: too many weight-gradient GEMMs");
    // Early / late launch.  The trunk's backward chain is a latency chain on trunk_tiles CTAs that moves no HBM bytes, and this
    // kernel is bound by HBM: the CTAs that do not fit beside the trunk chain's would otherwise idle, and so would the HBM.  So
    // the first n_early CTAs of the plan are launched on a side stream as soon as the heads' backward has finished (they run
    // beside sum_feet and the trunk chain, as 2-CTA clusters so that they fill whole TPCs and leave whole TPCs to the trunk's
    // CTA pairs), the others after the trunk chain; an early CTA runs longer and with fewer competitors, so its share of the
    // timeline is early_weight times a late CTA's.  The head GEMMs come first in the timeline; the plan is only split when
    // the early CTAs' share ends inside them (the trunk GEMMs need the trunk chain's output).
    if (g_dw_early_weight < 0.f) {
        const char* e = getenv("FINDB200_DW_SPLIT");        // early CTA's share relative to a late CTA's; 0 = one launch
        g_dw_early_weight = e ? (float)atof(e) : 1.65f;
        if (g_dw_early_weight < 0.f) g_dw_early_weight = 0.f;
    }
    const float early_weight = g_dw_early_weight;
    int n_early = 0;
    if (!frozen && early_weight > 0.f && chain_pairs_enabled() && n_ctas == sms && w.trunk_tiles <= 2 * sms) {
        n_early = (sms - w.trunk_tiles) & ~1;
        if (n_early < 16 || n_early > n_ctas - 16) n_early = 0;
    }
    int n_seg = 0;
  plan_again:
    n_seg = 0;
    {
        // walk the timeline: CTA b ends at cost tick total * (b + 1) / n_ctas, rounded to a unit boundary of the GEMM it falls in
        int gi = 0, u = 0;                      // current position: GEMM gi, unit u
        long long pos = 0;                      // cost tick of the current position
        int slot[kMaxDwGemms];
        for (int i = 0; i < ng; ++i) { slot[i] = 0; D.g[i].ncta = 0; }
        for (int b = 0; b < n_ctas; ++b) {
            D.cta_seg[b] = (unsigned short)n_seg;
            long long target = (b == n_ctas - 1) ? total_cost : total_cost * (b + 1) / n_ctas;
            if (n_early > 0 && b < n_ctas - 1) {
                const double sumw = (double)n_early * early_weight + (n_ctas - n_early);
                const double cum = b + 1 <= n_early ? (double)(b + 1) * early_weight : (double)n_early * early_weight + (b + 1 - n_early);
                target = (long long)((double)total_cost * cum / sumw);
            }
            const bool last = b == n_ctas - 1;
            while (gi < ng) {
                const int units = D.g[gi].feet * D.g[gi].tpf;
                const bool first = n_seg == (int)D.cta_seg[b];                // this CTA has no segment yet
                const long long p0 = pos + seg_cost;                          // opening a segment costs its epilogue
                if (!first && !last && p0 + unit_cost[gi] / 2 >= target) break;     // no room for another segment
                const long long end_g = p0 + (long long)(units - u) * unit_cost[gi];
                int u_end;
                if (end_g <= target || last) u_end = units;
                else {
                    u_end = u + (int)((target - p0 + unit_cost[gi] / 2) / unit_cost[gi]);       // nearest unit boundary
                    if (u_end <= u) u_end = u + 1;
                    if (u_end > units) u_end = units;
                }
                if (u_end > u) {
                    DwSeg& sg = D.seg[n_seg++];
                    sg.gemm = (short)gi; sg.slot = (short)slot[gi]++; sg.u0 = u; sg.u1 = u_end;
                    ++D.g[gi].ncta;
                    pos = p0 + (long long)(u_end - u) * unit_cost[gi];
                    u = u_end;
                }
                if (u < units) break;                   // this CTA's share ends inside GEMM gi
                ++gi; u = 0;
                if (pos >= target && b < n_ctas - 1) break;
            }
        }
        D.cta_seg[n_ctas] = (unsigned short)n_seg;
        FB_REQUIRE(gi == ng, FINDB200_ERR_UNSUPPORTED, "field_backward: internal: dW work split did not cover every GEMM");
        if (n_early > 0) {
            bool ok = true;
            for (int sgi = 0; sgi < D.cta_seg[n_early]; ++sgi) ok = ok && D.seg[sgi].gemm < n_head_gemms;
            if (!ok) { n_early = 0; goto plan_again; }
        }
    }
    int c0 = 0;
    for (int i = 0; i < ng; ++i) {
        D.g[i].cta0 = c0;
        D.g[i].partial = partial + (size_t)c0 * kW * kW;
        c0 += D.g[i].ncta;
    }
    FB_REQUIRE(c0 <= w.n_partials, FINDB200_ERR_UNSUPPORTED, "field_backward: internal: more dW partials than workspace slots");
    if (g_chain_dbg && (g_chain_dbg_flags & 512))             // diagnostics (tools/dw_trace.py): the static CTA allocation
        for (int i = 0; i < ng; ++i)
            fprintf(stderr, "dw gemm %d: param %d kblock %d feet %d a_imgs %d b_imgs %d N %d cta0 %d ncta %d\n", i, gemm_param[i], gemm_kblock[i],
                    D.g[i].feet, D.g[i].a_imgs, D.g[i].b_imgs, D.g[i].N, D.g[i].cta0, D.g[i].ncta);
    if (g_chain_dbg && (g_chain_dbg_flags & 512))
        for (int b = 0; b < n_ctas; ++b)
            for (int sgi = D.cta_seg[b]; sgi < D.cta_seg[b + 1]; ++sgi)
                fprintf(stderr, "dw cta %d: gemm %d units %d %d cost %lld early %d\n", b, D.seg[sgi].gemm, D.seg[sgi].u0, D.seg[sgi].u1,
                        unit_cost[D.seg[sgi].gemm], b < n_early ? 1 : 0);
    cudaStream_t side = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
    if (n_early > 0) {
        rc = fb::side_stream(1, &side, &ev_fork, &ev_join);
        if (rc) return rc;
        FB_CUDA(cudaEventRecord(ev_fork, st));
        FB_CUDA(cudaStreamWaitEvent(side, ev_fork, 0));
        D.cta_base = 0;
        rc = launch_dw(D, n_early, side, true);
        if (rc) return rc;
        FB_MARK_END(side);
        FB_CUDA(cudaEventRecord(ev_join, side));
    }
    if (!frozen) {
        rc = 0;
        if (pos_shared && B > 1) {
            FB_MARK("sum_feet", st);
            FB_CUDA(fb::launch_pdl(sum_feet_kernel, dim3(w.tpf2 * (kTileBytes / 16) / 256, n_heads), 256, 0, st, ws + w.hdz, ws + w.hsum, B, w.tpf2,
                                                                                             heads_mask));
            FB_CUDA(cudaGetLastError());
            for (int hd = 0; hd < 2; ++hd) C.tsrc[hd] = ws + w.hsum + (size_t)hd * w.trunk_tiles * kTileBytes;
        } else {
            for (int hd = 0; hd < 2; ++hd) C.tsrc[hd] = ws + w.hdz + (size_t)hd * 3 * head_tiles * kTileBytes;
        }
        C.n_items = w.trunk_tiles / ((w.trunk_tiles <= 2 * sms) ? 1 : 2);
        rc = chain_pairs_enabled() ? launch_chain2(3, C, sms, st) : launch_chain(3, (w.trunk_tiles <= 2 * sms) ? 1 : 2, C, sms, st);
        if (rc) return rc;
    }
    D.cta_base = n_early;
    rc = launch_dw(D, n_ctas - n_early, st);
    if (rc) return rc;
    if (n_early > 0) FB_CUDA(cudaStreamWaitEvent(st, ev_join, 0));

    ReduceParams R;
    memset(&R, 0, sizeof(R));
    if (!frozen) {
        R.maxabs = (const uint32_t*)(ws + w.scal);
        const int in_dims[P_COUNT] = {kEnc, kW, kW, kW, kW, kW + p->shapevec_size + p->posevec_size, kW, kW, kW,
                                      kW + p->texvec_size, kW, kW, kW};
        int nj = 0;
        for (int i = 0; i < ng; ++i) {
            const int prm = gemm_param[i];
            if (prm < 0) continue;
            ReduceJob& J = R.j[nj++];
            J.dst = gw->weight[prm]; J.partial = D.g[i].partial; J.dst_ld = in_dims[prm]; J.ncta = D.g[i].ncta;
            J.rows = (prm == P_D4 || prm == P_C4) ? 3 : kW;
            J.cols = kW; J.dst_col0 = 0; J.perm_kbase = -1; J.src_col0 = 0;
            if (prm == P_T0) { J.perm_kbase = gemm_kblock[i] * 256; J.cols = gemm_kblock[i] < 2 ? 256 : 3; }
        }
        // parameters of a head that received no gradient: write zeros (ncta = 0 sums nothing)
        for (int head = 0; head < 2; ++head) {
            if (heads_mask & (1 << head)) continue;
            const int base = head ? P_C1 : P_D1;
            for (int l = 0; l < 4; ++l) {
                ReduceJob& J = R.j[nj++];
                J.dst = gw->weight[base + l]; J.partial = partial; J.dst_ld = in_dims[base + l]; J.ncta = 0;
                J.rows = l == 3 ? 3 : kW; J.cols = kW; J.dst_col0 = 0; J.perm_kbase = -1; J.src_col0 = 0;
            }
        }
        R.n = nj;
    }

    FinishParams F;
    memset(&F, 0, sizeof(F));
    F.maxabs = (const uint32_t*)(ws + w.scal);
    F.colsum = colsum; F.regsums = (const long long*)(ws + w.regsums);
    F.Wd = p->weight[P_D1]; F.Wc = p->weight[P_C1];
    F.z_shape = z_shape; F.z_pose = z_pose; F.z_tex = z_tex; F.reg = reg;
    F.B = B; F.S = p->shapevec_size; F.Pz = p->posevec_size; F.T = p->texvec_size;
    F.heads_mask = heads_mask; F.frozen = frozen; F.have_reg = (reg && (grad_verts || g_grad_verts_fx)) ? 1 : 0;
    if (gw) for (int i = 0; i < P_COUNT; ++i) { F.gW[i] = gw->weight[i]; F.gb[i] = gw->bias[i]; }
    F.gz_shape = gz_shape; F.gz_pose = gz_pose; F.gz_tex = gz_tex; F.greg = greg;
    int maxid = P_COUNT * kW;
    const int Lmax = (F.S + F.Pz > F.T ? F.S + F.Pz : F.T);
    if (kW * Lmax > maxid) maxid = kW * Lmax;
    if (B * (F.S + F.Pz + F.T) * 8 > maxid) maxid = B * (F.S + F.Pz + F.T) * 8;     // section 3: 32 (b, j) pairs per block
    if (B > maxid) maxid = B;
    FB_MARK("reduce_finish", st);
    {
        int gx = cdiv(maxid, 256);
        if (R.n > 0 && gx < 64) gx = 64;
        FB_CUDA(fb::launch_pdl(reduce_finish_kernel, dim3(gx, R.n + 5), 256, 0, st, R, F));
    }
    FB_CUDA(cudaGetLastError());
    FB_MARK_END(st);
    return 0;
}

int findb200_set_step_overlap(int presort, float dw_early_weight) {
    fb::chamfer::g_presort = presort < 0 ? -1 : (presort ? 1 : 0);
    fb::field::g_dw_early_weight = dw_early_weight < 0.f ? -1.f : dw_early_weight;
    return 0;
}

}  // extern "C"
