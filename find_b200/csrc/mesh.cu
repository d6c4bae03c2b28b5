// find_b200/csrc/mesh.cu — the mesh-side pieces of FIND's 3-D training step, sm_100a.
//
//   (1) surface sampling: pytorch3d.ops.sample_points_from_meshes as called at src/model/losses.py:39,63,67 and
//       src/eval/eval_3d.py:149-150 — area-weighted face choice with replacement + uniform barycentric coordinates
//       (w0 = 1 - sqrt(u), w1 = sqrt(u)(1 - v), w2 = sqrt(u) v), differentiable w.r.t. the vertices.
//   (2) MeshSmoothnessLoss (src/model/losses.py:93-99): mesh_edge_loss + mesh_laplacian_smoothing(method="cot"),
//       forward and backward in one enqueue over the static template topology.
//
// These are HBM/latency-bound gather kernels over a few 10^4 faces per foot; nothing here is GEMM-shaped.
//
// Sampling is restated, not translated: torch's multinomial / rand streams cannot be reproduced, so the draw is defined
// here in a way a CPU oracle can restate bit for bit (oracle/mesh_oracle.py):
//   area_f  = 0.5 * sqrt(|e1 x e2|^2) with every fp32 operation individually rounded (no FMA contraction);
//   q_f     = floor(double(area_f) / double(max_f area_f) * 2^32)  as uint64  -> integer CDF, order independent;
//   r0..r3  = Philox4x32-10(counter = (sample, mesh, draw_lo, draw_hi), key = seed), draw = *draw_counter (0 if NULL):
//             a device counter so that a captured CUDA graph draws fresh samples on every replay;  target = mulhi64(r0:r1, total);
//   face    = first f with cdf_incl[f] > target;   u = (r2 >> 8) 2^-24,  v = (r3 >> 8) 2^-24.
// Callers that need PyTorch3D's exact draw pass face indices and (u, v) in; the interpolation then follows the same
// operation order as PyTorch3D's `w0*a + w1*b + w2*c` (each product and sum rounded once).
#include "common.cuh"
#include "../../include/find_b200.h"
#include <math_constants.h>

namespace fb {
namespace mesh {

// ---- Philox4x32-10 (Salmon et al. 2011; the constants are the published ones) --------------------------------------
FB_DEVINL void philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1, uint32_t out[4]) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
        const uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

FB_DEVINL float face_area_exact(const float* a, const float* b, const float* c) {
    const float e1x = __fsub_rn(b[0], a[0]), e1y = __fsub_rn(b[1], a[1]), e1z = __fsub_rn(b[2], a[2]);
    const float e2x = __fsub_rn(c[0], a[0]), e2y = __fsub_rn(c[1], a[1]), e2z = __fsub_rn(c[2], a[2]);
    const float cx = __fsub_rn(__fmul_rn(e1y, e2z), __fmul_rn(e1z, e2y));
    const float cy = __fsub_rn(__fmul_rn(e1z, e2x), __fmul_rn(e1x, e2z));
    const float cz = __fsub_rn(__fmul_rn(e1x, e2y), __fmul_rn(e1y, e2x));
    const float n2 = __fadd_rn(__fadd_rn(__fmul_rn(cx, cx), __fmul_rn(cy, cy)), __fmul_rn(cz, cz));
    return __fmul_rn(0.5f, __fsqrt_rn(n2));
}

struct SampleGeom {
    const float* verts;        // [N][V][3] (vert_stride = 0: one shared vertex array)
    const int32_t* faces;      // [F][3] shared, or [N][F][3]
    const int64_t* num_faces;  // [N] valid faces per mesh, nullable (= F)
    long long vert_stride, face_stride;   // elements between meshes (0 = shared)
    int N, V, F;
};

// areas [N][F] fp32 (0 for padding / out-of-range faces) and the per-mesh maximum (uint bits: areas are >= +0)
__global__ void face_area_kernel(const SampleGeom g, float* __restrict__ area, unsigned* __restrict__ amax_bits) {
    const int n = blockIdx.y;
    const int f = blockIdx.x * blockDim.x + threadIdx.x;
    const int nf = g.num_faces ? (int)min((long long)g.num_faces[n], (long long)g.F) : g.F;
    float a = 0.f;
    if (f < nf) {
        const int32_t* fc = g.faces + (size_t)n * g.face_stride + (size_t)f * 3;
        const int i0 = fc[0], i1 = fc[1], i2 = fc[2];
        if ((unsigned)i0 < (unsigned)g.V && (unsigned)i1 < (unsigned)g.V && (unsigned)i2 < (unsigned)g.V) {
            const float* vb = g.verts + (size_t)n * g.vert_stride;
            a = face_area_exact(vb + (size_t)i0 * 3, vb + (size_t)i1 * 3, vb + (size_t)i2 * 3);
            if (!(a >= 0.f) || a == CUDART_INF_F) a = 0.f;       // NaN / inf geometry is never drawn
        }
    }
    if (f < g.F) area[(size_t)n * g.F + f] = a;
    unsigned m = __float_as_uint(a);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(amax_bits + n, m);
}

// integer CDF: one CTA per mesh walks its faces 1024 at a time (block scan + running carry)
__global__ void __launch_bounds__(1024) face_cdf_kernel(const float* __restrict__ area, const unsigned* __restrict__ amax_bits,
                                                        int F, unsigned long long* __restrict__ cdf) {
    __shared__ unsigned long long s_warp[32];
    __shared__ unsigned long long s_carry;
    const int n = blockIdx.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const double amax = (double)__uint_as_float(amax_bits[n]);
    if (threadIdx.x == 0) s_carry = 0ull;
    __syncthreads();
    for (int f0 = 0; f0 < F; f0 += 1024) {
        const int f = f0 + threadIdx.x;
        unsigned long long q = 0ull;
        if (f < F && amax > 0.0) q = (unsigned long long)(__ddiv_rn((double)area[(size_t)n * F + f], amax) * 4294967296.0);
        unsigned long long x = q;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_warp[wid] = x;
        __syncthreads();
        if (wid == 0) {
            unsigned long long w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const unsigned long long y = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += y;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const unsigned long long incl = s_carry + (wid ? s_warp[wid - 1] : 0ull) + x;
        if (f < F) cdf[(size_t)n * F + f] = incl;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = incl;
        __syncthreads();
    }
}

struct SampleOut {
    float* samples;            // [N][S][3]
    int64_t* face_idx;         // [N][S]   (mesh-local face index), nullable
    float* bary;               // [N][S][3], nullable
    const float* attr;         // [N][V][C] per-vertex attribute (TexturesVertex colours), nullable
    float* attr_out;           // [N][S][C]
    long long attr_stride;
    int C;
};

__global__ void sample_points_kernel(const SampleGeom g, const unsigned long long* __restrict__ cdf,
                                     const int64_t* __restrict__ face_in, const float* __restrict__ uv_in, int S,
                                     unsigned long long seed, const unsigned long long* __restrict__ draw_counter,
                                     const SampleOut o) {
    const int n = blockIdx.y;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    uint32_t r[4] = {0u, 0u, 0u, 0u};
    if (!face_in || !uv_in) {
        const unsigned long long draw = draw_counter ? *draw_counter : 0ull;
        philox4x32_10((uint32_t)s, (uint32_t)n, (uint32_t)draw, (uint32_t)(draw >> 32), (uint32_t)seed, (uint32_t)(seed >> 32), r);
    }
    int f;
    if (face_in) {
        const long long fi = face_in[(size_t)n * S + s];
        f = (int)min(max(fi, 0ll), (long long)g.F - 1);
    } else {
        const unsigned long long* c = cdf + (size_t)n * g.F;
        const unsigned long long total = c[g.F - 1];
        const unsigned long long target = __umul64hi(((unsigned long long)r[0] << 32) | r[1], total);
        int lo = 0, hi = g.F - 1;              // first f with c[f] > target (total = 0: f = 0)
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(c + mid) > target) hi = mid; else lo = mid + 1;
        }
        f = total ? lo : 0;
    }
    float u, v;
    if (uv_in) {
        u = uv_in[(size_t)n * S + s];
        v = uv_in[(size_t)g.N * S + (size_t)n * S + s];
    } else {
        u = (float)(r[2] >> 8) * 5.9604644775390625e-8f;      // 2^-24: [0, 1)
        v = (float)(r[3] >> 8) * 5.9604644775390625e-8f;
    }
    const float us = __fsqrt_rn(u);
    const float w0 = __fsub_rn(1.0f, us), w1 = __fmul_rn(us, __fsub_rn(1.0f, v)), w2 = __fmul_rn(us, v);
    const int32_t* fc = g.faces + (size_t)n * g.face_stride + (size_t)f * 3;
    const int i0 = fc[0], i1 = fc[1], i2 = fc[2];
    const float* vb = g.verts + (size_t)n * g.vert_stride;
    const size_t at = (size_t)n * S + s;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float a = vb[(size_t)i0 * 3 + k], b = vb[(size_t)i1 * 3 + k], c = vb[(size_t)i2 * 3 + k];
        o.samples[at * 3 + k] = __fadd_rn(__fadd_rn(__fmul_rn(w0, a), __fmul_rn(w1, b)), __fmul_rn(w2, c));
    }
    if (o.face_idx) o.face_idx[at] = f;
    if (o.bary) { o.bary[at * 3] = w0; o.bary[at * 3 + 1] = w1; o.bary[at * 3 + 2] = w2; }
    if (o.attr && o.attr_out) {
        const float* ab = o.attr + (size_t)n * o.attr_stride;
        for (int k = 0; k < o.C; ++k) {
            const float a = ab[(size_t)i0 * o.C + k], b = ab[(size_t)i1 * o.C + k], c = ab[(size_t)i2 * o.C + k];
            o.attr_out[at * o.C + k] = __fadd_rn(__fadd_rn(__fmul_rn(w0, a), __fmul_rn(w1, b)), __fmul_rn(w2, c));
        }
    }
}

// d samples -> d verts: grad_verts[n, face[k]] += bary[k] * grad_sample (index_put(accumulate) in PyTorch3D's autograd graph)
__global__ void sample_points_bwd_kernel(const float* __restrict__ grad_samples, const int32_t* __restrict__ faces,
                                         long long face_stride, const int64_t* __restrict__ face_idx,
                                         const float* __restrict__ bary, int S, int V, int F, float scale,
                                         float* __restrict__ grad_verts) {
    const int n = blockIdx.y;
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= S) return;
    const size_t at = (size_t)n * S + s;
    const long long f = face_idx[at];
    if (f < 0 || f >= F) return;
    const int32_t* fc = faces + (size_t)n * face_stride + (size_t)f * 3;
    const float gx = grad_samples[at * 3] * scale, gy = grad_samples[at * 3 + 1] * scale, gz = grad_samples[at * 3 + 2] * scale;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float w = bary[at * 3 + k];
        float* dst = grad_verts + ((size_t)n * V + fc[k]) * 3;
        atomicAdd(dst, w * gx); atomicAdd(dst + 1, w * gy); atomicAdd(dst + 2, w * gz);
    }
}

}  // namespace mesh
}  // namespace fb

// =====================================================================================================================
// MeshSmoothnessLoss  (src/model/losses.py:93-99):  0.1 * mesh_laplacian_smoothing(meshes, "cot") + 10 * mesh_edge_loss(meshes)
// Static topology shared by all N meshes of the batch (the template): unique edges [E][2], faces [F][3] and a
// vertex -> (face, corner) incidence list in CSR form (built once per template by the host side, find_b200/mesh.py).
//
// PyTorch3D's definitions (restated; the package is not on this box):
//   edge:      mean over edges of (|v0 - v1| - target)^2 per mesh, then mean over meshes.
//   laplacian: L = cot-weight matrix built under no_grad (L_ij = sum over faces containing edge ij of cot(angle opposite
//              ij) / 4, area by Heron's formula clamped at 1e-12), norm_w_i = 1 / rowsum_i where rowsum_i > 0 (else the
//              rowsum itself), loss_i = |(L v)_i * norm_w_i - v_i|, mean over vertices, mean over meshes.
//              Gradient flows only through the explicit `v`:  d/dv_j = sum_i L_ij (norm_w_i g_i) - g_j,
//              g_i = (L v norm_w - v)_i / loss_i / (V N)   (0 where loss_i = 0, as torch's norm backward).
namespace fb {
namespace mesh {

struct Topo {
    const int32_t* faces;      // [F][3]
    const int32_t* edges;      // [E][2]
    const int32_t* vf_ptr;     // [V+1]
    const int32_t* vf_face;    // [nnz]
    const int32_t* vf_corner;  // [nnz]  which corner of vf_face the vertex is
    int V, F, E, nnz;
};

FB_DEVINL float block_sum_to(float v, double* dst) {
    __shared__ float s_red[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) s_red[wid] = v;
    __syncthreads();
    if (wid == 0) {
        v = lane < (int)(blockDim.x >> 5) ? s_red[lane] : 0.f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v != 0.f) atomicAdd(dst, (double)v);
    }
    return v;
}

// edge term: loss and gradient (atomics: an edge's two endpoints are shared with other edges)
__global__ void edge_loss_kernel(const float* __restrict__ verts, const Topo t, float target, float gscale,
                                 double* __restrict__ sum, float* __restrict__ grad) {
    const int n = blockIdx.y;
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    float l = 0.f;
    if (e < t.E) {
        const int i0 = t.edges[2 * e], i1 = t.edges[2 * e + 1];
        const float* a = verts + ((size_t)n * t.V + i0) * 3;
        const float* b = verts + ((size_t)n * t.V + i1) * 3;
        const float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
        const float len = sqrtf(dx * dx + dy * dy + dz * dz);
        const float r = len - target;
        l = r * r;
        if (grad && len > 0.f) {
            const float c = gscale * 2.f * r / len;
            float* ga = grad + ((size_t)n * t.V + i0) * 3;
            float* gb = grad + ((size_t)n * t.V + i1) * 3;
            atomicAdd(ga, c * dx); atomicAdd(ga + 1, c * dy); atomicAdd(ga + 2, c * dz);
            atomicAdd(gb, -c * dx); atomicAdd(gb + 1, -c * dy); atomicAdd(gb + 2, -c * dz);
        }
    }
    block_sum_to(l, sum);
}

// cot weights of face (p0,p1,p2): cot[k] = cot(angle at corner k) / 4
FB_DEVINL void face_cots(const float* p0, const float* p1, const float* p2, float cot[3]) {
    auto dist = [](const float* a, const float* b) {
        const float x = a[0] - b[0], y = a[1] - b[1], z = a[2] - b[2];
        return sqrtf(x * x + y * y + z * z);
    };
    const float A = dist(p1, p2), B = dist(p0, p2), C = dist(p0, p1);
    const float s = 0.5f * (A + B + C);
    const float area = sqrtf(fmaxf(s * (s - A) * (s - B) * (s - C), 1e-12f));
    const float A2 = A * A, B2 = B * B, C2 = C * C;
    cot[0] = (B2 + C2 - A2) / area * 0.25f;
    cot[1] = (A2 + C2 - B2) / area * 0.25f;
    cot[2] = (A2 + B2 - C2) / area * 0.25f;
}

// pass 1, one thread per (mesh, vertex): gather L v over the incident faces, loss_i, and h_i = norm_w_i g_i, g_i
__global__ void lap_gather_kernel(const float* __restrict__ verts, const Topo t, float gscale, double* __restrict__ sum,
                                  float2* __restrict__ wts, float* __restrict__ h, float* __restrict__ g) {
    const int n = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float li = 0.f;
    if (i < t.V) {
        const float* vb = verts + (size_t)n * t.V * 3;
        float lx = 0.f, ly = 0.f, lz = 0.f, rs = 0.f;
        for (int k = t.vf_ptr[i]; k < t.vf_ptr[i + 1]; ++k) {
            const int f = t.vf_face[k], c = t.vf_corner[k];
            const int32_t* fc = t.faces + (size_t)f * 3;
            float cot[3];
            face_cots(vb + (size_t)fc[0] * 3, vb + (size_t)fc[1] * 3, vb + (size_t)fc[2] * 3, cot);
            const int c1 = c == 2 ? 0 : c + 1, c2 = c == 0 ? 2 : c - 1;
            const float w1 = cot[c2], w2 = cot[c1];          // edge (i, corner c1) is opposite corner c2, and vice versa
            const float* q1 = vb + (size_t)fc[c1] * 3;
            const float* q2 = vb + (size_t)fc[c2] * 3;
            lx += w1 * q1[0] + w2 * q2[0]; ly += w1 * q1[1] + w2 * q2[1]; lz += w1 * q1[2] + w2 * q2[2];
            rs += w1 + w2;
            if (wts) wts[(size_t)n * t.nnz + k] = make_float2(w1, w2);
        }
        const float nw = rs > 0.f ? 1.f / rs : rs;
        const float dx = lx * nw - vb[(size_t)i * 3], dy = ly * nw - vb[(size_t)i * 3 + 1], dz = lz * nw - vb[(size_t)i * 3 + 2];
        li = sqrtf(dx * dx + dy * dy + dz * dz);
        if (g) {
            const float inv = li > 0.f ? gscale / li : 0.f;
            const size_t at = ((size_t)n * t.V + i) * 3;
            g[at] = dx * inv; g[at + 1] = dy * inv; g[at + 2] = dz * inv;
            h[at] = dx * inv * nw; h[at + 1] = dy * inv * nw; h[at + 2] = dz * inv * nw;
        }
    }
    block_sum_to(li, sum);
}

// pass 2: grad_j += sum_i L_ji h_i - g_j   (L is symmetric, so the same incidence gather serves the transpose)
__global__ void lap_bwd_kernel(const Topo t, const float2* __restrict__ wts, const float* __restrict__ h,
                               const float* __restrict__ g, float* __restrict__ grad) {
    const int n = blockIdx.y;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= t.V) return;
    const float* hb = h + (size_t)n * t.V * 3;
    float ax = 0.f, ay = 0.f, az = 0.f;
    for (int k = t.vf_ptr[i]; k < t.vf_ptr[i + 1]; ++k) {
        const int f = t.vf_face[k], c = t.vf_corner[k];
        const int32_t* fc = t.faces + (size_t)f * 3;
        const int c1 = c == 2 ? 0 : c + 1, c2 = c == 0 ? 2 : c - 1;
        const float2 w = wts[(size_t)n * t.nnz + k];
        const float* q1 = hb + (size_t)fc[c1] * 3;
        const float* q2 = hb + (size_t)fc[c2] * 3;
        ax += w.x * q1[0] + w.y * q2[0]; ay += w.x * q1[1] + w.y * q2[1]; az += w.x * q1[2] + w.y * q2[2];
    }
    const size_t at = ((size_t)n * t.V + i) * 3;
    grad[at] += ax - g[at]; grad[at + 1] += ay - g[at + 1]; grad[at + 2] += az - g[at + 2];
}

__global__ void smooth_finish_kernel(const double* __restrict__ sums, int N, int V, int E, float* __restrict__ loss2) {
    loss2[0] = (float)(sums[0] / ((double)(E < 1 ? 1 : E) * (double)(N < 1 ? 1 : N)));
    loss2[1] = (float)(sums[1] / ((double)(V < 1 ? 1 : V) * (double)(N < 1 ? 1 : N)));
}

}  // namespace mesh
}  // namespace fb

using namespace fb;
using namespace fb::mesh;

static inline size_t al256(size_t x) { return (x + 255) & ~(size_t)255; }

extern "C" {

size_t findb200_sample_workspace_bytes(int N, int F) {
    if (N < 1 || F < 1) return 0;
    return al256((size_t)N * 4) + al256((size_t)N * F * 4) + al256((size_t)N * F * 8);
}

int findb200_sample_points_forward(const float* verts, int verts_shared, const int32_t* faces, int faces_shared,
                                   const int64_t* num_faces, int N, int V, int F, int S, unsigned long long seed,
                                   const unsigned long long* draw_counter,
                                   const int64_t* face_idx_in, const float* uv_in, float* samples, int64_t* face_idx,
                                   float* bary, const float* vert_attr, int attr_shared, int C, float* attr_out,
                                   void* workspace, size_t workspace_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(verts && faces && samples, FINDB200_ERR_ARG, "sample_points: verts, faces and samples are required");
    FB_REQUIRE(N >= 1 && V >= 1 && F >= 1 && S >= 0 && C >= 0, FINDB200_ERR_SHAPE, "sample_points: bad sizes N=%d V=%d F=%d S=%d", N, V, F, S);
    FB_REQUIRE((vert_attr == nullptr) == (attr_out == nullptr) || C == 0, FINDB200_ERR_ARG, "sample_points: vert_attr and attr_out go together");
    if (S == 0) return 0;
    SampleGeom g{verts, faces, num_faces, verts_shared ? 0ll : (long long)V * 3, faces_shared ? 0ll : (long long)F * 3, N, V, F};
    unsigned long long* cdf = nullptr;
    if (!face_idx_in) {
        const size_t need = findb200_sample_workspace_bytes(N, F);
        FB_REQUIRE(workspace && workspace_bytes >= need, FINDB200_ERR_WORKSPACE, "sample_points: workspace too small (%zu < %zu bytes)",
                   workspace_bytes, need);
        uint8_t* ws = (uint8_t*)workspace;
        unsigned* amax = (unsigned*)ws;
        const size_t o_area = al256((size_t)N * 4);
        float* area = (float*)(ws + o_area);
        cdf = (unsigned long long*)(ws + o_area + al256((size_t)N * F * 4));
        FB_MARK("sample_area", st);
        FB_CUDA(cudaMemsetAsync(amax, 0, o_area, st));
        face_area_kernel<<<dim3((F + 255) / 256, N), 256, 0, st>>>(g, area, amax);
        FB_CUDA(cudaGetLastError());
        FB_MARK("sample_cdf", st);
        face_cdf_kernel<<<N, 1024, 0, st>>>(area, amax, F, cdf);
        FB_CUDA(cudaGetLastError());
    }
    SampleOut o{samples, face_idx, bary, C > 0 ? vert_attr : nullptr, C > 0 ? attr_out : nullptr,
                attr_shared ? 0ll : (long long)V * C, C};
    FB_MARK("sample_points", st);
    sample_points_kernel<<<dim3((S + 127) / 128, N), 128, 0, st>>>(g, cdf, face_idx_in, uv_in, S, seed, draw_counter, o);
    FB_CUDA(cudaGetLastError());
    FB_MARK_END(st);
    return 0;
}

int findb200_sample_points_backward(const float* grad_samples, const int32_t* faces, int faces_shared,
                                    const int64_t* face_idx, const float* bary, int N, int V, int F, int S,
                                    float grad_scale, int accumulate, float* grad_verts, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(grad_samples && faces && face_idx && bary && grad_verts, FINDB200_ERR_ARG, "sample_points_backward: null argument");
    FB_REQUIRE(N >= 1 && V >= 1 && F >= 1 && S >= 0, FINDB200_ERR_SHAPE, "sample_points_backward: bad sizes");
    FB_MARK("sample_bwd", st);
    if (!accumulate) FB_CUDA(cudaMemsetAsync(grad_verts, 0, (size_t)N * V * 3 * sizeof(float), st));
    if (S > 0) {
        sample_points_bwd_kernel<<<dim3((S + 127) / 128, N), 128, 0, st>>>(grad_samples, faces, faces_shared ? 0ll : (long long)F * 3,
                                                                          face_idx, bary, S, V, F, grad_scale, grad_verts);
        FB_CUDA(cudaGetLastError());
    }
    FB_MARK_END(st);
    return 0;
}

size_t findb200_mesh_smooth_workspace_bytes(int N, int V, int nnz) {
    if (N < 1 || V < 1 || nnz < 0) return 0;
    return 256 + al256((size_t)N * nnz * 8) + 2 * al256((size_t)N * V * 3 * 4);
}

int findb200_mesh_smooth(const float* verts, int N, int V, const int32_t* faces, int F, const int32_t* edges, int E,
                         const int32_t* vf_ptr, const int32_t* vf_face, const int32_t* vf_corner, int nnz,
                         float target_length, float w_edge, float w_lap, const float* grad_loss, float* loss2,
                         float* grad_verts, int accumulate, void* workspace, size_t workspace_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(verts && faces && edges && vf_ptr && vf_face && vf_corner && loss2, FINDB200_ERR_ARG, "mesh_smooth: null argument");
    FB_REQUIRE(N >= 1 && V >= 1 && F >= 1 && E >= 1 && nnz >= 1, FINDB200_ERR_SHAPE, "mesh_smooth: bad sizes");
    FB_REQUIRE(grad_loss == nullptr, FINDB200_ERR_UNSUPPORTED, "mesh_smooth: pass the upstream gradient folded into w_edge / w_lap");
    const size_t need = findb200_mesh_smooth_workspace_bytes(N, V, nnz);
    FB_REQUIRE(workspace && workspace_bytes >= need, FINDB200_ERR_WORKSPACE, "mesh_smooth: workspace too small (%zu < %zu bytes)",
               workspace_bytes, need);
    uint8_t* ws = (uint8_t*)workspace;
    double* sums = (double*)ws;
    float2* wts = (float2*)(ws + 256);
    float* h = (float*)(ws + 256 + al256((size_t)N * nnz * 8));
    float* g = h + al256((size_t)N * V * 3 * 4) / 4;
    Topo t{faces, edges, vf_ptr, vf_face, vf_corner, V, F, E, nnz};
    FB_MARK("smooth", st);
    FB_CUDA(cudaMemsetAsync(sums, 0, 256, st));
    if (grad_verts && !accumulate) FB_CUDA(cudaMemsetAsync(grad_verts, 0, (size_t)N * V * 3 * sizeof(float), st));
    edge_loss_kernel<<<dim3((E + 255) / 256, N), 256, 0, st>>>(verts, t, target_length, w_edge / ((float)E * (float)N), sums,
                                                               grad_verts);
    FB_CUDA(cudaGetLastError());
    lap_gather_kernel<<<dim3((V + 127) / 128, N), 128, 0, st>>>(verts, t, w_lap / ((float)V * (float)N), sums + 1,
                                                                grad_verts ? wts : nullptr, grad_verts ? h : nullptr,
                                                                grad_verts ? g : nullptr);
    FB_CUDA(cudaGetLastError());
    if (grad_verts) {
        lap_bwd_kernel<<<dim3((V + 127) / 128, N), 128, 0, st>>>(t, wts, h, g, grad_verts);
        FB_CUDA(cudaGetLastError());
    }
    smooth_finish_kernel<<<1, 1, 0, st>>>(sums, N, V, E, loss2);
    FB_CUDA(cudaGetLastError());
    FB_MARK_END(st);
    return 0;
}

}  // extern "C"
