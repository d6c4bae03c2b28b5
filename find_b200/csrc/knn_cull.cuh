// find_b200/csrc/knn_cull.cuh — exact 1-NN with bounding-box culling (included by chamfer.cu).
//
// Same contract as the brute-force tile kernel (dist = fmaf(dz,dz,fmaf(dy,dy,dx*dx)), lowest ORIGINAL index on ties,
// bit-exact vs oracle/chamfer_oracle.c), but most (query group, target chunk) pairs are never evaluated:
//
//   1. cloud_{bbox,hist,scan,scatter,boxes}_kernel: a counting sort of every cloud into the cells of a cubic grid (G^3 cells
//      over the cloud's bounding box, Hilbert-curve cell order), written as SoA rows padded with +inf to a multiple of
//      32, plus the permutation back to original indices and the bounding box of every 32-point chunk.
//   2. knn_cull_kernel: persistent warps.  A warp owns 32*kQ consecutive sorted queries (spatially compact), kQ per
//      lane in registers.  It evaluates the chunk nearest to the group first, then walks the chunk boxes 32 at a time
//      (one box per lane): a chunk survives level 1 if LB(group box, chunk box) <= U, U = max over the group's queries
//      of their current best distance, and is evaluated if, at level 2, LB(query, chunk box) <= best[query] for some
//      query of the group.  Chunk evaluation is the packed-fp32x2 brute-force inner loop.
//
// Why the cull is exact: LB is computed with the SAME operation sequence as the distance, on the per-axis gaps
// g = max(0, tlo - qhi, qlo - thi).  Round-to-nearest subtraction, multiplication and fma are monotone, so for every
// query in the group box and every target in the chunk box the COMPUTED fp32 distance is >= LB.  A chunk is skipped
// only when LB > best[q] for all q, so it holds neither a closer point nor a tie.  Ties across chunks (equal
// distance, lower original index elsewhere) set a per-query flag in the hot loop and take a slow exact path.
#pragma once
#include "common.cuh"
#include <math_constants.h>

namespace fb {
namespace chamfer {

constexpr int kCullChunk = 32;

// One cloud set (x or y): [N] clouds of up to P points.
struct SortSet {
    const float* aos;       // [N][P][3]
    const int64_t* len;     // nullable
    float* soa;             // [N][3][Ppad]   sorted, padded with +inf
    int* perm;              // [N][Ppad]      sorted position -> original index
    float* box;             // [N][6][Cpad]   lo xyz, hi xyz per 32-point chunk
    float* sbox;            // [N][6][Spad]   lo xyz, hi xyz per 32-chunk super-chunk
    unsigned* bb_min;       // [N][3]  order-encoded floats, identity 0xFFFFFFFF
    unsigned* bb_max;       // [N][3]  order-encoded floats, identity 0
    int* hist;              // [N][ncell]     zeroed by the host
    int* cell;              // [N][P]         cell of point i
    int* rank;              // [N][P]         arrival rank of point i inside its cell
    int P, Ppad, Cpad, Spad, glog2;
    int N;                  // clouds of THIS set (the two sets of one launch may differ: kernels return for n >= N)
};
// zero[k]: an optional region the FIRST sort kernel clears (4-byte words) while it computes the bounding boxes: buffers that
// later kernels of a fused training step accumulate into, cleared here instead of by separate memset nodes
struct SortParams { SortSet set[2]; unsigned* zero_ptr[2]; unsigned zero_words[2]; };

FB_DEVINL unsigned enc_ord(float f) {       // monotone float -> uint
    const unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
FB_DEVINL float dec_ord(unsigned e) { return __uint_as_float((e & 0x80000000u) ? (e & 0x7fffffffu) : ~e); }

FB_DEVINL uint32_t spread3(uint32_t v) {   // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
// Hilbert index of an integer cell (bits per axis = b): consecutive indices are face-adjacent cells, so a run of the
// sorted order is a compact blob (no Morton jumps).  Skilling's axes-to-transpose, then bit interleave.
FB_DEVINL uint32_t hilbert3(uint32_t x, uint32_t y, uint32_t z, int b) {
    uint32_t X[3] = {x, y, z};
    const uint32_t M = 1u << (b - 1);
    for (uint32_t Q = M; Q > 1; Q >>= 1) {
        const uint32_t Pm = Q - 1;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            if (X[i] & Q) X[0] ^= Pm;
            else { const uint32_t t = (X[0] ^ X[i]) & Pm; X[0] ^= t; X[i] ^= t; }
        }
    }
    X[1] ^= X[0]; X[2] ^= X[1];
    uint32_t t = 0;
    for (uint32_t Q = M; Q > 1; Q >>= 1)
        if (X[2] & Q) t ^= Q - 1;
    X[0] ^= t; X[1] ^= t; X[2] ^= t;
    return (spread3(X[0]) << 2) | (spread3(X[1]) << 1) | spread3(X[2]);
}
FB_DEVINL int cell_coord(float p, float lo, float inv_h, int G) {
    const float f = (p - lo) * inv_h;
    int c = 0;
    if (f >= 0.f) c = f < (float)G ? (int)f : G - 1;   // NaN -> 0
    return c;
}
FB_DEVINL int set_len(const SortSet& S, int n) {
    int len = S.len ? (int)S.len[n] : S.P;
    return len < 0 ? 0 : (len > S.P ? S.P : len);
}

// 1. bounding box per cloud.  grid (ceil(maxP/1024), N, 2), block 256, 4 points per thread.
__global__ void __launch_bounds__(256) cloud_bbox_kernel(const SortParams prm) {
    fb::pdl_wait_trigger();
    {
        const unsigned nthreads = gridDim.x * gridDim.y * gridDim.z * 256u;
        const unsigned t = (blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) * 256u + threadIdx.x;
#pragma unroll
        for (int k = 0; k < 2; ++k)
            if (prm.zero_ptr[k])
                for (unsigned i = t; i < prm.zero_words[k]; i += nthreads) prm.zero_ptr[k][i] = 0u;
    }
    const SortSet& S = prm.set[blockIdx.z];
    const int n = blockIdx.y;
    if (n >= S.N) return;
    const int len = set_len(S, n);
    const int i0 = blockIdx.x * 1024;
    if (i0 >= len) return;
    const float* aos = S.aos + (size_t)n * S.P * 3;
    float lo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, hi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
#pragma unroll
    for (int r = 0; r < 4; ++r) {
        const int i = i0 + r * 256 + threadIdx.x;
        if (i < len) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { const float v = aos[(size_t)i * 3 + k]; lo[k] = fminf(lo[k], v); hi[k] = fmaxf(hi[k], v); }
        }
    }
    __shared__ float s_red[8][6];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { s_red[wid][k] = lo[k]; s_red[wid][3 + k] = hi[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        const int k = threadIdx.x;
        float v = s_red[0][k];
        for (int w = 1; w < 8; ++w) v = k < 3 ? fminf(v, s_red[w][k]) : fmaxf(v, s_red[w][k]);
        if (k < 3) { if (v < CUDART_INF_F) atomicMin(S.bb_min + n * 3 + k, enc_ord(v)); }
        else if (v > -CUDART_INF_F) atomicMax(S.bb_max + n * 3 + (k - 3), enc_ord(v));
    }
}

struct CloudGrid { float lo[3]; float inv_h; };
FB_DEVINL CloudGrid cloud_grid(const SortSet& S, int n) {
    CloudGrid g;
    float ext = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const unsigned a = S.bb_min[n * 3 + k], b = S.bb_max[n * 3 + k];
        const float lo = a == 0xFFFFFFFFu ? 0.f : dec_ord(a), hi = b == 0u ? 0.f : dec_ord(b);
        g.lo[k] = lo;
        ext = fmaxf(ext, hi - lo);
    }
    const int G = 1 << S.glog2;
    g.inv_h = (ext > 0.f && ext < CUDART_INF_F) ? (float)G / ext : 0.f;     // cubic cells; degenerate cloud -> one cell
    if (!(g.inv_h < CUDART_INF_F)) g.inv_h = 0.f;
    return g;
}

// 2. cell + arrival rank of every point.  grid (ceil(maxP/256), N, 2), block 256.
__global__ void __launch_bounds__(256) cloud_hist_kernel(const SortParams prm) {
    fb::pdl_wait_trigger();
    const SortSet& S = prm.set[blockIdx.z];
    const int n = blockIdx.y;
    if (n >= S.N) return;
    const int len = set_len(S, n);
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= len) return;
    const CloudGrid g = cloud_grid(S, n);
    const int G = 1 << S.glog2;
    const float* p = S.aos + ((size_t)n * S.P + i) * 3;
    const uint32_t cx = cell_coord(p[0], g.lo[0], g.inv_h, G), cy = cell_coord(p[1], g.lo[1], g.inv_h, G),
                   cz = cell_coord(p[2], g.lo[2], g.inv_h, G);
    const int cell = (int)hilbert3(cx, cy, cz, S.glog2);
    S.cell[(size_t)n * S.P + i] = cell;
    S.rank[(size_t)n * S.P + i] = atomicAdd(S.hist + ((size_t)n << (3 * S.glog2)) + cell, 1);
}

// 3. in-place exclusive scan of every cloud's histogram.  grid (N, 2), block 1024: warp w owns the contiguous segment
//    [w*seg, (w+1)*seg) and walks it in coalesced 32-entry rows.
__global__ void __launch_bounds__(1024) cloud_scan_kernel(const SortParams prm) {
    fb::pdl_wait_trigger();
    __shared__ int s_wsum[32];
    const SortSet& S = prm.set[blockIdx.y];
    if (S.P <= 0 || (int)blockIdx.x >= S.N) return;
    const int n = blockIdx.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int ncell = 1 << (3 * S.glog2), seg = ncell / 32;
    int* hist = S.hist + ((size_t)n << (3 * S.glog2)) + wid * seg;
    int tot = 0;
    for (int r = lane; r < seg; r += 32) tot += hist[r];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    if (lane == 0) s_wsum[wid] = tot;
    __syncthreads();
    if (wid == 0) {
        const int v = s_wsum[lane];
        int w = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, w, o);
            if (lane >= o) w += u;
        }
        s_wsum[lane] = w - v;   // exclusive over warps
    }
    __syncthreads();
    int carry = s_wsum[wid];
    for (int r0 = 0; r0 < seg; r0 += 32) {
        const int r = r0 + lane;
        const int v = r < seg ? hist[r] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (r < seg) hist[r] = carry + inc - v;
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
}

// 4. scatter into curve order: pos = start[cell] + rank.  grid (ceil(maxPpad/256), N, 2), block 256.
__global__ void __launch_bounds__(256) cloud_scatter_kernel(const SortParams prm) {
    fb::pdl_wait_trigger();
    const SortSet& S = prm.set[blockIdx.z];
    const int n = blockIdx.y;
    if (n >= S.N) return;
    const int len = set_len(S, n);
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= S.Ppad || S.P <= 0) return;
    float* soa = S.soa + (size_t)n * 3 * S.Ppad;
    int* perm = S.perm + (size_t)n * S.Ppad;
    if (i < len) {
        const float* p = S.aos + ((size_t)n * S.P + i) * 3;
        const int pos = S.hist[((size_t)n << (3 * S.glog2)) + S.cell[(size_t)n * S.P + i]] + S.rank[(size_t)n * S.P + i];
        soa[pos] = p[0]; soa[S.Ppad + pos] = p[1]; soa[2 * S.Ppad + pos] = p[2];
        perm[pos] = i;
    } else {
        soa[i] = CUDART_INF_F; soa[S.Ppad + i] = CUDART_INF_F; soa[2 * S.Ppad + i] = CUDART_INF_F;
        perm[i] = 0;
    }
}

// 5. bounding box of every 32-point chunk of the sorted cloud.  grid (ceil(maxC/8), N, 2), block 256 (warp per chunk).
__global__ void __launch_bounds__(256) cloud_boxes_kernel(const SortParams prm) {
    fb::pdl_wait_trigger();
    const SortSet& S = prm.set[blockIdx.z];
    const int n = blockIdx.y, lane = threadIdx.x & 31;
    if (n >= S.N) return;
    const int len = set_len(S, n);
    const int c = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (c * kCullChunk >= len) return;
    const float* soa = S.soa + (size_t)n * 3 * S.Ppad;
    float* box = S.box + (size_t)n * 6 * S.Cpad;
    const int i = c * kCullChunk + lane;
    float a[3], b[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float v = soa[k * S.Ppad + i];
        a[k] = i < len ? v : CUDART_INF_F;
        b[k] = i < len ? v : -CUDART_INF_F;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a[k] = fminf(a[k], __shfl_xor_sync(0xffffffffu, a[k], o));
            b[k] = fmaxf(b[k], __shfl_xor_sync(0xffffffffu, b[k], o));
        }
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { box[k * S.Cpad + c] = a[k]; box[(3 + k) * S.Cpad + c] = b[k]; }
    }
}

// 6. bounding box of every super-chunk (32 chunks).  grid (ceil(maxS/8), N, 2), block 256 (warp per super-chunk).
__global__ void __launch_bounds__(256) cloud_superboxes_kernel(const SortParams prm) {
    fb::pdl_wait_trigger();
    const SortSet& S = prm.set[blockIdx.z];
    const int n = blockIdx.y, lane = threadIdx.x & 31;
    if (n >= S.N) return;
    const int nchunk = (set_len(S, n) + kCullChunk - 1) / kCullChunk;
    const int sc = blockIdx.x * 8 + (threadIdx.x >> 5);
    if (sc * 32 >= nchunk) return;
    const float* box = S.box + (size_t)n * 6 * S.Cpad;
    float* sbox = S.sbox + (size_t)n * 6 * S.Spad;
    const int c = sc * 32 + lane;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float v = k < 3 ? CUDART_INF_F : -CUDART_INF_F;
        if (c < nchunk) v = box[k * S.Cpad + c];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float u = __shfl_xor_sync(0xffffffffu, v, o);
            v = k < 3 ? fminf(v, u) : fmaxf(v, u);
        }
        if (lane == 0) sbox[k * S.Spad + sc] = v;
    }
}

// ---- a cloud set laid out in a GIVEN order ----------------------------------------------------------------------------
// The deformed template of a training step moves a little per step and smoothly over the surface: the curve order of the
// TEMPLATE's vertices (one cloud, sorted once per step beside the forward pass) keeps every 32-point chunk of every deformed foot
// compact.  This kernel replaces steps 1-6 for such a set: points are gathered in the given order into the SoA rows (+inf
// padding), the permutation is the order itself, and chunk and super-chunk boxes are taken on the way.  Any permutation gives
// the same kNN result bit for bit (the boxes are true bounds of their members; ties resolve on ORIGINAL indices): a poor order
// only costs speed.  grid (ceil(Ppad / 1024), N), block 1024 = one super-chunk of 32 chunks; all clouds are full (no lengths).
struct OrderParams {
    const float* aos;       // [N][P][3]
    const int* order;       // [Ppad] sorted position -> original index, shared by all clouds (entries >= P unused)
    float* soa; int* perm; float* box; float* sbox;
    int P, Ppad, Cpad, Spad;
};
__global__ void __launch_bounds__(1024) cloud_apply_order_kernel(const OrderParams prm) {
    fb::pdl_wait_trigger();
    __shared__ float s_box[32][6];
    const int n = blockIdx.y, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int pos = blockIdx.x * 1024 + threadIdx.x;
    const int c = blockIdx.x * 32 + wid;                       // chunk
    const int nchunk = (prm.P + kCullChunk - 1) / kCullChunk;
    float a[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, b[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
    if (pos < prm.Ppad) {
        float* soa = prm.soa + (size_t)n * 3 * prm.Ppad;
        int src = 0;
        float v[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F};
        if (pos < prm.P) {
            src = prm.order[pos];
            const float* p = prm.aos + ((size_t)n * prm.P + src) * 3;
#pragma unroll
            for (int k = 0; k < 3; ++k) { v[k] = p[k]; a[k] = v[k]; b[k] = v[k]; }
        }
        soa[pos] = v[0]; soa[prm.Ppad + pos] = v[1]; soa[2 * prm.Ppad + pos] = v[2];
        prm.perm[(size_t)n * prm.Ppad + pos] = src;
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            a[k] = fminf(a[k], __shfl_xor_sync(0xffffffffu, a[k], o));
            b[k] = fmaxf(b[k], __shfl_xor_sync(0xffffffffu, b[k], o));
        }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            s_box[wid][k] = a[k]; s_box[wid][3 + k] = b[k];
            if (c < nchunk) {
                float* box = prm.box + (size_t)n * 6 * prm.Cpad;
                box[k * prm.Cpad + c] = a[k]; box[(3 + k) * prm.Cpad + c] = b[k];
            }
        }
    }
    __syncthreads();
    if (wid < 6 && blockIdx.x * 32 < nchunk) {                 // warp k reduces coordinate k over the 32 chunks
        float v = s_box[lane][wid];                            // chunks past the end hold the identities
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float u = __shfl_xor_sync(0xffffffffu, v, o);
            v = wid < 3 ? fminf(v, u) : fmaxf(v, u);
        }
        if (lane == 0) prm.sbox[(size_t)n * 6 * prm.Spad + wid * prm.Spad + blockIdx.x] = v;
    }
}

// ---- the whole sort of one cloud in ONE launch: a cluster of 8 CTAs per cloud -----------------------------------------
// Steps 1-6 above need a grid-wide dependency between them, i.e. six launches (plus the memsets of the histogram and the
// bounding-box cells) of a few microseconds each, all on the critical path of a training step.  For clouds of up to a few
// ten thousand points one thread-block cluster per (cloud, set) does all of it: 8 x 1024 threads hold one or two points
// each, the histogram is distributed over the eight shared memories (CTA r owns cells [r, r + 1) * ncell / 8; arrival ranks
// come from atom.shared::cluster), and barrier.cluster separates the phases.  The result equals the six-kernel path up to
// the arrival order inside a grid cell, which the kNN does not depend on.  Also clears the caller's zero regions and the
// 256-byte header of the kNN workspace (work queue, statistics, tail counter).
// grid (8 N, 2), cluster (8, 1, 1), block 1024, dynamic shared memory = 4 * ncell / 8 bytes; needs glog2 <= 4.
constexpr int kSortCluster = 8;
FB_DEVINL int dsmem_atomic_add(uint32_t cluster_addr, int v) {
    int old;
    asm volatile("atom.shared::cluster.add.u32 %0, [%1], %2;" : "=r"(old) : "r"(cluster_addr), "r"(v) : "memory");
    return old;
}
FB_DEVINL int dsmem_ld_i32(uint32_t cluster_addr) {
    int v;
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(cluster_addr) : "memory");
    return v;
}
FB_DEVINL void dsmem_st_b32(uint32_t cluster_addr, uint32_t v) {
    asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(cluster_addr), "r"(v) : "memory");
}

__global__ void __launch_bounds__(1024) cloud_sort_cluster_kernel(const SortParams prm, unsigned* __restrict__ header64, unsigned long long* dbg) {
    fb::pdl_wait_trigger();
    auto stamp = [&](int k) { if (dbg && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); dbg[400 + k] = t; } };
    stamp(0);
    extern __shared__ int s_hist[];                 // the ncell / 8 cells this CTA owns: counts, then start offsets
    __shared__ float s_red[32][6];
    __shared__ float s_bb8[kSortCluster][6];        // every CTA's partial bounding box, broadcast to every CTA
    __shared__ int s_tot8[kSortCluster];            // every CTA's cell-count total, broadcast to every CTA
    __shared__ int s_wsum[32];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int crank = (int)cluster_ctarank();
    const int ctid = crank * 1024 + tid, cthreads = kSortCluster * 1024;
    {
        const unsigned nthreads = gridDim.x * gridDim.y * 1024u;
        const unsigned t = (blockIdx.x + gridDim.x * blockIdx.y) * 1024u + tid;
#pragma unroll
        for (int k = 0; k < 2; ++k)
            if (prm.zero_ptr[k])
                for (unsigned i = t; i < prm.zero_words[k]; i += nthreads) prm.zero_ptr[k][i] = 0u;
        if (header64 && t < 64u) header64[t] = 0u;
    }
    stamp(1);
    const SortSet& S = prm.set[blockIdx.y];
    if (S.P <= 0) return;                            // the whole cluster leaves together
    const int n = blockIdx.x / kSortCluster;
    const int len = set_len(S, n);
    const int ncell = 1 << (3 * S.glog2), G = 1 << S.glog2;
    const int cells_per = ncell / kSortCluster;      // 64 or 512 (<= 1024 threads)
    const float* aos = S.aos + (size_t)n * S.P * 3;
    if (tid < cells_per) s_hist[tid] = 0;

    // 1. bounding box: local reduction, then all-to-all through distributed shared memory
    float lo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, hi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
    for (int i = ctid; i < len; i += cthreads) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { const float v = aos[(size_t)i * 3 + k]; lo[k] = fminf(lo[k], v); hi[k] = fmaxf(hi[k], v); }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { s_red[wid][k] = lo[k]; s_red[wid][3 + k] = hi[k]; }
    }
    __syncthreads();
    if (tid < 6 * kSortCluster) {
        const int k = tid % 6, r = tid / 6;          // value k of this CTA's box -> CTA r
        float v = s_red[0][k];
        for (int w = 1; w < 32; ++w) v = k < 3 ? fminf(v, s_red[w][k]) : fmaxf(v, s_red[w][k]);
        dsmem_st_b32(mapa_u32(smem_u32(&s_bb8[crank][k]), (uint32_t)r), __float_as_uint(v));
    }
    cluster_sync_all();
    CloudGrid g;
    {
        float ext = 0.f;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float a = s_bb8[0][k], b = s_bb8[0][3 + k];
#pragma unroll
            for (int r = 1; r < kSortCluster; ++r) { a = fminf(a, s_bb8[r][k]); b = fmaxf(b, s_bb8[r][3 + k]); }
            const float l = a < CUDART_INF_F ? a : 0.f, h = b > -CUDART_INF_F ? b : 0.f;    // empty / NaN-only cloud -> 0, as cloud_grid
            g.lo[k] = l;
            ext = fmaxf(ext, h - l);
        }
        g.inv_h = (ext > 0.f && ext < CUDART_INF_F) ? (float)G / ext : 0.f;
        if (!(g.inv_h < CUDART_INF_F)) g.inv_h = 0.f;
    }

    stamp(2);
    // 2. cell + arrival rank.  Histogram cell c lives in CTA c / cells_per (atom.shared::cluster).  Neighbouring points of a
    //    mesh-ordered cloud fall into the same cell, and same-address atomics serialise, so a warp first groups its lanes by
    //    cell (match.any) and issues ONE atomic per distinct cell; the lanes of a group take consecutive ranks.
    const uint32_t hist_s = smem_u32(s_hist);
    int* cellv = S.cell + (size_t)n * S.P;
    int* rankv = S.rank + (size_t)n * S.P;
    for (int it = 0; it * cthreads < len; ++it) {
        const int i = it * cthreads + ctid;
        const bool valid = i < len;
        int cell = 0x40000000 | lane;                 // a key of its own for idle lanes
        if (valid) {
            const float* p = aos + (size_t)i * 3;
            const uint32_t cx = cell_coord(p[0], g.lo[0], g.inv_h, G), cy = cell_coord(p[1], g.lo[1], g.inv_h, G),
                           cz = cell_coord(p[2], g.lo[2], g.inv_h, G);
            cell = (int)hilbert3(cx, cy, cz, S.glog2);
        }
        const unsigned m = __match_any_sync(0xffffffffu, cell);
        const int leader = __ffs(m) - 1;
        int base = 0;
        if (valid && lane == leader)
            base = dsmem_atomic_add(mapa_u32(hist_s + (uint32_t)(cell % cells_per) * 4u, (uint32_t)(cell / cells_per)), __popc(m));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (valid) {
            cellv[i] = cell;
            rankv[i] = base + __popc(m & ((1u << lane) - 1u));
        }
    }
    cluster_sync_all();

    stamp(3);
    // 3. exclusive scan: inside the CTA (one cell per thread), then the CTA totals
    {
        const int v = tid < cells_per ? s_hist[tid] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += u;
        }
        if (lane == 31) s_wsum[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            const int t = s_wsum[lane];
            int w = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += u;
            }
            s_wsum[lane] = w - t;
            if (lane == 31) {                        // w = this CTA's total -> every CTA
                for (int r = 0; r < kSortCluster; ++r) dsmem_st_b32(mapa_u32(smem_u32(&s_tot8[crank]), (uint32_t)r), (uint32_t)w);
            }
        }
        cluster_sync_all();
        int off = 0;
        for (int r = 0; r < crank; ++r) off += s_tot8[r];
        if (tid < cells_per) s_hist[tid] = off + s_wsum[wid] + inc - v;
    }
    cluster_sync_all();     // start offsets of every CTA are final before anyone reads them remotely

    stamp(4);
    // 4. scatter into curve order (start offsets read from the owning CTA, once per distinct cell of a warp); +inf padding
    float* soa = S.soa + (size_t)n * 3 * S.Ppad;
    int* perm = S.perm + (size_t)n * S.Ppad;
    for (int it = 0; it * cthreads < len; ++it) {
        const int i = it * cthreads + ctid;
        const bool valid = i < len;
        const int cell = valid ? cellv[i] : (0x40000000 | lane);
        const unsigned m = __match_any_sync(0xffffffffu, cell);
        const int leader = __ffs(m) - 1;
        int start = 0;
        if (valid && lane == leader) start = dsmem_ld_i32(mapa_u32(hist_s + (uint32_t)(cell % cells_per) * 4u, (uint32_t)(cell / cells_per)));
        start = __shfl_sync(0xffffffffu, start, leader);
        if (valid) {
            const float* p = aos + (size_t)i * 3;
            const int pos = start + rankv[i];
            soa[pos] = p[0]; soa[S.Ppad + pos] = p[1]; soa[2 * S.Ppad + pos] = p[2];
            perm[pos] = i;
        }
    }
    for (int i = len + ctid; i < S.Ppad; i += cthreads) {
        soa[i] = CUDART_INF_F; soa[S.Ppad + i] = CUDART_INF_F; soa[2 * S.Ppad + i] = CUDART_INF_F;
        perm[i] = 0;
    }
    __threadfence();
    cluster_sync_all();

    stamp(5);
    // 5. chunk boxes (warp per chunk, 8 x 32 warps)
    float* box = S.box + (size_t)n * 6 * S.Cpad;
    const int nchunk = (len + kCullChunk - 1) / kCullChunk;
    for (int c = crank * 32 + wid; c < nchunk; c += kSortCluster * 32) {
        const int i = c * kCullChunk + lane;
        float a[3], b[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float v = soa[k * S.Ppad + i];
            a[k] = i < len ? v : CUDART_INF_F;
            b[k] = i < len ? v : -CUDART_INF_F;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                a[k] = fminf(a[k], __shfl_xor_sync(0xffffffffu, a[k], o));
                b[k] = fmaxf(b[k], __shfl_xor_sync(0xffffffffu, b[k], o));
            }
        }
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { box[k * S.Cpad + c] = a[k]; box[(3 + k) * S.Cpad + c] = b[k]; }
        }
    }
    __threadfence();
    cluster_sync_all();

    stamp(6);
    // 6. super-chunk boxes (warp per 32 chunks)
    float* sbox = S.sbox + (size_t)n * 6 * S.Spad;
    for (int sc = crank * 32 + wid; sc * 32 < nchunk; sc += kSortCluster * 32) {
        const int c = sc * 32 + lane;
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            float v = k < 3 ? CUDART_INF_F : -CUDART_INF_F;
            if (c < nchunk) v = box[k * S.Cpad + c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float u = __shfl_xor_sync(0xffffffffu, v, o);
                v = k < 3 ? fminf(v, u) : fmaxf(v, u);
            }
            if (lane == 0) sbox[k * S.Spad + sc] = v;
        }
    }
    stamp(7);
}

// ---- main kernel ------------------------------------------------------------------------------------------
struct CullDir {
    const float* q_soa;   // sorted [N][3][Pq_pad]
    const int* q_perm;    // [N][Pq_pad]
    const float* t_soa;   // sorted [N][3][Pt_pad]
    const int* t_perm;    // [N][Pt_pad]
    const float* t_box;   // [N][6][Ct_pad]
    const float* t_sbox;  // [N][6][St_pad]   boxes of 32-chunk super-chunks
    const int64_t* len_q; // nullable
    const int64_t* len_t; // nullable
    unsigned long long* keys;   // [N][Pq], original query order
    int Pq, Pq_pad, Pt, Pt_pad, Ct_pad, St_pad, n_groups;
};
struct CullParams {
    CullDir dir[2];
    int ndir, N, items_dir0, total_items;
    unsigned int* counter;            // work queue
    unsigned long long* stats;        // [0] pair evaluations executed, [1] cross-chunk-tie queries, [2] max chunk evaluations of one group, [3] groups
};

FB_DEVINL float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
FB_DEVINL float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
FB_DEVINL float gap(float alo, float ahi, float blo, float bhi) {   // distance between intervals, rounded like d = q - t
    return fmaxf(0.f, fmaxf(__fsub_rn(blo, ahi), __fsub_rn(alo, bhi)));
}
FB_DEVINL void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
FB_DEVINL float lb_dist(float gx, float gy, float gz) { return __fmaf_rn(gz, gz, __fmaf_rn(gy, gy, __fmul_rn(gx, gx))); }

template <int kQ>
__global__ void __launch_bounds__(128) knn_cull_kernel(const CullParams prm) {
    fb::pdl_wait_trigger();
    __shared__ float4 s_stage[4][24];      // per warp: one target chunk (x row | y row | z row)
    const int lane = threadIdx.x & 31;
    float4* stage = s_stage[threadIdx.x >> 5];
    unsigned long long n_eval = 0, n_slow = 0, n_eval_max = 0, n_items = 0;

    for (;;) {
        int item = 0;
        if (lane == 0) item = (int)atomicAdd(prm.counter, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= prm.total_items) break;

        const int d = item >= prm.items_dir0 ? 1 : 0;
        const CullDir& D = prm.dir[d];
        const int r = item - (d ? prm.items_dir0 : 0);
        const int grp = r % D.n_groups, n = r / D.n_groups;
        const int len_q = D.len_q ? min((int)D.len_q[n], D.Pq) : D.Pq;
        const int len_t = D.len_t ? min((int)D.len_t[n], D.Pt) : D.Pt;
        const int q0 = grp * 32 * kQ;
        if (q0 >= len_q || len_t <= 0) continue;

        const float* qsoa = D.q_soa + (size_t)n * 3 * D.Pq_pad;
        const float* tsoa = D.t_soa + (size_t)n * 3 * D.Pt_pad;
        const int* tperm = D.t_perm + (size_t)n * D.Pt_pad;
        const float* box = D.t_box + (size_t)n * 6 * D.Ct_pad;
        const int nchunk = (len_t + kCullChunk - 1) / kCullChunk;

        // queries -> registers; lanes past the end of the cloud repeat the group's first query (never published)
        float fx[kQ], fy[kQ], fz[kQ], best[kQ];
        uint64_t qx[kQ], qy[kQ], qz[kQ];
        int bchunk[kQ];
        uint32_t tie = 0;
        float glo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, ghi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            int i = q0 + q * 32 + lane;
            if (i >= len_q) i = q0;
            fx[q] = __ldg(qsoa + i); fy[q] = __ldg(qsoa + D.Pq_pad + i); fz[q] = __ldg(qsoa + 2 * D.Pq_pad + i);
            qx[q] = pack2(fx[q], fx[q]); qy[q] = pack2(fy[q], fy[q]); qz[q] = pack2(fz[q], fz[q]);
            best[q] = CUDART_INF_F; bchunk[q] = 0;
            glo[0] = fminf(glo[0], fx[q]); ghi[0] = fmaxf(ghi[0], fx[q]);
            glo[1] = fminf(glo[1], fy[q]); ghi[1] = fmaxf(ghi[1], fy[q]);
            glo[2] = fminf(glo[2], fz[q]); ghi[2] = fmaxf(ghi[2], fz[q]);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) { glo[k] = warp_min(glo[k]); ghi[k] = warp_max(ghi[k]); }

        const unsigned long long n_eval_before = n_eval;
        ++n_items;
        // a chunk travels global -> one float4 per lane (24 lanes) -> the warp's shared-memory stage -> broadcast reads:
        // one LDG per chunk, issued one chunk ahead of its use
        auto load_part = [&](int c) -> float4 {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (lane < 24)
                v = __ldg(reinterpret_cast<const float4*>(tsoa + (size_t)(lane >> 3) * D.Pt_pad + (size_t)c * kCullChunk) + (lane & 7));
            return v;
        };
        auto evaluate = [&](int c, const float4& part) {
            __syncwarp();
            if (lane < 24) stage[lane] = part;
            __syncwarp();
            const float4 *sx = stage, *sy = stage + 8, *sz = stage + 16;
            float cmin[kQ];
#pragma unroll
            for (int q = 0; q < kQ; ++q) cmin[q] = CUDART_INF_F;
#pragma unroll
            for (int g = 0; g < kCullChunk / 4; ++g) {
                const float4 X = sx[g], Y = sy[g], Z = sz[g];
                const uint64_t x01 = pack2(X.x, X.y), x23 = pack2(X.z, X.w);
                const uint64_t y01 = pack2(Y.x, Y.y), y23 = pack2(Y.z, Y.w);
                const uint64_t z01 = pack2(Z.x, Z.y), z23 = pack2(Z.z, Z.w);
#pragma unroll
                for (int q = 0; q < kQ; ++q) {
                    uint64_t dx = sub2(qx[q], x01), dy = sub2(qy[q], y01), dz = sub2(qz[q], z01);
                    uint64_t dd = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
                    float lo, hi;
                    unpack2(dd, lo, hi);
                    cmin[q] = min3(cmin[q], lo, hi);
                    dx = sub2(qx[q], x23); dy = sub2(qy[q], y23); dz = sub2(qz[q], z23);
                    dd = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
                    unpack2(dd, lo, hi);
                    cmin[q] = min3(cmin[q], lo, hi);
                }
            }
#pragma unroll
            for (int q = 0; q < kQ; ++q) {
                if (cmin[q] < best[q]) { best[q] = cmin[q]; bchunk[q] = c; tie &= ~(1u << q); }
                else if (cmin[q] == best[q]) tie |= 1u << q;
            }
            ++n_eval;
        };

        // seed: nearest super-chunk (32 chunks) by box centre, then the nearest chunk inside it
        const float* sbox = D.t_sbox + (size_t)n * 6 * D.St_pad;
        const int nsuper = (nchunk + 31) / 32;
        auto nearest = [&](const float* bxs, int pad, int first, int last) -> int {
            const float cx = 0.5f * (glo[0] + ghi[0]), cy = 0.5f * (glo[1] + ghi[1]), cz = 0.5f * (glo[2] + ghi[2]);
            unsigned long long bk = ~0ull;
            for (int c = first + lane; c < last; c += 32) {
                const float dx = 0.5f * (bxs[c] + bxs[3 * pad + c]) - cx;
                const float dy = 0.5f * (bxs[pad + c] + bxs[4 * pad + c]) - cy;
                const float dz = 0.5f * (bxs[2 * pad + c] + bxs[5 * pad + c]) - cz;
                const float k2 = dx * dx + dy * dy + dz * dz;        // >= 0 or NaN
                const unsigned long long key = ((unsigned long long)__float_as_uint(k2) << 32) | (unsigned)c;
                if (k2 >= 0.f && key < bk) bk = key;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, bk, o);
                bk = other < bk ? other : bk;
            }
            return bk != ~0ull ? (int)(bk & 0xffffffffu) : first;
        };
        const int seed_super = nearest(sbox, D.St_pad, 0, nsuper);
        const int seed = nearest(box, D.Ct_pad, seed_super * 32, min(nchunk, seed_super * 32 + 32));
        evaluate(seed, load_part(seed));

        // cull passes over a two-level box hierarchy (super-chunk = 32 chunks = 1024 targets; chunk = 32 targets), one box
        // per lane at either level.  Level 1 tests the group's box: pass 0 takes the chunks whose box overlaps it (LB == 0:
        // where the neighbours almost always are), pass 1 the remaining chunks with LB <= U, U = max over the group's
        // queries of their best so far; a super-chunk is opened when its own box passes the same test.  Level 2, for the
        // level-1 survivors in ascending chunk order: every lane tests its own kQ queries against the chunk box
        // (point-to-box bound vs that query's best); the chunk is evaluated if any query may gain or tie.
        // Survivors' box rows and target rows are prefetched into L1 while the tests run.
        auto group_U = [&]() -> float {
            float mb = 0.f;
#pragma unroll
            for (int q = 0; q < kQ; ++q) mb = fmaxf(mb, best[q]);
            return warp_max(mb);
        };
#pragma unroll 1
        for (int pass = 0; pass < 2; ++pass) {
#pragma unroll 1
            for (int s0 = 0; s0 < nsuper; s0 += 32) {
                const int sidx = s0 + lane;
                float slb = CUDART_INF_F;
                if (sidx < nsuper) {
                    slb = lb_dist(gap(glo[0], ghi[0], sbox[sidx], sbox[3 * D.St_pad + sidx]),
                                  gap(glo[1], ghi[1], sbox[D.St_pad + sidx], sbox[4 * D.St_pad + sidx]),
                                  gap(glo[2], ghi[2], sbox[2 * D.St_pad + sidx], sbox[5 * D.St_pad + sidx]));
                    if (!(slb >= 0.f)) slb = 0.f;
                }
                bool scand = slb == 0.f;
                if (pass) { const float U = group_U(); scand = slb <= U; }
                uint32_t ms = __ballot_sync(0xffffffffu, scand);
                if (lane < 6) {
                    for (uint32_t t = ms; t; t &= t - 1) prefetch_l1(box + lane * D.Ct_pad + (s0 + __ffs(t) - 1) * 32);
                }
                while (ms) {
                    const int sb = __ffs(ms) - 1;
                    ms &= ms - 1;
                    const int c0 = (s0 + sb) * 32;
                    const int c = c0 + lane;
                    float lb = CUDART_INF_F;
                    float bx[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    if (c < nchunk && c != seed) {
#pragma unroll
                        for (int k = 0; k < 6; ++k) bx[k] = box[k * D.Ct_pad + c];
                        lb = lb_dist(gap(glo[0], ghi[0], bx[0], bx[3]), gap(glo[1], ghi[1], bx[1], bx[4]),
                                     gap(glo[2], ghi[2], bx[2], bx[5]));
                        if (!(lb >= 0.f)) lb = 0.f;          // NaN box -> always a candidate
                    }
                    bool cand = lb == 0.f;
                    if (pass) { const float U = group_U(); cand = lb > 0.f && lb <= U; }   // U first: all lanes shuffle
                    uint32_t m = __ballot_sync(0xffffffffu, cand);
                    float4 cur = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (m) cur = load_part(c0 + __ffs(m) - 1);
                    while (m) {
                        const int b = __ffs(m) - 1;
                        m &= m - 1;
                        float4 nxt = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (m) nxt = load_part(c0 + __ffs(m) - 1);
                        float cb[6];
#pragma unroll
                        for (int k = 0; k < 6; ++k) cb[k] = __shfl_sync(0xffffffffu, bx[k], b);
                        bool want = false;
#pragma unroll
                        for (int q = 0; q < kQ; ++q) {
                            const float lq = lb_dist(gap(fx[q], fx[q], cb[0], cb[3]), gap(fy[q], fy[q], cb[1], cb[4]),
                                                     gap(fz[q], fz[q], cb[2], cb[5]));
                            want |= !(lq > best[q]);           // NaN -> evaluate
                        }
                        if (__any_sync(0xffffffffu, want)) evaluate(c0 + b, cur);
                        cur = nxt;
                    }
                }
            }
        }

        if (n_eval - n_eval_before > n_eval_max) n_eval_max = n_eval - n_eval_before;

        // resolve the lowest original index among the targets at distance best[q], publish
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int i = q0 + q * 32 + lane;
            if (i >= len_q) continue;
            int bj = 0x7fffffff;
            auto scan_chunk = [&](int c) {
                const int j0 = c * kCullChunk;
                const float4* sx = reinterpret_cast<const float4*>(tsoa + j0);
                const float4* sy = reinterpret_cast<const float4*>(tsoa + D.Pt_pad + j0);
                const float4* sz = reinterpret_cast<const float4*>(tsoa + 2 * (size_t)D.Pt_pad + j0);
                for (int g = 0; g < kCullChunk / 4; ++g) {
                    const float4 X = __ldg(sx + g), Y = __ldg(sy + g), Z = __ldg(sz + g);
                    const float tx[4] = {X.x, X.y, X.z, X.w}, ty[4] = {Y.x, Y.y, Y.z, Y.w}, tz[4] = {Z.x, Z.y, Z.z, Z.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float dd = dist_exact(fx[q], fy[q], fz[q], tx[e], ty[e], tz[e]);
                        if (dd == best[q] && j0 + g * 4 + e < len_t) bj = min(bj, __ldg(tperm + j0 + g * 4 + e));
                    }
                }
            };
            if (!((tie >> q) & 1u)) {
                scan_chunk(bchunk[q]);
            } else {
                ++n_slow;
                for (int c = 0; c < nchunk; ++c) {
                    const float gx = gap(fx[q], fx[q], box[c], box[3 * D.Ct_pad + c]);
                    const float gy = gap(fy[q], fy[q], box[D.Ct_pad + c], box[4 * D.Ct_pad + c]);
                    const float gz = gap(fz[q], fz[q], box[2 * D.Ct_pad + c], box[5 * D.Ct_pad + c]);
                    const float lb = lb_dist(gx, gy, gz);
                    if (!(lb > best[q])) scan_chunk(c);
                }
            }
            if (bj == 0x7fffffff) bj = 0;   // no finite distance at all (NaN / inf input)
            const int orig = __ldg(D.q_perm + (size_t)n * D.Pq_pad + i);
            D.keys[(size_t)n * D.Pq + orig] =
                ((unsigned long long)__float_as_uint(best[q]) << 32) | (unsigned long long)(unsigned)bj;
        }
    }
    if (prm.stats) {
        n_slow = __reduce_add_sync(0xffffffffu, (unsigned)n_slow);
        if (lane == 0) {
            if (n_eval) atomicAdd(prm.stats, n_eval * (unsigned long long)(32 * kQ * kCullChunk));
            if (n_slow) atomicAdd(prm.stats + 1, n_slow);
            atomicMax(prm.stats + 2, n_eval_max);
            atomicAdd(prm.stats + 3, n_items);
        }
    }
}

// ---- CTA-cooperative variant ----------------------------------------------------------------------------------
// Same walk, but ONE group of 32*kQ queries is worked on by the four warps of a CTA: every warp keeps all the group's queries
// in registers, evaluates the seed chunk (a tight first bound for everyone) and then only the target chunks whose index is
// its own residue mod 4; the four partial results are merged through shared memory before the index resolution.  A launch
// of the one-warp-per-group kernel lasts as long as its slowest group (cfg2: 53 chunk evaluations against an average of 17,
// SMs 46 % busy); here the slowest group is four times shorter and the 2 560 groups are pulled from the queue by ~740
// resident CTAs, so the tail evens out.  Exactness is unchanged: a chunk is skipped by warp w only if it cannot beat or tie
// w's own best, which is >= the merged best.
template <int kQ>
__global__ void __launch_bounds__(128) knn_cull_cta_kernel(const CullParams prm) {
    fb::pdl_wait_trigger();
    __shared__ float4 s_stage[4][24];      // per warp: one target chunk (x row | y row | z row)
    __shared__ float s_best[4][kQ][32];    // per warp: its best distance / chunk / tie bits for the group's queries
    __shared__ int s_chunk[4][kQ][32];
    __shared__ uint32_t s_tie[4][32];
    __shared__ int s_item;
    const int lane = threadIdx.x & 31, wq = threadIdx.x >> 5;
    float4* stage = s_stage[wq];
    unsigned long long n_eval = 0, n_slow = 0, n_eval_max = 0, n_items = 0;

    for (;;) {
        __syncthreads();                   // the previous item's merged results have been read
        if (threadIdx.x == 0) s_item = (int)atomicAdd(prm.counter, 1u);
        __syncthreads();
        const int item = s_item;
        if (item >= prm.total_items) break;

        const int d = item >= prm.items_dir0 ? 1 : 0;
        const CullDir& D = prm.dir[d];
        const int r = item - (d ? prm.items_dir0 : 0);
        const int grp = r % D.n_groups, n = r / D.n_groups;
        const int len_q = D.len_q ? min((int)D.len_q[n], D.Pq) : D.Pq;
        const int len_t = D.len_t ? min((int)D.len_t[n], D.Pt) : D.Pt;
        const int q0 = grp * 32 * kQ;
        if (q0 >= len_q || len_t <= 0) continue;

        const float* qsoa = D.q_soa + (size_t)n * 3 * D.Pq_pad;
        const float* tsoa = D.t_soa + (size_t)n * 3 * D.Pt_pad;
        const int* tperm = D.t_perm + (size_t)n * D.Pt_pad;
        const float* box = D.t_box + (size_t)n * 6 * D.Ct_pad;
        const int nchunk = (len_t + kCullChunk - 1) / kCullChunk;

        // queries -> registers; lanes past the end of the cloud repeat the group's first query (never published)
        float fx[kQ], fy[kQ], fz[kQ], best[kQ];
        uint64_t qx[kQ], qy[kQ], qz[kQ];
        int bchunk[kQ];
        uint32_t tie = 0;
        float glo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, ghi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            int i = q0 + q * 32 + lane;
            if (i >= len_q) i = q0;
            fx[q] = __ldg(qsoa + i); fy[q] = __ldg(qsoa + D.Pq_pad + i); fz[q] = __ldg(qsoa + 2 * D.Pq_pad + i);
            qx[q] = pack2(fx[q], fx[q]); qy[q] = pack2(fy[q], fy[q]); qz[q] = pack2(fz[q], fz[q]);
            best[q] = CUDART_INF_F; bchunk[q] = 0;
            glo[0] = fminf(glo[0], fx[q]); ghi[0] = fmaxf(ghi[0], fx[q]);
            glo[1] = fminf(glo[1], fy[q]); ghi[1] = fmaxf(ghi[1], fy[q]);
            glo[2] = fminf(glo[2], fz[q]); ghi[2] = fmaxf(ghi[2], fz[q]);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) { glo[k] = warp_min(glo[k]); ghi[k] = warp_max(ghi[k]); }

        const unsigned long long n_eval_before = n_eval;
        ++n_items;
        // a chunk travels global -> one float4 per lane (24 lanes) -> the warp's shared-memory stage -> broadcast reads:
        // one LDG per chunk, issued one chunk ahead of its use
        auto load_part = [&](int c) -> float4 {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (lane < 24)
                v = __ldg(reinterpret_cast<const float4*>(tsoa + (size_t)(lane >> 3) * D.Pt_pad + (size_t)c * kCullChunk) + (lane & 7));
            return v;
        };
        auto evaluate = [&](int c, const float4& part) {
            __syncwarp();
            if (lane < 24) stage[lane] = part;
            __syncwarp();
            const float4 *sx = stage, *sy = stage + 8, *sz = stage + 16;
            float cmin[kQ];
#pragma unroll
            for (int q = 0; q < kQ; ++q) cmin[q] = CUDART_INF_F;
#pragma unroll
            for (int g = 0; g < kCullChunk / 4; ++g) {
                const float4 X = sx[g], Y = sy[g], Z = sz[g];
                const uint64_t x01 = pack2(X.x, X.y), x23 = pack2(X.z, X.w);
                const uint64_t y01 = pack2(Y.x, Y.y), y23 = pack2(Y.z, Y.w);
                const uint64_t z01 = pack2(Z.x, Z.y), z23 = pack2(Z.z, Z.w);
#pragma unroll
                for (int q = 0; q < kQ; ++q) {
                    uint64_t dx = sub2(qx[q], x01), dy = sub2(qy[q], y01), dz = sub2(qz[q], z01);
                    uint64_t dd = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
                    float lo, hi;
                    unpack2(dd, lo, hi);
                    cmin[q] = min3(cmin[q], lo, hi);
                    dx = sub2(qx[q], x23); dy = sub2(qy[q], y23); dz = sub2(qz[q], z23);
                    dd = fma2(dz, dz, fma2(dy, dy, mul2(dx, dx)));
                    unpack2(dd, lo, hi);
                    cmin[q] = min3(cmin[q], lo, hi);
                }
            }
#pragma unroll
            for (int q = 0; q < kQ; ++q) {
                if (cmin[q] < best[q]) { best[q] = cmin[q]; bchunk[q] = c; tie &= ~(1u << q); }
                else if (cmin[q] == best[q]) tie |= 1u << q;
            }
            ++n_eval;
        };

        // seed: nearest super-chunk (32 chunks) by box centre, then the nearest chunk inside it
        const float* sbox = D.t_sbox + (size_t)n * 6 * D.St_pad;
        const int nsuper = (nchunk + 31) / 32;
        auto nearest = [&](const float* bxs, int pad, int first, int last) -> int {
            const float cx = 0.5f * (glo[0] + ghi[0]), cy = 0.5f * (glo[1] + ghi[1]), cz = 0.5f * (glo[2] + ghi[2]);
            unsigned long long bk = ~0ull;
            for (int c = first + lane; c < last; c += 32) {
                const float dx = 0.5f * (bxs[c] + bxs[3 * pad + c]) - cx;
                const float dy = 0.5f * (bxs[pad + c] + bxs[4 * pad + c]) - cy;
                const float dz = 0.5f * (bxs[2 * pad + c] + bxs[5 * pad + c]) - cz;
                const float k2 = dx * dx + dy * dy + dz * dz;        // >= 0 or NaN
                const unsigned long long key = ((unsigned long long)__float_as_uint(k2) << 32) | (unsigned)c;
                if (k2 >= 0.f && key < bk) bk = key;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, bk, o);
                bk = other < bk ? other : bk;
            }
            return bk != ~0ull ? (int)(bk & 0xffffffffu) : first;
        };
        const int seed_super = nearest(sbox, D.St_pad, 0, nsuper);
        const int seed = nearest(box, D.Ct_pad, seed_super * 32, min(nchunk, seed_super * 32 + 32));
        evaluate(seed, load_part(seed));

        // cull passes over a two-level box hierarchy (super-chunk = 32 chunks = 1024 targets; chunk = 32 targets), one box
        // per lane at either level.  Level 1 tests the group's box: pass 0 takes the chunks whose box overlaps it (LB == 0:
        // where the neighbours almost always are), pass 1 the remaining chunks with LB <= U, U = max over the group's
        // queries of their best so far; a super-chunk is opened when its own box passes the same test.  Level 2, for the
        // level-1 survivors in ascending chunk order: every lane tests its own kQ queries against the chunk box
        // (point-to-box bound vs that query's best); the chunk is evaluated if any query may gain or tie.
        // Survivors' box rows and target rows are prefetched into L1 while the tests run.
        auto group_U = [&]() -> float {
            float mb = 0.f;
#pragma unroll
            for (int q = 0; q < kQ; ++q) mb = fmaxf(mb, best[q]);
            return warp_max(mb);
        };
#pragma unroll 1
        for (int pass = 0; pass < 2; ++pass) {
#pragma unroll 1
            for (int s0 = 0; s0 < nsuper; s0 += 32) {
                const int sidx = s0 + lane;
                float slb = CUDART_INF_F;
                if (sidx < nsuper) {
                    slb = lb_dist(gap(glo[0], ghi[0], sbox[sidx], sbox[3 * D.St_pad + sidx]),
                                  gap(glo[1], ghi[1], sbox[D.St_pad + sidx], sbox[4 * D.St_pad + sidx]),
                                  gap(glo[2], ghi[2], sbox[2 * D.St_pad + sidx], sbox[5 * D.St_pad + sidx]));
                    if (!(slb >= 0.f)) slb = 0.f;
                }
                bool scand = slb == 0.f;
                if (pass) { const float U = group_U(); scand = slb <= U; }
                uint32_t ms = __ballot_sync(0xffffffffu, scand);
                if (lane < 6) {
                    for (uint32_t t = ms; t; t &= t - 1) prefetch_l1(box + lane * D.Ct_pad + (s0 + __ffs(t) - 1) * 32);
                }
                while (ms) {
                    const int sb = __ffs(ms) - 1;
                    ms &= ms - 1;
                    const int c0 = (s0 + sb) * 32;
                    const int c = c0 + lane;
                    float lb = CUDART_INF_F;
                    float bx[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                    if (c < nchunk && c != seed) {
#pragma unroll
                        for (int k = 0; k < 6; ++k) bx[k] = box[k * D.Ct_pad + c];
                        lb = lb_dist(gap(glo[0], ghi[0], bx[0], bx[3]), gap(glo[1], ghi[1], bx[1], bx[4]),
                                     gap(glo[2], ghi[2], bx[2], bx[5]));
                        if (!(lb >= 0.f)) lb = 0.f;          // NaN box -> always a candidate
                    }
                    bool cand = lb == 0.f;
                    if (pass) { const float U = group_U(); cand = lb > 0.f && lb <= U; }   // U first: all lanes shuffle
                    cand = cand && ((c & 3) == wq);          // this warp's quarter of the target chunks
                    uint32_t m = __ballot_sync(0xffffffffu, cand);
                    float4 cur = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (m) cur = load_part(c0 + __ffs(m) - 1);
                    while (m) {
                        const int b = __ffs(m) - 1;
                        m &= m - 1;
                        float4 nxt = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (m) nxt = load_part(c0 + __ffs(m) - 1);
                        float cb[6];
#pragma unroll
                        for (int k = 0; k < 6; ++k) cb[k] = __shfl_sync(0xffffffffu, bx[k], b);
                        bool want = false;
#pragma unroll
                        for (int q = 0; q < kQ; ++q) {
                            const float lq = lb_dist(gap(fx[q], fx[q], cb[0], cb[3]), gap(fy[q], fy[q], cb[1], cb[4]),
                                                     gap(fz[q], fz[q], cb[2], cb[5]));
                            want |= !(lq > best[q]);           // NaN -> evaluate
                        }
                        if (__any_sync(0xffffffffu, want)) evaluate(c0 + b, cur);
                        cur = nxt;
                    }
                }
            }
        }

        if (n_eval - n_eval_before > n_eval_max) n_eval_max = n_eval - n_eval_before;

        // merge the four warps' results: the minimum distance; a tie when a warp that attains it saw one, or when two warps
        // attain it in different chunks (every warp evaluates the seed chunk, so the same chunk twice is not a tie)
#pragma unroll
        for (int q = 0; q < kQ; ++q) { s_best[wq][q][lane] = best[q]; s_chunk[wq][q][lane] = bchunk[q]; }
        s_tie[wq][lane] = tie;
        __syncthreads();
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            float m = s_best[0][q][lane];
            int mc = s_chunk[0][q][lane];
            bool t = (s_tie[0][lane] >> q) & 1u;
#pragma unroll
            for (int w2 = 1; w2 < 4; ++w2) {
                const float b2 = s_best[w2][q][lane];
                const int c2 = s_chunk[w2][q][lane];
                const bool t2 = (s_tie[w2][lane] >> q) & 1u;
                if (b2 < m) { m = b2; mc = c2; t = t2; }
                else if (b2 == m) { t = t || t2 || (c2 != mc); if (c2 < mc) mc = c2; }
            }
            best[q] = m; bchunk[q] = mc;
            tie = t ? (tie | (1u << q)) : (tie & ~(1u << q));
        }

        // resolve the lowest original index among the targets at distance best[q], publish (query slot q by warp q mod 4)
#pragma unroll
        for (int q = 0; q < kQ; ++q) {
            const int i = q0 + q * 32 + lane;
            if ((q & 3) != wq) continue;
            if (i >= len_q) continue;
            int bj = 0x7fffffff;
            auto scan_chunk = [&](int c) {
                const int j0 = c * kCullChunk;
                const float4* sx = reinterpret_cast<const float4*>(tsoa + j0);
                const float4* sy = reinterpret_cast<const float4*>(tsoa + D.Pt_pad + j0);
                const float4* sz = reinterpret_cast<const float4*>(tsoa + 2 * (size_t)D.Pt_pad + j0);
                for (int g = 0; g < kCullChunk / 4; ++g) {
                    const float4 X = __ldg(sx + g), Y = __ldg(sy + g), Z = __ldg(sz + g);
                    const float tx[4] = {X.x, X.y, X.z, X.w}, ty[4] = {Y.x, Y.y, Y.z, Y.w}, tz[4] = {Z.x, Z.y, Z.z, Z.w};
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const float dd = dist_exact(fx[q], fy[q], fz[q], tx[e], ty[e], tz[e]);
                        if (dd == best[q] && j0 + g * 4 + e < len_t) bj = min(bj, __ldg(tperm + j0 + g * 4 + e));
                    }
                }
            };
            if (!((tie >> q) & 1u)) {
                scan_chunk(bchunk[q]);
            } else {
                ++n_slow;
                for (int c = 0; c < nchunk; ++c) {
                    const float gx = gap(fx[q], fx[q], box[c], box[3 * D.Ct_pad + c]);
                    const float gy = gap(fy[q], fy[q], box[D.Ct_pad + c], box[4 * D.Ct_pad + c]);
                    const float gz = gap(fz[q], fz[q], box[2 * D.Ct_pad + c], box[5 * D.Ct_pad + c]);
                    const float lb = lb_dist(gx, gy, gz);
                    if (!(lb > best[q])) scan_chunk(c);
                }
            }
            if (bj == 0x7fffffff) bj = 0;   // no finite distance at all (NaN / inf input)
            const int orig = __ldg(D.q_perm + (size_t)n * D.Pq_pad + i);
            D.keys[(size_t)n * D.Pq + orig] =
                ((unsigned long long)__float_as_uint(best[q]) << 32) | (unsigned long long)(unsigned)bj;
        }
    }
    if (prm.stats) {
        n_slow = __reduce_add_sync(0xffffffffu, (unsigned)n_slow);
        if (lane == 0) {
            if (n_eval) atomicAdd(prm.stats, n_eval * (unsigned long long)(32 * kQ * kCullChunk));
            if (n_slow) atomicAdd(prm.stats + 1, n_slow);
            atomicMax(prm.stats + 2, n_eval_max);
            atomicAdd(prm.stats + 3, n_items);
        }
    }
}

// ---- grid search: one thread per query --------------------------------------------------------------------------------
// The same sorted clouds, read through a table of the grid's cells (cloud_lut_kernel: linear cell index -> [start, end) in the
// sorted order).  A thread owns ONE query: it visits the cell the query falls in, then the shells of cells around it, keeping
// the lexicographic minimum of (distance, original index), and stops when nothing outside the visited block can beat or tie it:
//   every point outside the block [x0..x1] x [y0..y1] x [z0..z1] lies beyond one of the block's faces that are interior to the
//   grid, so its COMPUTED distance is >= the squared gap from the query to the nearest such face — taken with a relative
//   safety margin that covers the rounding of the cell assignment (p -> floor((p - lo) * inv_h)) and of the distance itself —
// or when every point of the cloud has been visited (degenerate grids, far-away queries).  A surface cloud of 10 k points in
// a 16^3 grid has ~2.5 points per occupied cell: a query evaluates the ~30-70 points of its 27-cell neighbourhood instead of
// the ~540 of the box-hierarchy walk, there is no per-group serial chain, and 160 k independent threads hide the latency.
// Exactness: same distance expression and tie rule as everywhere else; bit-identical to the other kernels (tests).
// lower bound of |q - p| along one axis for any point p assigned to cell `cell` of that axis (0 if q may be level with it),
// with the same safety margins as the block faces below
FB_DEVINL float cell_gap(float q, float lo, float h, int cell) {
    const float a = lo + (float)cell * h * (1.0f - 0x1p-18f) - fabsf(lo) * 0x1p-20f;            // cell's points are >= a
    const float b = lo + (float)(cell + 1) * h * (1.0f + 0x1p-18f) + fabsf(lo) * 0x1p-20f;      // and < b
    return fmaxf(0.f, fmaxf(a - q, q - b));
}

struct GridDir {
    const float* q_soa;   // sorted queries [N][3][Pq_pad]
    const int* q_perm;
    const float* t_soa;   // sorted targets [N][3][Pt_pad]
    const int* t_perm;
    const int2* t_lut;    // [N][G^3] linear cell -> [start, end)
    const unsigned *t_bbmin, *t_bbmax;   // [N][3] order-encoded bounding box of the targets
    const int64_t *len_q, *len_t;
    unsigned long long* keys;            // [N][Pq] original query order
    int Pq, Pq_pad, Pt, Pt_pad, glog2, blocks_per_cloud;
};
struct GridParams { GridDir dir[2]; int ndir, N, blocks_dir0; unsigned long long* stats; };

// 6. (grid mode) linear cell table.  grid (ceil(ncell / 256), N, 2), block 256; after cloud_scan_kernel (hist = exclusive starts).
__global__ void __launch_bounds__(256) cloud_lut_kernel(const SortParams prm, int2* __restrict__ lut_x, int2* __restrict__ lut_y) {
    fb::pdl_wait_trigger();
    const SortSet& S = prm.set[blockIdx.z];
    if (S.P <= 0 || (int)blockIdx.y >= S.N) return;
    const int n = blockIdx.y, G = 1 << S.glog2, ncell = 1 << (3 * S.glog2);
    const int lin = blockIdx.x * 256 + threadIdx.x;
    if (lin >= ncell) return;
    const int cx = lin & (G - 1), cy = (lin >> S.glog2) & (G - 1), cz = lin >> (2 * S.glog2);
    const int h = (int)hilbert3((uint32_t)cx, (uint32_t)cy, (uint32_t)cz, S.glog2);
    const int* hist = S.hist + ((size_t)n << (3 * S.glog2));
    const int start = hist[h], end = (h + 1 < ncell) ? hist[h + 1] : set_len(S, n);
    (blockIdx.z ? lut_y : lut_x)[((size_t)n << (3 * S.glog2)) + lin] = make_int2(start, end);
}

__global__ void __launch_bounds__(256) knn_grid_kernel(const GridParams prm) {
    fb::pdl_wait_trigger();
    const int d = (int)blockIdx.x >= prm.blocks_dir0 ? 1 : 0;
    const GridDir& D = prm.dir[d];
    const int bl = (int)blockIdx.x - (d ? prm.blocks_dir0 : 0);
    const int n = bl / D.blocks_per_cloud;
    const int i = (bl - n * D.blocks_per_cloud) * 256 + threadIdx.x;          // sorted query index
    const int len_q = D.len_q ? min((int)D.len_q[n], D.Pq) : D.Pq;
    const int len_t = D.len_t ? min((int)D.len_t[n], D.Pt) : D.Pt;
    __shared__ unsigned long long s_eval;
    if (threadIdx.x == 0) s_eval = 0ull;
    __syncthreads();
    unsigned long long n_eval = 0;
    if (i < len_q && len_t > 0) {
    const float* qs = D.q_soa + (size_t)n * 3 * D.Pq_pad;
    const float fx = __ldg(qs + i), fy = __ldg(qs + D.Pq_pad + i), fz = __ldg(qs + 2 * D.Pq_pad + i);
    const float* ts = D.t_soa + (size_t)n * 3 * D.Pt_pad;
    const int* tperm = D.t_perm + (size_t)n * D.Pt_pad;
    const int G = 1 << D.glog2;
    const int2* lut = D.t_lut + ((size_t)n << (3 * D.glog2));
    // the targets' grid (same arithmetic as cloud_grid)
    float lo[3], ext = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const unsigned a = D.t_bbmin[n * 3 + k], b = D.t_bbmax[n * 3 + k];
        const float l = a == 0xFFFFFFFFu ? 0.f : dec_ord(a), h = b == 0u ? 0.f : dec_ord(b);
        lo[k] = l;
        ext = fmaxf(ext, h - l);
    }
    float inv_h = (ext > 0.f && ext < CUDART_INF_F) ? (float)G / ext : 0.f;
    if (!(inv_h < CUDART_INF_F)) inv_h = 0.f;
    const float hcell = inv_h > 0.f ? 1.0f / inv_h : 0.f;
    const int cx = cell_coord(fx, lo[0], inv_h, G), cy = cell_coord(fy, lo[1], inv_h, G), cz = cell_coord(fz, lo[2], inv_h, G);
    const float q[3] = {fx, fy, fz};

    float best = CUDART_INF_F;
    int bj = 0x7fffffff;
    int visited = 0;
    for (int r = 0; r < G; ++r) {
        const int x0 = max(cx - r, 0), x1 = min(cx + r, G - 1), y0 = max(cy - r, 0), y1 = min(cy + r, G - 1), z0 = max(cz - r, 0),
                  z1 = min(cz + r, G - 1);
        for (int z = z0; z <= z1; ++z)
            for (int y = y0; y <= y1; ++y) {
                const bool face = (abs(z - cz) == r) || (abs(y - cy) == r);
                // on a face row every x belongs to the shell; otherwise only the two end cells do
                // squared gap from the query to this row of cells along y and z (0 when the query is level with the row)
                float gyz = 0.f;
                if (r > 0 && inv_h > 0.f) {
                    const float gy = cell_gap(q[1], lo[1], hcell, y), gz = cell_gap(q[2], lo[2], hcell, z);
                    gyz = gy * gy + gz * gz;
                    if (gyz * (1.0f - 0x1p-18f) > best) continue;          // the whole row is farther than the best so far
                }
                for (int x = x0; x <= x1; x += (face || r == 0) ? 1 : max(x1 - x0, 1)) {
                    if (!face && abs(x - cx) != r) continue;
                    if (r > 0 && inv_h > 0.f) {                                // this cell's box cannot beat or tie the best: skip it
                        const float gx = cell_gap(q[0], lo[0], hcell, x);
                        if ((gyz + gx * gx) * (1.0f - 0x1p-18f) > best) continue;
                    }
                    const int2 se = __ldg(lut + (((z << D.glog2) + y) << D.glog2) + x);
                    for (int j = se.x; j < se.y; ++j) {
                        const float dd = dist_exact(fx, fy, fz, __ldg(ts + j), __ldg(ts + D.Pt_pad + j), __ldg(ts + 2 * D.Pt_pad + j));
                        if (dd <= best) {                       // NaN never enters
                            const int oj = __ldg(tperm + j);
                            if (dd < best || oj < bj) { best = dd; bj = oj; }
                        }
                    }
                    visited += se.y - se.x;
                    n_eval += (unsigned long long)(se.y - se.x);
                }
            }
        if (visited >= len_t) break;
        if (inv_h == 0.f) continue;                               // single-cell grid: geometry says nothing
        // lower bound of the computed distance of any point outside the block
        float lb = CUDART_INF_F;
        const int b0[3] = {x0, y0, z0}, b1[3] = {x1, y1, z1};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (b0[k] > 0) {           // points in cells < b0: p < lo + b0 * h (1 + margin)
                const float face_hi = lo[k] + (float)b0[k] * hcell * (1.0f + 0x1p-18f) + fabsf(lo[k]) * 0x1p-20f;
                const float g = q[k] - face_hi;
                lb = fminf(lb, g > 0.f ? g * g * (1.0f - 0x1p-18f) : 0.f);
            }
            if (b1[k] < G - 1) {       // points in cells > b1: p >= lo + (b1 + 1) * h (1 - margin)
                const float face_lo = lo[k] + (float)(b1[k] + 1) * hcell * (1.0f - 0x1p-18f) - fabsf(lo[k]) * 0x1p-20f;
                const float g = face_lo - q[k];
                lb = fminf(lb, g > 0.f ? g * g * (1.0f - 0x1p-18f) : 0.f);
            }
        }
        if (lb > best) break;                                     // strictly: an equal distance outside could carry a lower index
    }
    if (bj == 0x7fffffff) bj = 0;                                 // no finite distance at all (NaN / inf input)
    const int orig = __ldg(D.q_perm + (size_t)n * D.Pq_pad + i);
    D.keys[(size_t)n * D.Pq + orig] = ((unsigned long long)__float_as_uint(best) << 32) | (unsigned long long)(unsigned)bj;
    }
    if (prm.stats) {                                              // executed pair evaluations, one global atomic per block
        unsigned ne = (unsigned)n_eval;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ne += __shfl_xor_sync(0xffffffffu, ne, o);
        if ((threadIdx.x & 31) == 0 && ne) atomicAdd(&s_eval, (unsigned long long)ne);
        __syncthreads();
        if (threadIdx.x == 0 && s_eval) atomicAdd(prm.stats, s_eval);
    }
}

}  // namespace chamfer
}  // namespace fb
