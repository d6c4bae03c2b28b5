// find_b200/csrc/knn_cull.cuh — exact 1-NN with bounding-box culling (included by chamfer.cu).
//
// Same contract as the brute-force tile kernel (dist = fmaf(dz,dz,fmaf(dy,dy,dx*dx)), lowest ORIGINAL index on ties,
// bit-exact vs oracle/chamfer_oracle.c), but most (query group, target chunk) pairs are never evaluated:
//
//   1. sort_cloud_kernel: one CTA per cloud bins the points into a cubic grid (G^3 cells over the cloud's bounding
//      box, Morton cell order, counting sort in shared memory) and writes them as SoA rows padded with +inf to a
//      multiple of 32, the permutation back to original indices, and the bounding box of every 32-point chunk.
//   2. knn_cull_kernel: persistent warps.  A warp owns 32*kQ consecutive sorted queries (spatially compact), kQ per
//      lane in registers.  It evaluates the chunk nearest to the group first, then walks the chunk boxes 32 at a time
//      (one box per lane), and evaluates a chunk only if LB(group box, chunk box) <= U, U = max over the group's
//      queries of their current best distance.  Chunk evaluation is the packed-fp32x2 brute-force inner loop.
//
// Why the cull is exact: LB is computed with the SAME operation sequence as the distance, on the per-axis gaps
// g = max(0, tlo - qhi, qlo - thi).  Round-to-nearest subtraction, multiplication and fma are monotone, so for every
// query in the group box and every target in the chunk box the COMPUTED fp32 distance is >= LB.  A chunk is skipped
// only when LB > U >= best[q] for all q, so it holds neither a closer point nor a tie.  Ties across chunks (equal
// distance, lower original index elsewhere) set a per-query flag in the hot loop and take a slow exact path.
#pragma once
#include "common.cuh"
#include <math_constants.h>

namespace fb {
namespace chamfer {

constexpr int kCullChunk = 32;
constexpr int kSortThreads = 1024;

struct SortSet {
    const float* aos;       // [N][P][3]
    const int64_t* len;     // nullable
    float* soa;             // [N][3][Ppad]
    int* perm;              // [N][Ppad]   sorted position -> original index
    float* box;             // [N][6][Cpad]  lo xyz, hi xyz per 32-point chunk
    int P, Ppad, Cpad, glog2;
};
struct SortParams { SortSet set[2]; };

FB_DEVINL uint32_t spread3(uint32_t v) {   // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
FB_DEVINL int cell_coord(float p, float lo, float inv_h, int G) {
    const float f = (p - lo) * inv_h;
    int c = 0;
    if (f >= 0.f) c = f < (float)G ? (int)f : G - 1;   // NaN -> 0
    return c;
}

// grid (N, nsets), block kSortThreads, dynamic smem = (1 << 3*glog2) * 4 bytes
__global__ void __launch_bounds__(kSortThreads) sort_cloud_kernel(const SortParams prm) {
    extern __shared__ int s_hist[];
    __shared__ float s_red[32][6];
    __shared__ float s_bb[6];
    __shared__ int s_wsum[32];

    const SortSet& S = prm.set[blockIdx.y];
    const int n = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    if (S.P <= 0) return;
    int len = S.len ? (int)S.len[n] : S.P;
    len = len < 0 ? 0 : (len > S.P ? S.P : len);
    const float* aos = S.aos + (size_t)n * S.P * 3;
    float* soa = S.soa + (size_t)n * 3 * S.Ppad;
    int* perm = S.perm + (size_t)n * S.Ppad;
    float* box = S.box + (size_t)n * 6 * S.Cpad;
    const int G = 1 << S.glog2, ncell = 1 << (3 * S.glog2);

    // 1. bounding box of the cloud (NaNs ignored by fminf/fmaxf)
    float lo[3] = {CUDART_INF_F, CUDART_INF_F, CUDART_INF_F}, hi[3] = {-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F};
    for (int i = tid; i < len; i += kSortThreads) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { const float v = aos[(size_t)i * 3 + k]; lo[k] = fminf(lo[k], v); hi[k] = fmaxf(hi[k], v); }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k)
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) { s_red[wid][k] = lo[k]; s_red[wid][3 + k] = hi[k]; }
    }
    for (int c = tid; c < ncell; c += kSortThreads) s_hist[c] = 0;
    __syncthreads();
    if (wid == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float a = s_red[lane][k], b = s_red[lane][3 + k];
            for (int o = 16; o > 0; o >>= 1) {
                a = fminf(a, __shfl_xor_sync(0xffffffffu, a, o));
                b = fmaxf(b, __shfl_xor_sync(0xffffffffu, b, o));
            }
            if (lane == 0) { s_bb[k] = a; s_bb[3 + k] = b; }
        }
    }
    __syncthreads();
    const float blo[3] = {s_bb[0], s_bb[1], s_bb[2]};
    float ext = fmaxf(fmaxf(s_bb[3] - s_bb[0], s_bb[4] - s_bb[1]), s_bb[5] - s_bb[2]);
    float inv_h = (ext > 0.f && ext < CUDART_INF_F) ? (float)G / ext : 0.f;   // cubic cells; degenerate cloud -> one cell
    if (!(inv_h < CUDART_INF_F)) inv_h = 0.f;

    auto cell_of = [&](int i) -> int {
        const float x = aos[(size_t)i * 3], y = aos[(size_t)i * 3 + 1], z = aos[(size_t)i * 3 + 2];
        const uint32_t cx = cell_coord(x, blo[0], inv_h, G), cy = cell_coord(y, blo[1], inv_h, G),
                       cz = cell_coord(z, blo[2], inv_h, G);
        return (int)(spread3(cx) | (spread3(cy) << 1) | (spread3(cz) << 2));
    };

    // 2. histogram
    for (int i = tid; i < len; i += kSortThreads) atomicAdd(&s_hist[cell_of(i)], 1);
    __syncthreads();

    // 3. exclusive scan over cells (each thread owns `per` consecutive cells)
    {
        const int per = ncell >= kSortThreads ? ncell / kSortThreads : 1;
        const int c0 = tid * per;
        int sum = 0;
        if (c0 < ncell)
            for (int j = 0; j < per; ++j) sum += s_hist[c0 + j];
        int inc = sum;
        for (int o = 1; o < 32; o <<= 1) {
            const int v = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += v;
        }
        if (lane == 31) s_wsum[wid] = inc;
        __syncthreads();
        if (wid == 0) {
            int w = s_wsum[lane];
            for (int o = 1; o < 32; o <<= 1) {
                const int v = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += v;
            }
            s_wsum[lane] = w;   // inclusive over warps
        }
        __syncthreads();
        int run = inc - sum + (wid ? s_wsum[wid - 1] : 0);
        if (c0 < ncell)
            for (int j = 0; j < per; ++j) { const int h = s_hist[c0 + j]; s_hist[c0 + j] = run; run += h; }
    }
    __syncthreads();

    // 4. scatter into cell order (order inside a cell is arbitrary; results do not depend on it)
    for (int i = tid; i < len; i += kSortThreads) {
        const int pos = atomicAdd(&s_hist[cell_of(i)], 1);
        soa[pos] = aos[(size_t)i * 3];
        soa[S.Ppad + pos] = aos[(size_t)i * 3 + 1];
        soa[2 * S.Ppad + pos] = aos[(size_t)i * 3 + 2];
        perm[pos] = i;
    }
    for (int i = len + tid; i < S.Ppad; i += kSortThreads) {
        soa[i] = CUDART_INF_F; soa[S.Ppad + i] = CUDART_INF_F; soa[2 * S.Ppad + i] = CUDART_INF_F;
        perm[i] = 0;
    }
    __syncthreads();

    // 5. chunk boxes (plain loads: written by this block just above)
    const int nchunk = (len + kCullChunk - 1) / kCullChunk;
    for (int c = wid; c < nchunk; c += kSortThreads / 32) {
        const int i = c * kCullChunk + lane;
        float v[3], a[3], b[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            v[k] = soa[k * S.Ppad + i];
            a[k] = i < len ? v[k] : CUDART_INF_F;
            b[k] = i < len ? v[k] : -CUDART_INF_F;
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
}

// ---- main kernel ------------------------------------------------------------------------------------------
struct CullDir {
    const float* q_soa;   // sorted [N][3][Pq_pad]
    const int* q_perm;    // [N][Pq_pad]
    const float* t_soa;   // sorted [N][3][Pt_pad]
    const int* t_perm;    // [N][Pt_pad]
    const float* t_box;   // [N][6][Ct_pad]
    const int64_t* len_q; // nullable
    const int64_t* len_t; // nullable
    unsigned long long* keys;   // [N][Pq], original query order
    int Pq, Pq_pad, Pt, Pt_pad, Ct_pad, n_groups;
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
FB_DEVINL float lb_dist(float gx, float gy, float gz) { return __fmaf_rn(gz, gz, __fmaf_rn(gy, gy, __fmul_rn(gx, gx))); }

template <int kQ>
__global__ void __launch_bounds__(128) knn_cull_kernel(const CullParams prm) {
    const int lane = threadIdx.x & 31;
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
        float U = CUDART_INF_F;
        auto evaluate = [&](int c) {
            const float4* sx = reinterpret_cast<const float4*>(tsoa + (size_t)c * kCullChunk);
            const float4* sy = reinterpret_cast<const float4*>(tsoa + D.Pt_pad + (size_t)c * kCullChunk);
            const float4* sz = reinterpret_cast<const float4*>(tsoa + 2 * (size_t)D.Pt_pad + (size_t)c * kCullChunk);
            float cmin[kQ];
#pragma unroll
            for (int q = 0; q < kQ; ++q) cmin[q] = CUDART_INF_F;
#pragma unroll
            for (int g = 0; g < kCullChunk / 4; ++g) {
                const float4 X = __ldg(sx + g), Y = __ldg(sy + g), Z = __ldg(sz + g);
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
            float m = 0.f;
#pragma unroll
            for (int q = 0; q < kQ; ++q) {
                if (cmin[q] < best[q]) { best[q] = cmin[q]; bchunk[q] = c; tie &= ~(1u << q); }
                else if (cmin[q] == best[q]) tie |= 1u << q;
                m = fmaxf(m, best[q]);
            }
            U = warp_max(m);
            ++n_eval;
        };

        // seed: the chunk whose box centre is nearest to the group's centre
        int seed = 0;
        {
            const float cx = 0.5f * (glo[0] + ghi[0]), cy = 0.5f * (glo[1] + ghi[1]), cz = 0.5f * (glo[2] + ghi[2]);
            unsigned long long bk = ~0ull;
            for (int c = lane; c < nchunk; c += 32) {
                const float dx = 0.5f * (box[c] + box[3 * D.Ct_pad + c]) - cx;
                const float dy = 0.5f * (box[D.Ct_pad + c] + box[4 * D.Ct_pad + c]) - cy;
                const float dz = 0.5f * (box[2 * D.Ct_pad + c] + box[5 * D.Ct_pad + c]) - cz;
                const float k2 = dx * dx + dy * dy + dz * dz;        // >= 0 or NaN
                const unsigned long long key = ((unsigned long long)__float_as_uint(k2) << 32) | (unsigned)c;
                if (k2 >= 0.f && key < bk) bk = key;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, bk, o);
                bk = other < bk ? other : bk;
            }
            if (bk != ~0ull) seed = (int)(bk & 0xffffffffu);
        }
        evaluate(seed);

        // cull pass: one chunk box per lane.  Level 1: group box vs chunk box against U (one test per chunk, lane
        // parallel).  Level 2, for the survivors in ascending chunk order: every lane tests its own kQ queries against
        // the chunk box (point-to-box bound vs that query's best); the chunk is evaluated if any query may gain or tie.
        for (int c0 = 0; c0 < nchunk; c0 += 32) {
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
            uint32_t m = __ballot_sync(0xffffffffu, lb <= U);
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                const float lbc = __shfl_sync(0xffffffffu, lb, b);
                if (lbc > U) continue;
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
                if (__any_sync(0xffffffffu, want)) evaluate(c0 + b);
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

}  // namespace chamfer
}  // namespace fb
