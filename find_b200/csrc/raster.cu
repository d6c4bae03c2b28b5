// find_b200/csrc/raster.cu — soft-silhouette rasteriser (forward + backward), the first slice of SURVEY.md §8(f) row 4.
//
// Replaces, for the silhouette used by FIND's 2-D losses (/root/reference/src/model/renderer.py:122-128, 208-245, 309-311),
// PyTorch3D's rasterize_meshes (faces_per_pixel = 100, blur_radius = log(1/1e-4 - 1) * sigma, clipped + perspective-correct
// barycentrics) followed by SoftSilhouetteShader:  alpha = 1 - prod_k (1 - sigmoid(-signed_dist_k / sigma))  over the K
// candidates of smallest depth.  The arithmetic is restated in oracle/raster_oracle.py (header there lists the rules).
//
// Design (no graphics API: plain CUDA):
//   face_setup_kernel     one thread per (mesh x view, face): the 9 vertex floats + face id (48 B record) and the blur-grown
//                         bounding box (float4); a degenerate or behind-the-camera face gets an empty box
//   sil_tile_fwd_kernel   persistent CTAs (4 per SM) pull 16 x 16 pixel tiles from a device queue (boundary tiles cost ~10x a
//                         background tile); one thread per pixel, a warp covers 8 x 4 pixels.  The CTA streams the mesh's boxes
//                         (1024 per round, two barriers), compacts the faces that touch the tile IN FACE ORDER into shared memory
//                         and evaluates them for its 256 pixels from shared memory.  No global binning pass.
//                         K truncation is exact without a per-pixel heap: every candidate's depth bits and blend factor go to a
//                         per-CTA store [cap][256]; a pixel with more than K candidates finds its K-th smallest depth by a 16-ary
//                         search on the float bit pattern over its own list and multiplies the kept factors (earlier face on
//                         ties).  A list longer than the capacity falls back to a block-wide search that re-evaluates the tile's
//                         faces.  The same walk yields the nearest candidate (pix_to_face / zbuf) and, in the <kFragments = true>
//                         instantiation used for hard rasterisation, its barycentrics and signed distance.
//   sil_tile_bwd_kernel   the same tile walk; d alpha / d d2_k = +-prod p_k / sigma needs only the saved product and selection.
//                         Per face the 256 pixels' contributions are reduced in the warp before 6 atomics per (warp, face).
//   feat_tile_fwd / feat_tile_bwd   FeatureShader (the reference's softmax_blend over the silhouette fragments), same walk; the
//                         backward sums a warp's contributions to a face with a halving butterfly before the atomics.
//   vert_normals / phong_shade (+ _bwd)   Phong image of the hard fragments and its hand-differentiated backward.
//   points_render_kernel  the keypoint overlay (PointsRenderer + AlphaCompositor), forward only (end of this file).
// Bound: latency / issue (divergent point-triangle evaluations, ~18 of 32 lanes active; profiles/r1_silhouette.md); DRAM
// traffic (the candidate store) is ~160 GB/s.
#include "common.cuh"
#include "../../include/find_b200.h"

namespace fb {
namespace raster {

constexpr int kTile = 16;
constexpr int kThreads = kTile * kTile;
constexpr int kListCap = 4096;           // faces touching one tile that the CTA keeps for the K-truncation passes
constexpr float kEps = 1e-8f;
constexpr float kZClip = 0.01f;          // znear / 2 of the reference's cameras (renderer.py:258, 228-236)

struct FaceRec {                          // 48 B, three float4 loads
    float x0, y0, z0, x1;
    float y1, z1, x2, y2;
    float z2, pad0, pad1, pad2;
};

struct SilParams {
    const float4* box;                    // [N, F]  (xmin, xmax, ymin, ymax) grown by sqrt(blur_radius); empty when invalid
    const FaceRec* rec;                   // [N, F]
    int N, F, H, W;
    float blur_radius, sigma;
    int K, perspective;
    float* alpha;                         // [N, H, W]
    float* prod;                          // [N, H, W]   prod_k (1 - p_k) of the kept candidates
    uint2* sel;                           // [N, H, W]   (depth-threshold bits, ties allowed); (0xffffffff, 0) = all kept
    int* pix_to_face;                     // [N, H, W]   optional: nearest candidate face of the pixel (per-mesh index), -1 = none
    float* zbuf;                          // [N, H, W]   optional: its depth, -1 = none
    float* bary;                          // [N, H, W, 3] optional: its barycentrics, -1 = none
    float* dists;                         // [N, H, W]   optional: its signed squared distance (negative inside), -1 = none
    int clip;                             // clip + renormalise the barycentrics (PyTorch3D: blur_radius > 0)
    const float* grad_alpha;              // backward
    float* grad_verts;                    // [N, V, 3]
    const int* faces;                     // [F, 3] (backward: vertex ids)
    int V;
    uint32_t* cand_bits;                  // [grid][cap][256]  per-pixel candidate depths of the tile a CTA is working on
    float* cand_fac;                      // [grid][cap][256]  their blend factors 1 - p
    int cap;
    unsigned int* tile_counter;           // device: next tile of the forward kernel's queue (zeroed per call)
    unsigned int* overflow;               // device flag: a tile exceeded kListCap while a pixel needed K truncation
};

__global__ void face_setup_kernel(const float* __restrict__ verts, const int* __restrict__ faces, int N, int V, int F,
                                  float blur, float4* __restrict__ box, FaceRec* __restrict__ rec) {
    pdl_wait_trigger();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)N * F) return;
    const int n = (int)(i / F), f = (int)(i - (long long)n * F);
    const int i0 = faces[3 * f], i1 = faces[3 * f + 1], i2 = faces[3 * f + 2];
    const float* v = verts + (size_t)n * V * 3;
    FaceRec r;
    r.x0 = v[3 * i0]; r.y0 = v[3 * i0 + 1]; r.z0 = v[3 * i0 + 2];
    r.x1 = v[3 * i1]; r.y1 = v[3 * i1 + 1]; r.z1 = v[3 * i1 + 2];
    r.x2 = v[3 * i2]; r.y2 = v[3 * i2 + 1]; r.z2 = v[3 * i2 + 2];
    r.pad0 = __int_as_float(f);              // face id travels with the record (pix_to_face output)
    r.pad1 = r.pad2 = 0.f;
    const float area = __fsub_rn(__fmul_rn(r.x2 - r.x0, r.y1 - r.y0), __fmul_rn(r.y2 - r.y0, r.x1 - r.x0));
    // PyTorch3D clips faces against z_clip = znear / 2 = 0.01 (renderer.py:228-236, znear = 0.02 renderer.py:258); here a face with
    // ANY vertex nearer than that is culled whole (conservative, no new vertices): a face that straddles the camera plane would
    // otherwise keep wrapped x/z, y/z coordinates and rasterise as garbage.  Feet seen from 0.1-0.35 m never come that close.
    const float zmin = fminf(fminf(r.z0, r.z1), r.z2);
    const bool ok = (fabsf(area) > kEps) && (zmin >= kZClip);  // NaN coordinates fail both tests
    float4 b;
    if (ok) {
        b.x = fminf(fminf(r.x0, r.x1), r.x2) - blur; b.y = fmaxf(fmaxf(r.x0, r.x1), r.x2) + blur;
        b.z = fminf(fminf(r.y0, r.y1), r.y2) - blur; b.w = fmaxf(fmaxf(r.y0, r.y1), r.y2) + blur;
    } else {
        b = make_float4(1e30f, -1e30f, 1e30f, -1e30f);
    }
    box[i] = b;
    rec[i] = r;
}

FB_DEVINL float edge_fn(float px, float py, float ax, float ay, float bx, float by) {
    return __fsub_rn(__fmul_rn(px - ax, by - ay), __fmul_rn(py - ay, bx - ax));
}

// squared distance from p to segment ab; also the clamped parameter t (t = 2 marks the degenerate-segment branch)
FB_DEVINL float seg_dist2(float px, float py, float ax, float ay, float bx, float by, float& t_out) {
    const float bax = bx - ax, bay = by - ay;
    const float l2 = __fadd_rn(__fmul_rn(bax, bax), __fmul_rn(bay, bay));
    if (l2 <= kEps) {
        t_out = 2.f;
        const float dx = px - bx, dy = py - by;
        return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
    }
    float t = __fdiv_rn(__fadd_rn(__fmul_rn(bax, px - ax), __fmul_rn(bay, py - ay)), l2);
    t = fminf(fmaxf(t, 0.f), 1.f);
    t_out = t;
    const float dx = px - __fadd_rn(ax, __fmul_rn(t, bax)), dy = py - __fadd_rn(ay, __fmul_rn(t, bay));
    return __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
}

struct Eval {
    bool cand;          // candidate of this pixel
    bool inside;
    float d2;           // squared distance to the nearest edge
    float pz;           // interpolated depth (>= 0 for a candidate)
    float w[3], wsum;   // barycentrics the depth was interpolated with, before the division by wsum (1 when not clipped)
    int edge;           // nearest edge: 0 = v0v1, 1 = v0v2, 2 = v1v2
    float t;            // its clamped parameter
};

FB_DEVINL Eval eval_face(const FaceRec& r, const float4& bx, float px, float py, float blur_radius, int perspective, int clip = 1) {
    Eval e;
    e.cand = false;
    // bounding-box test of THIS pixel against the grown box of face_setup_kernel (the tile test only said the face touches the tile)
    if (!(px >= bx.x && px <= bx.y && py >= bx.z && py <= bx.w)) return e;
    const float area = edge_fn(r.x2, r.y2, r.x0, r.y0, r.x1, r.y1);
    const float den = area + kEps;
    const float e0 = edge_fn(px, py, r.x1, r.y1, r.x2, r.y2), e1 = edge_fn(px, py, r.x2, r.y2, r.x0, r.y0), e2 = edge_fn(px, py, r.x0, r.y0, r.x1, r.y1);
    float t01, t02, t12;
    const float d01 = seg_dist2(px, py, r.x0, r.y0, r.x1, r.y1, t01);
    const float d02 = seg_dist2(px, py, r.x0, r.y0, r.x2, r.y2, t02);
    const float d12 = seg_dist2(px, py, r.x1, r.y1, r.x2, r.y2, t12);
    e.d2 = d01; e.edge = 0; e.t = t01;
    if (d02 < e.d2) { e.d2 = d02; e.edge = 1; e.t = t02; }
    if (d12 < e.d2) { e.d2 = d12; e.edge = 2; e.t = t12; }
    // w_i = e_i / den > 0 needs e_i and den of the same sign: when that already fails and the pixel is beyond the blur
    // radius, the face cannot be a candidate and the divisions below are skipped (the decision itself is unchanged)
    const bool maybe_inside = (den > 0.f) ? (e0 > 0.f && e1 > 0.f && e2 > 0.f) : (e0 < 0.f && e1 < 0.f && e2 < 0.f);
    if (!maybe_inside && !(e.d2 < blur_radius)) return e;
    const float w0 = __fdiv_rn(e0, den), w1 = __fdiv_rn(e1, den), w2 = __fdiv_rn(e2, den);
    e.inside = (w0 > 0.f) && (w1 > 0.f) && (w2 > 0.f);
    float b0 = w0, b1 = w1, b2 = w2;
    if (perspective) {
        const float t0 = __fmul_rn(__fmul_rn(w0, r.z1), r.z2), t1 = __fmul_rn(__fmul_rn(r.z0, w1), r.z2), t2 = __fmul_rn(__fmul_rn(r.z0, r.z1), w2);
        const float dn = fmaxf(__fadd_rn(__fadd_rn(t0, t1), t2), kEps);
        b0 = __fdiv_rn(t0, dn); b1 = __fdiv_rn(t1, dn); b2 = __fdiv_rn(t2, dn);
    }
    if (clip) {
        const float c0 = fminf(fmaxf(b0, 0.f), 1.f), c1 = fminf(fmaxf(b1, 0.f), 1.f), c2 = fminf(fmaxf(b2, 0.f), 1.f);
        const float cs = fmaxf(__fadd_rn(__fadd_rn(c0, c1), c2), 1e-5f);
        e.pz = __fdiv_rn(__fadd_rn(__fadd_rn(__fmul_rn(c0, r.z0), __fmul_rn(c1, r.z1)), __fmul_rn(c2, r.z2)), cs);
        e.w[0] = c0; e.w[1] = c1; e.w[2] = c2; e.wsum = cs;
    } else {
        e.pz = __fadd_rn(__fadd_rn(__fmul_rn(b0, r.z0), __fmul_rn(b1, r.z1)), __fmul_rn(b2, r.z2));
        e.w[0] = b0; e.w[1] = b1; e.w[2] = b2; e.wsum = 1.f;
    }
    if (!(e.pz >= 0.f)) return e;
    e.cand = e.inside || (e.d2 < blur_radius);
    return e;
}

FB_DEVINL float sigmoidf_(float x) { return __fdiv_rn(1.f, 1.f + __expf(-x)); }

struct TileShared {
    int ids[kListCap];
    FaceRec rec[kThreads];
    float4 box[kThreads];
    int vid[kThreads][3];
    int wcnt[4 * kThreads / 32];
    int next_tile;
    int nlist;
    int overflow;
};

// Faces of mesh n whose grown box touches the tile, in face order, into sh.ids.  kSub x 256 boxes are loaded per round (independent
// loads, two barriers per round).  On overflow of the list the caller's `spill` functor consumes the full list (evaluate what is
// there) and the list restarts.
constexpr int kSub = 4;
template <typename Spill>
FB_DEVINL void build_tile_list(const SilParams& P, int n, float txmin, float txmax, float tymin, float tymax, TileShared& sh, Spill spill) {
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int kWarps = kThreads / 32;
    if (tid == 0) sh.overflow = 0;
    int nl = 0;                                      // uniform across the CTA
    const float4* box = P.box + (size_t)n * P.F;
    for (int base = 0; base < P.F; base += kThreads * kSub) {
        float4 b[kSub];
#pragma unroll
        for (int u = 0; u < kSub; ++u) {
            const int f = base + u * kThreads + tid;
            b[u] = (f < P.F) ? __ldg(box + f) : make_float4(1e30f, -1e30f, 1e30f, -1e30f);
        }
        unsigned m[kSub];
        bool hit[kSub];
#pragma unroll
        for (int u = 0; u < kSub; ++u) {
            hit[u] = (b[u].x <= txmax) && (b[u].y >= txmin) && (b[u].z <= tymax) && (b[u].w >= tymin);
            m[u] = __ballot_sync(0xffffffffu, hit[u]);
            if (lane == 0) sh.wcnt[u * kWarps + warp] = __popc(m[u]);
        }
        __syncthreads();
        int before[kSub], total = 0;
#pragma unroll
        for (int u = 0; u < kSub; ++u) {
            before[u] = total;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) { const int c = sh.wcnt[u * kWarps + w]; if (w < warp) before[u] += c; total += c; }
        }
        if (nl + total > kListCap) {                 // uniform branch: spill what we have, restart the list
            spill(nl);
            if (tid == 0) sh.overflow = 1;
            nl = 0;
        }
#pragma unroll
        for (int u = 0; u < kSub; ++u)
            if (hit[u]) sh.ids[nl + before[u] + __popc(m[u] & ((1u << lane) - 1u))] = base + u * kThreads + tid;
        nl += total;
        __syncthreads();
    }
    if (tid == 0) sh.nlist = nl;
    __syncthreads();
}

// Evaluate the first `count` faces of the tile list for this thread's pixel: fn(j_in_list, face_id, rec, eval).
template <bool kNeedVid, typename Fn>
FB_DEVINL void for_each_listed(const SilParams& P, int n, int count, float px, float py, float blur, bool active, TileShared& sh, Fn fn,
                               int clip = 1) {
    const int tid = threadIdx.x;
    const FaceRec* rec = P.rec + (size_t)n * P.F;
    const bool warp_active = __any_sync(0xffffffffu, active);      // a warp with no pixel to serve only helps staging
    for (int base = 0; base < count; base += kThreads) {
        __syncthreads();
        const int m = min(kThreads, count - base);
        if (tid < m) {
            const int f = sh.ids[base + tid];
            const float4* src = reinterpret_cast<const float4*>(rec + f);
            float4* dst = reinterpret_cast<float4*>(&sh.rec[tid]);
            dst[0] = __ldg(src); dst[1] = __ldg(src + 1); dst[2] = __ldg(src + 2);
            sh.box[tid] = __ldg(P.box + (size_t)n * P.F + f);
            if (kNeedVid) { sh.vid[tid][0] = P.faces[3 * f]; sh.vid[tid][1] = P.faces[3 * f + 1]; sh.vid[tid][2] = P.faces[3 * f + 2]; }
        }
        __syncthreads();
        if (warp_active)
            for (int j = 0; j < m; ++j) {
                Eval e;
                e.cand = false;
                if (active) e = eval_face(sh.rec[j], sh.box[j], px, py, P.blur_radius, P.perspective, clip);
                fn(j, sh.rec[j], e);
            }
    }
    __syncthreads();
}

FB_DEVINL uint32_t depth_bits(float pz) { return __float_as_uint(pz) & 0x7fffffffu; }

// kFragments: also keep the nearest candidate's barycentrics and signed distance (hard rasterisation); the silhouette
// instantiation does not carry those registers.
template <bool kFragments>
__global__ void __launch_bounds__(kThreads, 4) sil_tile_fwd_kernel(const __grid_constant__ SilParams P) {
    pdl_wait_trigger();
    __shared__ TileShared sh;
    const int tid = threadIdx.x;
    const int tiles_x = (P.W + kTile - 1) / kTile, tiles_y = (P.H + kTile - 1) / kTile;
    const long long n_tiles = (long long)P.N * tiles_x * tiles_y;
    const float blur = sqrtf(P.blur_radius), inv_sigma = __fdiv_rn(1.f, P.sigma);
    // this CTA's slice of the candidate store: [cap][256 pixels] depth bits and blend factors, written once per candidate in the
    // first walk so that a pixel with more than K candidates selects its K nearest from its own list (no re-evaluation)
    uint32_t* cbits = P.cand_bits + (size_t)blockIdx.x * P.cap * kThreads;
    float* cfac = P.cand_fac + (size_t)blockIdx.x * P.cap * kThreads;
    for (;;) {
    // dynamic tile queue: tiles on the silhouette boundary cost 10x a background tile, so a static stride leaves a long tail
    __syncthreads();
    if (tid == 0) sh.next_tile = (int)atomicAdd(P.tile_counter, 1u);
    __syncthreads();
    const long long tile = sh.next_tile;
    if (tile >= n_tiles) break;
    const int n = (int)(tile / (tiles_x * tiles_y)), trem = (int)(tile - (long long)n * tiles_x * tiles_y);
    const int tile_y = trem / tiles_x, tile_x = trem - tile_y * tiles_x;
    // a warp covers 8 x 4 pixels (not 16 x 2): fewer faces touch a warp's footprint, so fewer divergent evaluations
    const int col = tile_x * kTile + ((tid >> 5) & 1) * 8 + (tid & 7), row = tile_y * kTile + (tid >> 6) * 4 + ((tid >> 3) & 3);
    const bool in_img = (col < P.W) && (row < P.H);
    const float px = 1.f - __fdiv_rn((float)(2 * col + 1), (float)P.W), py = 1.f - __fdiv_rn((float)(2 * row + 1), (float)P.H);
    const int c_hi = min(tile_x * kTile + kTile - 1, P.W - 1), r_hi = min(tile_y * kTile + kTile - 1, P.H - 1);
    const float txmax = 1.f - __fdiv_rn((float)(2 * tile_x * kTile + 1), (float)P.W), txmin = 1.f - __fdiv_rn((float)(2 * c_hi + 1), (float)P.W);
    const float tymax = 1.f - __fdiv_rn((float)(2 * tile_y * kTile + 1), (float)P.H), tymin = 1.f - __fdiv_rn((float)(2 * r_hi + 1), (float)P.H);

    int cnt = 0;
    float prod = 1.f;
    uint2 sel = make_uint2(0xffffffffu, 0u);
    uint32_t zlo = 0xffffffffu, zhi = 0u;           // depth range of this pixel's candidates (bit patterns)
    int near_face = -1;
    float near_w0 = -1.f, near_w1 = -1.f, near_w2 = -1.f, near_ws = 1.f, near_d = -1.f;
    auto accumulate = [&](int, const FaceRec& r, const Eval& e) {
        if (e.cand) {
            const float fac = 1.f - sigmoidf_((e.inside ? e.d2 : -e.d2) * inv_sigma);
            const uint32_t b = depth_bits(e.pz);
            if (cnt < P.cap && P.alpha) { cbits[(size_t)cnt * kThreads + tid] = b; cfac[(size_t)cnt * kThreads + tid] = fac; }
            ++cnt;
            prod = __fmul_rn(prod, fac);
            if (b < zlo) {                                        // strict: the earlier face wins a depth tie (faces arrive in order)
                near_face = __float_as_int(r.pad0);
                if (kFragments) { near_w0 = e.w[0]; near_w1 = e.w[1]; near_w2 = e.w[2]; near_ws = e.wsum; near_d = e.inside ? -e.d2 : e.d2; }
            }
            zlo = min(zlo, b); zhi = max(zhi, b);
        }
    };
    build_tile_list(P, n, txmin, txmax, tymin, tymax, sh,
                    [&](int count) { for_each_listed<false>(P, n, count, px, py, blur, in_img, sh, accumulate, kFragments ? P.clip : 1); });
    const int nlist = sh.nlist;
    const bool spilled = sh.overflow != 0;
    for_each_listed<false>(P, n, nlist, px, py, blur, in_img, sh, accumulate, kFragments ? P.clip : 1);   // the silhouette always clips

    const bool sat_any = in_img && P.alpha && cnt > P.K;
    if (sat_any && cnt <= P.cap) {
        // thread-local: K-th smallest depth of the stored list by 16-ary search on the bit pattern, then the product of the kept
        uint32_t lo = zlo, hi = zhi;
        int below = 0;
        while (lo < hi) {
            const uint32_t step = ((hi - lo) >> 4) + 1u;
            uint32_t t[15];
            int c[15];
#pragma unroll
            for (int i = 0; i < 15; ++i) {
                const unsigned long long v = (unsigned long long)lo + (unsigned long long)step * (i + 1) - 1ull;
                t[i] = v > hi ? hi : (uint32_t)v;
                c[i] = 0;
            }
            for (int q = 0; q < cnt; ++q) {
                const uint32_t b = cbits[(size_t)q * kThreads + tid];
                if (b < lo || b > hi) continue;
#pragma unroll
                for (int i = 0; i < 15; ++i) c[i] += (b <= t[i]) ? 1 : 0;
            }
            int k = 15;
#pragma unroll
            for (int i = 14; i >= 0; --i) if (below + c[i] >= P.K) k = i;
            uint32_t nlo = lo, nhi = hi;
            int add = 0;
#pragma unroll
            for (int i = 0; i < 15; ++i) {
                if (k == i) nhi = t[i];
                if (k == i + 1) { nlo = t[i] + 1u; add = c[i]; }
            }
            lo = nlo; hi = nhi; below += add;
        }
        const int ties = P.K - below;
        int seen = 0;
        float prod2 = 1.f;
        for (int q = 0; q < cnt; ++q) {
            const uint32_t b = cbits[(size_t)q * kThreads + tid];
            const bool keep = (b < lo) || (b == lo && seen++ < ties);
            if (keep) prod2 = __fmul_rn(prod2, cfac[(size_t)q * kThreads + tid]);
        }
        prod = prod2; sel = make_uint2(lo, (uint32_t)ties);
    }
    const bool sat = sat_any && cnt > P.cap;             // list did not fit: block-wide search by re-evaluation
    if (__syncthreads_or(sat)) {
        if (spilled) {
            if (sat) atomicExch(P.overflow, 1u);     // cannot truncate exactly: the host reports it (findb200_silhouette_status)
        } else {
            // K-th smallest depth of this pixel's candidates: 16-ary search on the bit pattern (non-negative floats order as
            // uints) inside [zlo, zhi], re-evaluating the tile's faces each round instead of storing per-pixel lists.
            // Invariant: #(bits < lo) = below < K <= #(bits <= hi).
            uint32_t lo = zlo, hi = zhi;
            int below = 0;
            bool searching = sat && lo < hi;
            while (__syncthreads_or(searching)) {
                const uint32_t step = ((hi - lo) >> 4) + 1u;
                uint32_t t[15];
                int c[15];
#pragma unroll
                for (int i = 0; i < 15; ++i) {
                    const unsigned long long v = (unsigned long long)lo + (unsigned long long)step * (i + 1) - 1ull;
                    t[i] = v > hi ? hi : (uint32_t)v;
                    c[i] = 0;
                }
                for_each_listed<false>(P, n, nlist, px, py, blur, searching, sh, [&](int, const FaceRec&, const Eval& e) {
                    if (!e.cand) return;
                    const uint32_t b = depth_bits(e.pz);
                    if (b < lo || b > hi) return;
#pragma unroll
                    for (int i = 0; i < 15; ++i) c[i] += (b <= t[i]) ? 1 : 0;
                });
                if (searching) {
                    int k = 15;
#pragma unroll
                    for (int i = 14; i >= 0; --i) if (below + c[i] >= P.K) k = i;
                    uint32_t nlo = lo, nhi = hi;
                    int add = 0;
#pragma unroll
                    for (int i = 0; i < 15; ++i) {
                        if (k == i) nhi = t[i];
                        if (k == i + 1) { nlo = t[i] + 1u; add = c[i]; }
                    }
                    lo = nlo; hi = nhi; below += add;
                    searching = lo < hi;
                }
            }
            const uint32_t thr = lo;
            const int ties = P.K - below;
            int seen = 0;
            float prod2 = 1.f;
            for_each_listed<false>(P, n, nlist, px, py, blur, sat, sh, [&](int, const FaceRec&, const Eval& e) {
                if (!e.cand) return;
                const uint32_t b = depth_bits(e.pz);
                const bool keep = (b < thr) || (b == thr && seen++ < ties);
                if (keep) prod2 = __fmul_rn(prod2, 1.f - sigmoidf_((e.inside ? e.d2 : -e.d2) * inv_sigma));
            });
            if (sat) { prod = prod2; sel = make_uint2(thr, (uint32_t)ties); }
        }
    }
    if (in_img) {
        const size_t o = ((size_t)n * P.H + row) * P.W + col;
        if (P.alpha) P.alpha[o] = 1.f - prod;
        if (P.prod) P.prod[o] = prod;
        if (P.sel) P.sel[o] = sel;
        if (P.pix_to_face) P.pix_to_face[o] = near_face;
        if (P.zbuf) P.zbuf[o] = near_face >= 0 ? __uint_as_float(zlo) : -1.f;
        if (kFragments && P.bary) {
            const bool hit = near_face >= 0 && near_ws != 1.f;
            P.bary[3 * o] = hit ? __fdiv_rn(near_w0, near_ws) : near_w0;
            P.bary[3 * o + 1] = hit ? __fdiv_rn(near_w1, near_ws) : near_w1;
            P.bary[3 * o + 2] = hit ? __fdiv_rn(near_w2, near_ws) : near_w2;
        }
        if (kFragments && P.dists) P.dists[o] = near_d;
    }
    __syncthreads();
    }   // tile loop
}

__global__ void __launch_bounds__(kThreads, 4) sil_tile_bwd_kernel(const __grid_constant__ SilParams P) {
    pdl_wait_trigger();
    __shared__ TileShared sh;
    const int n = blockIdx.z, tid = threadIdx.x, lane = tid & 31;
    const int col = blockIdx.x * kTile + ((tid >> 5) & 1) * 8 + (tid & 7), row = blockIdx.y * kTile + (tid >> 6) * 4 + ((tid >> 3) & 3);
    const bool in_img = (col < P.W) && (row < P.H);
    const float px = 1.f - __fdiv_rn((float)(2 * col + 1), (float)P.W), py = 1.f - __fdiv_rn((float)(2 * row + 1), (float)P.H);
    const int c_hi = min(blockIdx.x * kTile + kTile - 1, P.W - 1), r_hi = min(blockIdx.y * kTile + kTile - 1, P.H - 1);
    const float txmax = 1.f - __fdiv_rn((float)(2 * blockIdx.x * kTile + 1), (float)P.W), txmin = 1.f - __fdiv_rn((float)(2 * c_hi + 1), (float)P.W);
    const float tymax = 1.f - __fdiv_rn((float)(2 * blockIdx.y * kTile + 1), (float)P.H), tymin = 1.f - __fdiv_rn((float)(2 * r_hi + 1), (float)P.H);
    const float blur = sqrtf(P.blur_radius), inv_sigma = __fdiv_rn(1.f, P.sigma);

    float coef = 0.f;                      // grad_alpha * prod / sigma of this pixel
    uint2 sel = make_uint2(0xffffffffu, 0u);
    if (in_img) {
        const size_t o = ((size_t)n * P.H + row) * P.W + col;
        coef = P.grad_alpha[o] * P.prod[o] * inv_sigma;
        sel = P.sel[o];
    }
    // a tile none of whose pixels carries gradient (far inside or far outside the silhouette) skips the walk
    if (!__syncthreads_or(coef != 0.f)) return;
    const bool active = in_img && coef != 0.f;
    int seen = 0;
    float* gv = P.grad_verts + (size_t)n * P.V * 3;
    auto scatter = [&](int j, const FaceRec& r, const Eval& e) {
        float g[3][2] = {{0.f, 0.f}, {0.f, 0.f}, {0.f, 0.f}};
        bool contributes = false;
        if (e.cand) {
            const uint32_t b = depth_bits(e.pz);
            const bool keep = (sel.x == 0xffffffffu) || (b < sel.x) || (b == sel.x && seen++ < (int)sel.y);
            if (keep) {
                const float p = sigmoidf_((e.inside ? e.d2 : -e.d2) * inv_sigma);
                const float c = (e.inside ? coef : -coef) * p;          // d alpha / d d2
                if (c != 0.f) {
                    const int ia = (e.edge == 2) ? 1 : 0, ib = (e.edge == 0) ? 1 : 2;
                    const float ax = ia == 0 ? r.x0 : r.x1, ay = ia == 0 ? r.y0 : r.y1;
                    const float bx = ib == 1 ? r.x1 : r.x2, by = ib == 1 ? r.y1 : r.y2;
                    const float t = (e.t > 1.5f) ? 1.f : e.t;            // degenerate segment: distance to b
                    const float rx = px - (ax + t * (bx - ax)), ry = py - (ay + t * (by - ay));
                    const float ga = -2.f * (1.f - t) * c, gb = -2.f * t * c;
                    g[ia][0] = ga * rx; g[ia][1] = ga * ry;
                    g[ib][0] = gb * rx; g[ib][1] = gb * ry;
                    contributes = true;
                }
            }
        }
        if (__ballot_sync(0xffffffffu, contributes) == 0u) return;
#pragma unroll
        for (int k = 0; k < 3; ++k)
#pragma unroll
            for (int d = 0; d < 2; ++d) {
                float v = g[k][d];
#pragma unroll
                for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
                if (lane == 0 && v != 0.f) atomicAdd(gv + (size_t)sh.vid[j][k] * 3 + d, v);
            }
    };
    build_tile_list(P, n, txmin, txmax, tymin, tymax, sh,
                    [&](int count) { for_each_listed<true>(P, n, count, px, py, blur, active, sh, scatter); });
    for_each_listed<true>(P, n, sh.nlist, px, py, blur, active, sh, scatter);
}

// ---- FeatureShader: the reference's own softmax_blend over the silhouette fragments (renderer.py:23-105) -----------------
// out_c = (sum_i w_i f_ic + delta bg_c) / (sum_i w_i + delta) over the K kept candidates i of a pixel,
//   w_i = sigmoid(-sd_i / sigma) * exp((zinv_i - max zinv) / gamma),  zinv = (zfar - z) / (zfar - znear)  (softmax_blend's defaults
//   znear = 1, zfar = 100: the reference passes none), delta = max(exp((eps - max zinv) / gamma), eps), f_ic = the per-vertex
//   features interpolated with the CLIPPED perspective-correct barycentrics (TexturesVertex.sample_textures on those fragments).
// Both kernels re-walk the tile's faces like the silhouette backward and keep the same K candidates through the saved (depth
// threshold, ties) pair; the nearest kept depth (max zinv) is the forward's zbuf.
struct FeatParams {
    SilParams s;                          // box / rec / sizes / blur / sigma / perspective; s.sel = saved selection; s.faces
    const float* zbuf;                    // [N, H, W] nearest kept depth, -1 = none
    const int* p2f;                       // [N, H, W] its face (backward: the candidate that carries the max-zinv term)
    const float* feat;                    // [M, V, C]
    int C, views, c0;                     // channels [c0, c0 + 8) in this launch (forward)
    float gamma, znear, zfar;
    const float* bg;                      // [C] or null (zeros)
    float* out;                           // [N, H, W, C]
    float* den;                           // [N, H, W]  sum w + delta
    const float* grad_out;                // backward
    float* g_feat;                        // [M, V, C]  +=
    float* g_ndc;                         // [N, V, 3]  +=
};
constexpr int kFeatCh = 8;

__global__ void __launch_bounds__(kThreads, 4) feat_tile_fwd_kernel(const __grid_constant__ FeatParams Q) {
    pdl_wait_trigger();
    __shared__ TileShared sh;
    const SilParams& P = Q.s;
    const int n = blockIdx.z, tid = threadIdx.x;
    const int col = blockIdx.x * kTile + ((tid >> 5) & 1) * 8 + (tid & 7), row = blockIdx.y * kTile + (tid >> 6) * 4 + ((tid >> 3) & 3);
    const bool in_img = (col < P.W) && (row < P.H);
    const float px = 1.f - __fdiv_rn((float)(2 * col + 1), (float)P.W), py = 1.f - __fdiv_rn((float)(2 * row + 1), (float)P.H);
    const int c_hi = min(blockIdx.x * kTile + kTile - 1, P.W - 1), r_hi = min(blockIdx.y * kTile + kTile - 1, P.H - 1);
    const float txmax = 1.f - __fdiv_rn((float)(2 * blockIdx.x * kTile + 1), (float)P.W), txmin = 1.f - __fdiv_rn((float)(2 * c_hi + 1), (float)P.W);
    const float tymax = 1.f - __fdiv_rn((float)(2 * blockIdx.y * kTile + 1), (float)P.H), tymin = 1.f - __fdiv_rn((float)(2 * r_hi + 1), (float)P.H);
    const float blur = sqrtf(P.blur_radius), inv_sigma = __fdiv_rn(1.f, P.sigma);
    const float eps = 1e-10f, zr = Q.zfar - Q.znear;
    const size_t o = in_img ? ((size_t)n * P.H + row) * P.W + col : 0;
    uint2 sel = make_uint2(0xffffffffu, 0u);
    float zmax = eps;
    if (in_img) {
        sel = P.sel[o];
        const float zb = Q.zbuf[o];
        if (zb >= 0.f) zmax = fmaxf((Q.zfar - zb) / zr, eps);
    }
    const float* feat = Q.feat + (size_t)(n / Q.views) * P.V * Q.C;
    const int nch = min(kFeatCh, Q.C - Q.c0);
    float acc[kFeatCh], wsum = 0.f;
#pragma unroll
    for (int c = 0; c < kFeatCh; ++c) acc[c] = 0.f;
    int seen = 0;
    auto blend = [&](int j, const FaceRec&, const Eval& e) {
        if (!e.cand) return;
        const uint32_t b = depth_bits(e.pz);
        const bool keep = (sel.x == 0xffffffffu) || (b < sel.x) || (b == sel.x && seen++ < (int)sel.y);
        if (!keep) return;
        const float prob = sigmoidf_((e.inside ? e.d2 : -e.d2) * inv_sigma);
        const float w = prob * expf(((Q.zfar - e.pz) / zr - zmax) / Q.gamma);
        wsum += w;
        const float u0 = e.w[0] / e.wsum, u1 = e.w[1] / e.wsum, u2 = e.w[2] / e.wsum;
        const float* f0 = feat + (size_t)sh.vid[j][0] * Q.C + Q.c0;
        const float* f1 = feat + (size_t)sh.vid[j][1] * Q.C + Q.c0;
        const float* f2 = feat + (size_t)sh.vid[j][2] * Q.C + Q.c0;
#pragma unroll
        for (int c = 0; c < kFeatCh; ++c)
            if (c < nch) acc[c] += w * (u0 * __ldg(f0 + c) + u1 * __ldg(f1 + c) + u2 * __ldg(f2 + c));
    };
    build_tile_list(P, n, txmin, txmax, tymin, tymax, sh, [&](int count) { for_each_listed<true>(P, n, count, px, py, blur, in_img, sh, blend); });
    for_each_listed<true>(P, n, sh.nlist, px, py, blur, in_img, sh, blend);
    if (in_img) {
        const float delta = fmaxf(expf((eps - zmax) / Q.gamma), eps);
        const float den = wsum + delta;
        if (Q.c0 == 0) Q.den[o] = den;
        for (int c = 0; c < nch; ++c) Q.out[o * Q.C + Q.c0 + c] = (acc[c] + delta * (Q.bg ? Q.bg[Q.c0 + c] : 0.f)) / den;
    }
}

// d/d(NDC vertices of one face) from d/d(final barycentrics u), d/d(interpolated depth) and d/d(signed squared edge distance),
// for the rasteriser's forward as eval_face computes it (clip + renormalise when `clip`, perspective correction when `persp`).
FB_DEVINL void face_geometry_bwd(const FaceRec& r, float px, float py, int persp, int clip, const float (&g_u)[3], float g_z, float g_sd,
                                 float (&gX)[3], float (&gY)[3], float (&gZ)[3]) {
    const float X[3] = {r.x0, r.x1, r.x2}, Y[3] = {r.y0, r.y1, r.y2}, Z[3] = {r.z0, r.z1, r.z2};
#pragma unroll
    for (int k = 0; k < 3; ++k) { gX[k] = 0.f; gY[k] = 0.f; gZ[k] = 0.f; }
    const float area = edge_fn(X[2], Y[2], X[0], Y[0], X[1], Y[1]);
    const float dena = area + kEps;
    const float e0 = edge_fn(px, py, X[1], Y[1], X[2], Y[2]), e1 = edge_fn(px, py, X[2], Y[2], X[0], Y[0]), e2 = edge_fn(px, py, X[0], Y[0], X[1], Y[1]);
    const float ws[3] = {e0 / dena, e1 / dena, e2 / dena};
    float b[3] = {ws[0], ws[1], ws[2]};
    float t[3] = {0.f, 0.f, 0.f}, dn_raw = 1.f;
    if (persp) {
        t[0] = ws[0] * Z[1] * Z[2]; t[1] = Z[0] * ws[1] * Z[2]; t[2] = Z[0] * Z[1] * ws[2];
        dn_raw = t[0] + t[1] + t[2];
        const float dn = fmaxf(dn_raw, kEps);
        b[0] = t[0] / dn; b[1] = t[1] / dn; b[2] = t[2] / dn;
    }
    // final barycentrics u and depth z
    float g_b[3];
    if (clip) {
        float c[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) c[k] = fminf(fmaxf(b[k], 0.f), 1.f);
        const float cs_raw = c[0] + c[1] + c[2];
        const float cs = fmaxf(cs_raw, 1e-5f);
        const float z = (c[0] * Z[0] + c[1] * Z[1] + c[2] * Z[2]) / cs;
        const float u[3] = {c[0] / cs, c[1] / cs, c[2] / cs};
        float g_c[3];
        if (cs_raw > 1e-5f) {
            const float gu_dot = g_u[0] * u[0] + g_u[1] * u[1] + g_u[2] * u[2];
#pragma unroll
            for (int k = 0; k < 3; ++k) g_c[k] = (g_u[k] - gu_dot) / cs + g_z * (Z[k] - z) / cs;
        } else {
#pragma unroll
            for (int k = 0; k < 3; ++k) g_c[k] = g_u[k] / cs + g_z * Z[k] / cs;
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            gZ[k] += g_z * u[k];
            g_b[k] = (b[k] > 0.f && b[k] < 1.f) ? g_c[k] : 0.f;
        }
    } else {
#pragma unroll
        for (int k = 0; k < 3; ++k) { g_b[k] = g_u[k] + g_z * Z[k]; gZ[k] += g_z * b[k]; }
    }
    float g_ws[3];
    if (persp) {
        if (dn_raw > kEps) {
            const float dn = dn_raw;
            const float bsum = g_b[0] * b[0] + g_b[1] * b[1] + g_b[2] * b[2];
            const float g_t0 = (g_b[0] - bsum) / dn, g_t1 = (g_b[1] - bsum) / dn, g_t2 = (g_b[2] - bsum) / dn;
            g_ws[0] = g_t0 * Z[1] * Z[2]; g_ws[1] = g_t1 * Z[0] * Z[2]; g_ws[2] = g_t2 * Z[0] * Z[1];
            gZ[0] += g_t1 * ws[1] * Z[2] + g_t2 * Z[1] * ws[2];
            gZ[1] += g_t0 * ws[0] * Z[2] + g_t2 * Z[0] * ws[2];
            gZ[2] += g_t0 * ws[0] * Z[1] + g_t1 * Z[0] * ws[1];
        } else {
            const float dn = kEps;
            g_ws[0] = g_b[0] / dn * Z[1] * Z[2]; g_ws[1] = g_b[1] / dn * Z[0] * Z[2]; g_ws[2] = g_b[2] / dn * Z[0] * Z[1];
            gZ[0] += g_b[1] / dn * ws[1] * Z[2] + g_b[2] / dn * Z[1] * ws[2];
            gZ[1] += g_b[0] / dn * ws[0] * Z[2] + g_b[2] / dn * Z[0] * ws[2];
            gZ[2] += g_b[0] / dn * ws[0] * Z[1] + g_b[1] / dn * Z[0] * ws[1];
        }
    } else {
#pragma unroll
        for (int k = 0; k < 3; ++k) g_ws[k] = g_b[k];
    }
    const float g_e[3] = {g_ws[0] / dena, g_ws[1] / dena, g_ws[2] / dena};
    const float g_area = -(g_ws[0] * ws[0] + g_ws[1] * ws[1] + g_ws[2] * ws[2]) / dena;
    auto edge_bwd = [&](float ge, int a, int bb) {
        gX[a] += ge * (py - Y[bb]); gY[a] += ge * (X[bb] - px);
        gX[bb] += ge * -(py - Y[a]); gY[bb] += ge * (px - X[a]);
    };
    edge_bwd(g_e[0], 1, 2);
    edge_bwd(g_e[1], 2, 0);
    edge_bwd(g_e[2], 0, 1);
    gX[0] += g_area * (Y[2] - Y[1]); gY[0] += g_area * (X[1] - X[2]);
    gX[1] += g_area * -(Y[2] - Y[0]); gY[1] += g_area * (X[2] - X[0]);
    gX[2] += g_area * (Y[1] - Y[0]); gY[2] += g_area * -(X[1] - X[0]);
    if (g_sd != 0.f) {
        float tt[3];
        const float d01 = seg_dist2(px, py, X[0], Y[0], X[1], Y[1], tt[0]);
        const float d02 = seg_dist2(px, py, X[0], Y[0], X[2], Y[2], tt[1]);
        const float d12 = seg_dist2(px, py, X[1], Y[1], X[2], Y[2], tt[2]);
        int ed = 0;
        float dmin = d01;
        if (d02 < dmin) { dmin = d02; ed = 1; }
        if (d12 < dmin) { dmin = d12; ed = 2; }
        const int ia = ed == 2 ? 1 : 0, ib = ed == 0 ? 1 : 2;
        const bool inside = (ws[0] > 0.f) && (ws[1] > 0.f) && (ws[2] > 0.f);
        const float gs = inside ? -g_sd : g_sd;
        const float tp = tt[ed];
        if (tp == 2.f) {
            gX[ib] += gs * -2.f * (px - X[ib]); gY[ib] += gs * -2.f * (py - Y[ib]);
        } else {
            const float qx = X[ia] + tp * (X[ib] - X[ia]), qy = Y[ia] + tp * (Y[ib] - Y[ia]);
            const float cx = px - qx, cy = py - qy;
            gX[ia] += gs * -2.f * (1.f - tp) * cx; gY[ia] += gs * -2.f * (1.f - tp) * cy;
            gX[ib] += gs * -2.f * tp * cx; gY[ib] += gs * -2.f * tp * cy;
        }
    }
}

__global__ void __launch_bounds__(kThreads, 2) feat_tile_bwd_kernel(const __grid_constant__ FeatParams Q) {
    pdl_wait_trigger();
    __shared__ TileShared sh;
    const SilParams& P = Q.s;
    const int n = blockIdx.z, tid = threadIdx.x;
    const int col = blockIdx.x * kTile + ((tid >> 5) & 1) * 8 + (tid & 7), row = blockIdx.y * kTile + (tid >> 6) * 4 + ((tid >> 3) & 3);
    const bool in_img = (col < P.W) && (row < P.H);
    const float px = 1.f - __fdiv_rn((float)(2 * col + 1), (float)P.W), py = 1.f - __fdiv_rn((float)(2 * row + 1), (float)P.H);
    const int c_hi = min(blockIdx.x * kTile + kTile - 1, P.W - 1), r_hi = min(blockIdx.y * kTile + kTile - 1, P.H - 1);
    const float txmax = 1.f - __fdiv_rn((float)(2 * blockIdx.x * kTile + 1), (float)P.W), txmin = 1.f - __fdiv_rn((float)(2 * c_hi + 1), (float)P.W);
    const float tymax = 1.f - __fdiv_rn((float)(2 * blockIdx.y * kTile + 1), (float)P.H), tymin = 1.f - __fdiv_rn((float)(2 * r_hi + 1), (float)P.H);
    const float blur = sqrtf(P.blur_radius), inv_sigma = __fdiv_rn(1.f, P.sigma);
    const float eps = 1e-10f, zr = Q.zfar - Q.znear;
    const size_t o = in_img ? ((size_t)n * P.H + row) * P.W + col : 0;
    uint2 sel = make_uint2(0xffffffffu, 0u);
    float zmax = eps, den = 1.f, g_zmax = 0.f;
    int near = -1;
    bool active = false;
    if (in_img) {
        sel = P.sel[o];
        const float zb = Q.zbuf[o];
        near = Q.p2f[o];
        den = Q.den[o];
        const bool hit = zb >= 0.f;
        if (hit) zmax = fmaxf((Q.zfar - zb) / zr, eps);
        // gradient carried by max zinv: -(1/gamma) sum_i gw_i w_i plus delta's own term.  sum_i gw_i w_i collapses to
        // delta * sum_c g_c (out_c - bg_c) / den, and d(delta)/d(zmax) = -delta / gamma when delta is not clamped:
        //   G = -(delta / gamma) * sum_c g_c (out_c - bg_c) / den + [delta > eps] * (-(delta / gamma)) * sum_c g_c (bg_c - out_c) / den
        // so with an unclamped delta the two cancel exactly, and with the clamp only the first stays.
        float dot = 0.f, gsum = 0.f;
        for (int c = 0; c < Q.C; ++c) {
            const float g = Q.grad_out[o * Q.C + c];
            gsum += fabsf(g);
            dot += g * (Q.out[o * Q.C + c] - (Q.bg ? Q.bg[c] : 0.f));
        }
        active = hit && gsum != 0.f;
        const float dl_raw = expf((eps - zmax) / Q.gamma);
        const float delta = fmaxf(dl_raw, eps);
        if (hit && (Q.zfar - zb) / zr > eps) g_zmax = (dl_raw > eps) ? 0.f : -(delta / Q.gamma) * dot / den;
    }
    if (!__syncthreads_or(active)) return;
    const float* feat = Q.feat + (size_t)(n / Q.views) * P.V * Q.C;
    float* g_feat = Q.g_feat + (size_t)(n / Q.views) * P.V * Q.C;
    float* g_ndc = Q.g_ndc + (size_t)n * P.V * 3;
    const float* gout = Q.grad_out + o * Q.C;
    const float* outp = Q.out + o * Q.C;
    int seen = 0;
    const int lane = tid & 31;
    // Every lane of a warp sees the same face j at the same time (for_each_listed is warp-uniform), so the contributions of the
    // warp's 32 pixels to that face's three vertices are summed in the warp before they go to memory: up to 3 C + 9 atomics per
    // (warp, face) instead of per (pixel, face).  The 16-value halving butterfly below leaves value i of the 16 on the lanes
    // whose upper four bits spell i (31 shuffles per 16 values instead of 80).
    auto scatter = [&](int j, const FaceRec& r, const Eval& e) {
        bool contributes = false;
        float wd = 0.f, u[3] = {0.f, 0.f, 0.f};
        float gnd[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) gnd[i] = 0.f;
        if (e.cand) {
            const uint32_t b = depth_bits(e.pz);
            const bool keep = (sel.x == 0xffffffffu) || (b < sel.x) || (b == sel.x && seen++ < (int)sel.y);
            if (keep && active) {
                contributes = true;
                const float prob = sigmoidf_((e.inside ? e.d2 : -e.d2) * inv_sigma);
                const float E = expf(((Q.zfar - e.pz) / zr - zmax) / Q.gamma);
                const float w = prob * E;
                u[0] = e.w[0] / e.wsum; u[1] = e.w[1] / e.wsum; u[2] = e.w[2] / e.wsum;
                const int v0 = sh.vid[j][0], v1 = sh.vid[j][1], v2 = sh.vid[j][2];
                const float *f0 = feat + (size_t)v0 * Q.C, *f1 = feat + (size_t)v1 * Q.C, *f2 = feat + (size_t)v2 * Q.C;
                float gw = 0.f, g_u[3] = {0.f, 0.f, 0.f};
                wd = w / den;
                for (int c = 0; c < Q.C; ++c) {
                    const float a0 = __ldg(f0 + c), a1 = __ldg(f1 + c), a2 = __ldg(f2 + c);
                    const float g = gout[c];
                    gw += g * ((u[0] * a0 + u[1] * a1 + u[2] * a2) - outp[c]);
                    const float gf = g * wd;                                 // d/d(interpolated feature c)
                    g_u[0] += gf * a0; g_u[1] += gf * a1; g_u[2] += gf * a2;
                }
                gw /= den;
                // w = prob * E:  d/d(prob) = gw E;  d/d(zinv) = gw w / gamma (+ the max-zinv term on the nearest face)
                const float g_prob = gw * E;
                float g_zinv = gw * w / Q.gamma;
                if (__float_as_int(r.pad0) == near && depth_bits(e.pz) == depth_bits(Q.zbuf[o])) g_zinv += g_zmax - 0.f;
                // every kept candidate's exponent also holds -zmax: that part is inside g_zmax (collapsed sum), see above
                const float g_z = -g_zinv / zr;
                const float g_sd = g_prob * (-prob * (1.f - prob) * inv_sigma);      // prob = sigmoid(-sd / sigma)
                float gX[3], gY[3], gZ[3];
                face_geometry_bwd(r, px, py, P.perspective, 1, g_u, g_z, g_sd, gX, gY, gZ);
#pragma unroll
                for (int k = 0; k < 3; ++k) { gnd[3 * k] = gX[k]; gnd[3 * k + 1] = gY[k]; gnd[3 * k + 2] = gZ[k]; }
            }
        }
        if (__ballot_sync(0xffffffffu, contributes) == 0u) return;
        const int vv[3] = {sh.vid[j][0], sh.vid[j][1], sh.vid[j][2]};
        // v[0..15] of every lane -> the sum over the warp of v[idx], idx = lane >> 1, on both lanes of the pair
        auto reduce16 = [&](float (&v)[16]) -> float {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const bool hi = lane & 16;
                const float send = hi ? v[i] : v[i + 8], keepv = hi ? v[i + 8] : v[i];
                v[i] = keepv + __shfl_xor_sync(0xffffffffu, send, 16);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const bool hi = lane & 8;
                const float send = hi ? v[i] : v[i + 4], keepv = hi ? v[i + 4] : v[i];
                v[i] = keepv + __shfl_xor_sync(0xffffffffu, send, 8);
            }
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const bool hi = lane & 4;
                const float send = hi ? v[i] : v[i + 2], keepv = hi ? v[i + 2] : v[i];
                v[i] = keepv + __shfl_xor_sync(0xffffffffu, send, 4);
            }
            {
                const bool hi = lane & 2;
                const float send = hi ? v[0] : v[1], keepv = hi ? v[1] : v[0];
                v[0] = keepv + __shfl_xor_sync(0xffffffffu, send, 2);
            }
            return v[0] + __shfl_xor_sync(0xffffffffu, v[0], 1);
        };
        const int idx = lane >> 1;
        const bool writer = (lane & 1) == 0;
        {
            const float t = reduce16(gnd);                            // 9 of the 16: (vertex k, axis d) = idx / 3, idx % 3
            if (writer && idx < 9 && t != 0.f) atomicAdd(g_ndc + (size_t)vv[idx / 3] * 3 + idx % 3, t);
        }
        for (int c0 = 0; c0 < Q.C; c0 += 16) {
            float gf[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) gf[i] = (contributes && c0 + i < Q.C) ? gout[c0 + i] * wd : 0.f;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                float v[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) v[i] = u[k] * gf[i];
                const float t = reduce16(v);
                if (writer && c0 + idx < Q.C && t != 0.f) atomicAdd(g_feat + (size_t)vv[k] * Q.C + c0 + idx, t);
            }
        }
    };
    build_tile_list(P, n, txmin, txmax, tymin, tymax, sh, [&](int count) { for_each_listed<true>(P, n, count, px, py, blur, in_img, sh, scatter); });
    for_each_listed<true>(P, n, sh.nlist, px, py, blur, in_img, sh, scatter);
}

}  // namespace raster
}  // namespace fb

using namespace fb;
using namespace fb::raster;

namespace {
struct WsLayout { size_t box, rec, flag, cbits, cfac, total; };
constexpr int kFwdCtas = 148 * 4;             // persistent CTAs of the forward kernel (4 per SM: 64 registers, 36 KB shared)
int prefer_shared() {                         // 4 x 36 KB of static shared memory per SM needs the larger carve-out
    static int rc = [] {
        cudaError_t e = cudaFuncSetAttribute(sil_tile_fwd_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(sil_tile_fwd_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(sil_tile_bwd_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
        return (int)e;
    }();
    return rc;
}
int cand_cap() {
    const char* e = getenv("FINDB200_SIL_CAP");          // tests force the re-evaluation path with a small capacity
    const int c = e ? atoi(e) : 384;
    return c < 1 ? 1 : (c > 1024 ? 1024 : c);
}
WsLayout ws_layout(int N, int F) {
    WsLayout L;
    size_t o = 0;
    L.box = o; o += ((size_t)N * F * sizeof(float4) + 255) & ~(size_t)255;
    L.rec = o; o += ((size_t)N * F * sizeof(FaceRec) + 255) & ~(size_t)255;
    L.flag = o; o += 256;
    L.cbits = o; o += (size_t)kFwdCtas * cand_cap() * kThreads * sizeof(uint32_t);
    L.cfac = o; o += (size_t)kFwdCtas * cand_cap() * kThreads * sizeof(float);
    L.total = o;
    return L;
}
int fill(SilParams& P, void* ws, size_t ws_bytes, int N, int V, int F, int H, int W, float blur_radius, float sigma, int K, int perspective, bool scratch) {
    FB_REQUIRE(N >= 0 && V >= 1 && F >= 0 && H >= 1 && W >= 1, FINDB200_ERR_SHAPE, "silhouette: bad shape N=%d V=%d F=%d H=%d W=%d", N, V, F, H, W);
    FB_REQUIRE(N <= 65535 && (long long)N * ((W + 15) / 16) * ((H + 15) / 16) < 2000000000ll, FINDB200_ERR_SHAPE, "silhouette: too many meshes x views x tiles in one call (N=%d)", N);
    FB_REQUIRE(blur_radius >= 0.f && sigma > 0.f && K >= 1, FINDB200_ERR_ARG, "silhouette: blur_radius >= 0, sigma > 0 and faces_per_pixel >= 1 required");
    const WsLayout L = ws_layout(N, F);
    const size_t need = scratch ? L.total : L.cbits;
    FB_REQUIRE(ws && ws_bytes >= need, FINDB200_ERR_WORKSPACE, "silhouette: workspace too small (%zu < %zu)", ws_bytes, need);
    memset(&P, 0, sizeof(P));
    P.box = (const float4*)((char*)ws + L.box);
    P.rec = (const FaceRec*)((char*)ws + L.rec);
    P.overflow = (unsigned int*)((char*)ws + L.flag);
    P.tile_counter = P.overflow + 1;
    P.cand_bits = (uint32_t*)((char*)ws + L.cbits);
    P.cand_fac = (float*)((char*)ws + L.cfac);
    P.cap = cand_cap();
    P.N = N; P.F = F; P.H = H; P.W = W; P.V = V;
    P.blur_radius = blur_radius; P.sigma = sigma; P.K = K; P.perspective = perspective; P.clip = 1;
    return 0;
}
}  // namespace

extern "C" size_t findb200_silhouette_workspace_bytes(int N, int F) { return ws_layout(N > 0 ? N : 0, F > 0 ? F : 0).total; }
// the leading part of the workspace the backward reads (face records + status flag): keep this prefix, the rest is scratch
extern "C" size_t findb200_silhouette_saved_bytes(int N, int F) { return ws_layout(N > 0 ? N : 0, F > 0 ? F : 0).cbits; }

extern "C" int findb200_silhouette_forward(const float* verts_ndc, const int32_t* faces, int N, int V, int F, int H, int W,
                                           float blur_radius, float sigma, int faces_per_pixel, int perspective_correct,
                                           float* alpha, float* prod, void* sel, int32_t* pix_to_face, float* zbuf, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    SilParams P;
    int rc = fill(P, ws, ws_bytes, N, V, F, H, W, blur_radius, sigma, faces_per_pixel, perspective_correct, true);
    if (rc) return rc;
    FB_REQUIRE(verts_ndc && (faces || F == 0) && alpha, FINDB200_ERR_ARG, "silhouette_forward: null argument");
    if (N == 0) return 0;
    P.alpha = alpha; P.prod = prod; P.sel = (uint2*)sel; P.pix_to_face = pix_to_face; P.zbuf = zbuf;
    FB_CUDA(cudaMemsetAsync(P.overflow, 0, 2 * sizeof(unsigned int), st));      // status flag + tile queue
    if (F > 0) {
        FB_MARK("face_setup", st);
        const long long tot = (long long)N * F;
        FB_CUDA(launch_pdl(face_setup_kernel, dim3((unsigned)((tot + 255) / 256)), dim3(256), 0, st, verts_ndc, faces, N, V, F,
                           sqrtf(blur_radius), (float4*)P.box, (FaceRec*)P.rec));
    }
    FB_CUDA((cudaError_t)prefer_shared());
    FB_MARK("sil_tile_fwd", st);
    const long long n_tiles = (long long)N * ((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
    FB_CUDA(launch_pdl(sil_tile_fwd_kernel<false>, dim3((unsigned)(n_tiles < kFwdCtas ? n_tiles : kFwdCtas)), dim3(kThreads), 0, st, P));
    FB_MARK_END(st);
    return 0;
}

// Hard rasterisation (faces_per_pixel = 1): the nearest candidate face of every pixel with its depth, barycentrics and signed
// squared distance — PyTorch3D's Fragments of rasterize_meshes(faces_per_pixel=1), as the image rasteriser of
// renderer.py:116-120 produces them (blur_radius = 0: a candidate is a face that contains the pixel centre).
extern "C" int findb200_rasterize_nearest(const float* verts_ndc, const int32_t* faces, int N, int V, int F, int H, int W,
                                          float blur_radius, int perspective_correct, int clip_barycentric_coords,
                                          int32_t* pix_to_face, float* zbuf, float* bary, float* dists, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    SilParams P;
    int rc = fill(P, ws, ws_bytes, N, V, F, H, W, blur_radius, 1.f, 1, perspective_correct, true);
    if (rc) return rc;
    FB_REQUIRE(verts_ndc && (faces || F == 0) && pix_to_face, FINDB200_ERR_ARG, "rasterize_nearest: null argument");
    if (N == 0) return 0;
    P.pix_to_face = pix_to_face; P.zbuf = zbuf; P.bary = bary; P.dists = dists; P.clip = clip_barycentric_coords ? 1 : 0;
    FB_CUDA(cudaMemsetAsync(P.overflow, 0, 2 * sizeof(unsigned int), st));
    if (F > 0) {
        FB_MARK("face_setup", st);
        const long long tot = (long long)N * F;
        FB_CUDA(launch_pdl(face_setup_kernel, dim3((unsigned)((tot + 255) / 256)), dim3(256), 0, st, verts_ndc, faces, N, V, F,
                           sqrtf(blur_radius), (float4*)P.box, (FaceRec*)P.rec));
    }
    FB_CUDA((cudaError_t)prefer_shared());
    FB_MARK("raster_nearest", st);
    const long long n_tiles = (long long)N * ((W + kTile - 1) / kTile) * ((H + kTile - 1) / kTile);
    FB_CUDA(launch_pdl(sil_tile_fwd_kernel<true>, dim3((unsigned)(n_tiles < kFwdCtas ? n_tiles : kFwdCtas)), dim3(kThreads), 0, st, P));
    FB_MARK_END(st);
    return 0;
}

// The workspace must be the one the forward call filled (face records); grad_verts is accumulated into (+=) x and y, z gets none.
extern "C" int findb200_silhouette_backward(const float* grad_alpha, const int32_t* faces, int N, int V, int F, int H, int W,
                                            float blur_radius, float sigma, int faces_per_pixel, int perspective_correct,
                                            const float* prod, const void* sel, float* grad_verts_ndc, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    SilParams P;
    int rc = fill(P, ws, ws_bytes, N, V, F, H, W, blur_radius, sigma, faces_per_pixel, perspective_correct, false);
    if (rc) return rc;
    if (N == 0 || F == 0) return 0;
    FB_REQUIRE(grad_alpha && faces && prod && sel && grad_verts_ndc, FINDB200_ERR_ARG, "silhouette_backward: null argument");
    P.grad_alpha = grad_alpha; P.prod = (float*)prod; P.sel = (uint2*)sel; P.grad_verts = grad_verts_ndc; P.faces = faces;
    FB_CUDA((cudaError_t)prefer_shared());
    FB_MARK("sil_tile_bwd", st);
    FB_CUDA(launch_pdl(sil_tile_bwd_kernel, dim3((W + kTile - 1) / kTile, (H + kTile - 1) / kTile, N), dim3(kThreads), 0, st, P));
    FB_MARK_END(st);
    return 0;
}

// FeatureShader on the fragments of a findb200_silhouette_forward call (same workspace, its sel / zbuf / pix_to_face outputs).
// features [M, V, C] with M = N / num_views meshes; background [C] on the device or NULL (zeros); out [N, H, W, C], den [N, H, W].
static int fill_feat(FeatParams& Q, const float* features, int C, int num_views, const int32_t* faces, int N, int V, int F, int H, int W,
                     float blur_radius, float sigma, int K, int perspective, float gamma, float znear, float zfar, const float* background,
                     const void* sel, const float* zbuf, void* ws, size_t ws_bytes) {
    int rc = fill(Q.s, ws, ws_bytes, N, V, F, H, W, blur_radius, sigma, K, perspective, false);
    if (rc) return rc;
    FB_REQUIRE(features && faces && sel && zbuf && C >= 1 && num_views >= 1 && N % num_views == 0, FINDB200_ERR_ARG, "feature_blend: bad argument");
    FB_REQUIRE(gamma > 0.f && zfar > znear, FINDB200_ERR_ARG, "feature_blend: gamma > 0 and zfar > znear required");
    Q.s.sel = (uint2*)sel; Q.s.faces = faces;
    Q.zbuf = zbuf; Q.p2f = nullptr; Q.feat = features; Q.C = C; Q.views = num_views; Q.c0 = 0; Q.gamma = gamma; Q.znear = znear; Q.zfar = zfar;
    Q.bg = background; Q.out = nullptr; Q.den = nullptr; Q.grad_out = nullptr; Q.g_feat = nullptr; Q.g_ndc = nullptr;
    return 0;
}

extern "C" int findb200_feature_blend_forward(const float* features, int C, int num_views, const int32_t* faces, int N, int V, int F, int H,
                                              int W, float blur_radius, float sigma, int faces_per_pixel, int perspective_correct,
                                              float gamma, float znear, float zfar, const float* background, const void* sel,
                                              const float* zbuf, float* out, float* den, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0) return 0;
    FeatParams Q;
    int rc = fill_feat(Q, features, C, num_views, faces, N, V, F, H, W, blur_radius, sigma, faces_per_pixel, perspective_correct, gamma, znear,
                       zfar, background, sel, zbuf, ws, ws_bytes);
    if (rc) return rc;
    FB_REQUIRE(out && den, FINDB200_ERR_ARG, "feature_blend_forward: null output");
    Q.out = out; Q.den = den;
    FB_CUDA((cudaError_t)prefer_shared());
    FB_MARK("feat_tile_fwd", st);
    for (int c0 = 0; c0 < C; c0 += kFeatCh) {            // eight channels per walk (per-pixel accumulators live in registers)
        Q.c0 = c0;
        FB_CUDA(launch_pdl(feat_tile_fwd_kernel, dim3((W + kTile - 1) / kTile, (H + kTile - 1) / kTile, N), dim3(kThreads), 0, st, Q));
    }
    FB_MARK_END(st);
    return 0;
}

// grad_features [M, V, C] and grad_verts_ndc [N, V, 3] are ACCUMULATED into (the caller zeroes them)
extern "C" int findb200_feature_blend_backward(const float* grad_out, const float* features, int C, int num_views, const int32_t* faces,
                                               int N, int V, int F, int H, int W, float blur_radius, float sigma, int faces_per_pixel,
                                               int perspective_correct, float gamma, float znear, float zfar, const float* background,
                                               const void* sel, const float* zbuf, const int32_t* pix_to_face, const float* out,
                                               const float* den, float* grad_features, float* grad_verts_ndc, void* ws, size_t ws_bytes,
                                               void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (N == 0 || F == 0) return 0;
    FeatParams Q;
    int rc = fill_feat(Q, features, C, num_views, faces, N, V, F, H, W, blur_radius, sigma, faces_per_pixel, perspective_correct, gamma, znear,
                       zfar, background, sel, zbuf, ws, ws_bytes);
    if (rc) return rc;
    FB_REQUIRE(grad_out && pix_to_face && out && den && grad_features && grad_verts_ndc, FINDB200_ERR_ARG, "feature_blend_backward: null argument");
    Q.p2f = pix_to_face; Q.out = (float*)out; Q.den = (float*)den; Q.grad_out = grad_out; Q.g_feat = grad_features; Q.g_ndc = grad_verts_ndc;
    static bool carve = false;
    if (!carve) { cudaFuncSetAttribute(feat_tile_bwd_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared); carve = true; }
    FB_MARK("feat_tile_bwd", st);
    FB_CUDA(launch_pdl(feat_tile_bwd_kernel, dim3((W + kTile - 1) / kTile, (H + kTile - 1) / kTile, N), dim3(kThreads), 0, st, Q));
    FB_MARK_END(st);
    return 0;
}

// 1 when a tile of the last forward call on this workspace could not apply the K truncation exactly (more than 4096 faces on
// one 16 x 16 tile AND a pixel with more than faces_per_pixel candidates): alpha there multiplies ALL candidates.  Synchronises.
extern "C" int findb200_silhouette_status(const void* ws, int N, int F, void* stream) {
    unsigned int flag = 0;
    const WsLayout L = ws_layout(N, F);
    FB_CUDA(cudaMemcpyAsync(&flag, (const char*)ws + L.flag, sizeof(flag), cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    FB_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return (int)flag;
}

// ------------------------------------------------------------------------------------------------------------------------
// Phong image of the hard fragments (forward only): pytorch3d SoftPhongShader = TexturesVertex sampling + phong_shading
// (PointLights, default Materials) + softmax_rgb_blend, as FootRenderer sets it up (renderer.py:110-115, 133) and calls it
// (:273-275).  Restated in oracle/raster_oracle.py::phong_image.
namespace fb {
namespace raster {

// area-weighted vertex normals, PyTorch3D's per-corner cross products, gathered over the vertex -> (face, corner) CSR list
__global__ void vert_normals_kernel(const float* __restrict__ verts, const int* __restrict__ faces, const int* __restrict__ vf_ptr,
                                    const int* __restrict__ vf_face, const int* __restrict__ vf_corner, int N, int V, float* __restrict__ out) {
    pdl_wait_trigger();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)N * V) return;
    const int n = (int)(i / V), v = (int)(i - (long long)n * V);
    const float* vb = verts + (size_t)n * V * 3;
    float nx = 0.f, ny = 0.f, nz = 0.f;
    for (int k = vf_ptr[v]; k < vf_ptr[v + 1]; ++k) {
        const int f = vf_face[k], c = vf_corner[k];
        const int ia = faces[3 * f + c], ib = faces[3 * f + (c + 1) % 3], ic = faces[3 * f + (c + 2) % 3];
        const float ax = vb[3 * ia], ay = vb[3 * ia + 1], az = vb[3 * ia + 2];
        const float ux = vb[3 * ib] - ax, uy = vb[3 * ib + 1] - ay, uz = vb[3 * ib + 2] - az;     // next corner - this corner
        const float wx = vb[3 * ic] - ax, wy = vb[3 * ic + 1] - ay, wz = vb[3 * ic + 2] - az;     // previous corner - this corner
        nx += uy * wz - uz * wy; ny += uz * wx - ux * wz; nz += ux * wy - uy * wx;
    }
    const float len = fmaxf(sqrtf(nx * nx + ny * ny + nz * nz), 1e-6f);
    out[3 * i] = nx / len; out[3 * i + 1] = ny / len; out[3 * i + 2] = nz / len;
}

struct PhongParams {
    const int* p2f; const float* bary; const float* zbuf; const float* dists;      // [I, H, W (,3)]
    const float* verts; const float* normals; const float* colours;               // [M, V, 3]
    const int* faces;
    const float* cam;                                                              // [views, 3] camera centres (world)
    long long n_pix; int HW, V, views, H_;
    float light[3], amb[3], dif[3], spec[3], bg[3];
    float shininess, sigma, gamma, znear, zfar;
    float* out;                                                                    // [I, H, W, 3]
};

FB_DEVINL void normalize3(float& x, float& y, float& z) {
    const float l = fmaxf(sqrtf(x * x + y * y + z * z), 1e-6f);
    x /= l; y /= l; z /= l;
}

__global__ void phong_shade_kernel(const __grid_constant__ PhongParams P) {
    pdl_wait_trigger();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_pix) return;
    const int f = P.p2f[i];
    float r = P.bg[0], g = P.bg[1], b = P.bg[2];
    if (f >= 0) {
        const int img = (int)(i / P.HW), m = img / P.views, view = img - m * P.views;
        const float w0 = P.bary[3 * i], w1 = P.bary[3 * i + 1], w2 = P.bary[3 * i + 2];
        const int i0 = P.faces[3 * f], i1 = P.faces[3 * f + 1], i2 = P.faces[3 * f + 2];
        const size_t o = (size_t)m * P.V * 3;
        float p[3], nrm[3], tex[3];
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            p[d] = w0 * P.verts[o + 3 * i0 + d] + w1 * P.verts[o + 3 * i1 + d] + w2 * P.verts[o + 3 * i2 + d];
            nrm[d] = w0 * P.normals[o + 3 * i0 + d] + w1 * P.normals[o + 3 * i1 + d] + w2 * P.normals[o + 3 * i2 + d];
            tex[d] = w0 * P.colours[o + 3 * i0 + d] + w1 * P.colours[o + 3 * i1 + d] + w2 * P.colours[o + 3 * i2 + d];
        }
        normalize3(nrm[0], nrm[1], nrm[2]);
        float lx = P.light[0] - p[0], ly = P.light[1] - p[1], lz = P.light[2] - p[2];
        normalize3(lx, ly, lz);
        const float cosang = nrm[0] * lx + nrm[1] * ly + nrm[2] * lz;
        const float diff = fmaxf(cosang, 0.f);
        float vx = P.cam[3 * view] - p[0], vy = P.cam[3 * view + 1] - p[1], vz = P.cam[3 * view + 2] - p[2];
        normalize3(vx, vy, vz);
        const float rx = -lx + 2.f * cosang * nrm[0], ry = -ly + 2.f * cosang * nrm[1], rz = -lz + 2.f * cosang * nrm[2];
        const float al = (cosang > 0.f) ? fmaxf(vx * rx + vy * ry + vz * rz, 0.f) : 0.f;
        const float sp = powf(al, P.shininess);
        const float eps = 1e-10f;
        const float prob = __fdiv_rn(1.f, 1.f + expf(P.dists[i] / P.sigma));          // sigmoid(-dist / sigma)
        const float z_inv = (P.zfar - P.zbuf[i]) / (P.zfar - P.znear);
        const float z_max = fmaxf(z_inv, eps);
        const float wnum = prob * expf((z_inv - z_max) / P.gamma);
        const float delta = fmaxf(expf((eps - z_max) / P.gamma), eps);
        const float den = wnum + delta;
        const float c0 = (P.amb[0] + P.dif[0] * diff) * tex[0] + P.spec[0] * sp;
        const float c1 = (P.amb[1] + P.dif[1] * diff) * tex[1] + P.spec[1] * sp;
        const float c2 = (P.amb[2] + P.dif[2] * diff) * tex[2] + P.spec[2] * sp;
        r = (wnum * c0 + delta * P.bg[0]) / den; g = (wnum * c1 + delta * P.bg[1]) / den; b = (wnum * c2 + delta * P.bg[2]) / den;
    }
    P.out[3 * i] = r; P.out[3 * i + 1] = g; P.out[3 * i + 2] = b;
}

// ---- backward of the Phong image ----------------------------------------------------------------------------------------
// d(image)/d(world vertices, vertex colours): per pixel, the forward is recomputed and differentiated by hand —
//   blend (softmax_rgb_blend, one face per pixel)  ->  colour, and the sigmoid of the signed edge distance,
//   phong_shading                                   ->  interpolated position / normal / texel,
//   barycentric interpolation                       ->  the three vertices' attributes, and the barycentrics themselves,
//   rasterize_meshes backward                       ->  the face's NDC vertices: perspective correction, edge functions, and
//                                                       the squared distance to the nearest edge.
// pix_to_face and the depth ordering carry no gradient; neither does zbuf here: with gamma = 1e-4 the background weight
// delta = exp((eps - z_inv) / gamma) is clamped at eps for every visible depth, and (z_inv - max z_inv) is identically 0 with
// one face per pixel.  Same rules as oracle/raster_oracle.py::render_image under autograd.  Gradients are accumulated with fp32
// atomics into pre-zeroed buffers (PyTorch3D does the same).
struct PhongBwdParams {
    PhongParams f;
    const float* grad_image;              // [I, H, W, 3]
    const float* verts_ndc;               // [I, V, 3]
    float *g_verts, *g_normals, *g_colours;   // [M, V, 3]
    float* g_ndc;                         // [I, V, 3]
};

FB_DEVINL void normalize_bwd(const float (&x)[3], const float (&g)[3], float (&out)[3]) {
    // y = x / max(|x|, 1e-6):  dL/dx = (g - y (y . g)) / |x|   (|x| > eps), g / eps otherwise
    const float l = sqrtf(x[0] * x[0] + x[1] * x[1] + x[2] * x[2]);
    if (l > 1e-6f) {
        const float y0 = x[0] / l, y1 = x[1] / l, y2 = x[2] / l;
        const float d = y0 * g[0] + y1 * g[1] + y2 * g[2];
        out[0] = (g[0] - y0 * d) / l; out[1] = (g[1] - y1 * d) / l; out[2] = (g[2] - y2 * d) / l;
    } else {
        out[0] = g[0] / 1e-6f; out[1] = g[1] / 1e-6f; out[2] = g[2] / 1e-6f;
    }
}

__global__ void __launch_bounds__(128) phong_shade_bwd_kernel(const __grid_constant__ PhongBwdParams B) {
    pdl_wait_trigger();
    const PhongParams& P = B.f;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_pix) return;
    const int f = P.p2f[i];
    if (f < 0) return;
    const float g[3] = {B.grad_image[3 * i], B.grad_image[3 * i + 1], B.grad_image[3 * i + 2]};
    if (g[0] == 0.f && g[1] == 0.f && g[2] == 0.f) return;
    const int img = (int)(i / P.HW), m = img / P.views, view = img - m * P.views;
    const int pix = (int)(i - (long long)img * P.HW);
    const float w[3] = {P.bary[3 * i], P.bary[3 * i + 1], P.bary[3 * i + 2]};
    const int vid[3] = {P.faces[3 * f], P.faces[3 * f + 1], P.faces[3 * f + 2]};
    const size_t o = (size_t)m * P.V * 3;
    float Pv[3][3], Nv[3][3], Cv[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            Pv[k][d] = P.verts[o + 3 * vid[k] + d]; Nv[k][d] = P.normals[o + 3 * vid[k] + d]; Cv[k][d] = P.colours[o + 3 * vid[k] + d];
        }
    float p[3], nraw[3], tex[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        p[d] = w[0] * Pv[0][d] + w[1] * Pv[1][d] + w[2] * Pv[2][d];
        nraw[d] = w[0] * Nv[0][d] + w[1] * Nv[1][d] + w[2] * Nv[2][d];
        tex[d] = w[0] * Cv[0][d] + w[1] * Cv[1][d] + w[2] * Cv[2][d];
    }
    float n[3] = {nraw[0], nraw[1], nraw[2]};
    normalize3(n[0], n[1], n[2]);
    const float lraw[3] = {P.light[0] - p[0], P.light[1] - p[1], P.light[2] - p[2]};
    float l[3] = {lraw[0], lraw[1], lraw[2]};
    normalize3(l[0], l[1], l[2]);
    const float cosang = n[0] * l[0] + n[1] * l[1] + n[2] * l[2];
    const float diff = fmaxf(cosang, 0.f);
    const float vraw[3] = {P.cam[3 * view] - p[0], P.cam[3 * view + 1] - p[1], P.cam[3 * view + 2] - p[2]};
    float v[3] = {vraw[0], vraw[1], vraw[2]};
    normalize3(v[0], v[1], v[2]);
    const float r[3] = {-l[0] + 2.f * cosang * n[0], -l[1] + 2.f * cosang * n[1], -l[2] + 2.f * cosang * n[2]};
    const float vr = v[0] * r[0] + v[1] * r[1] + v[2] * r[2];
    const float al = (cosang > 0.f) ? fmaxf(vr, 0.f) : 0.f;
    const float sp = powf(al, P.shininess);
    const float eps = 1e-10f;
    const float dist = P.dists[i];
    const float prob = __fdiv_rn(1.f, 1.f + expf(dist / P.sigma));
    const float z_inv = (P.zfar - P.zbuf[i]) / (P.zfar - P.znear);
    const float z_max = fmaxf(z_inv, eps);
    const float ez = expf((z_inv - z_max) / P.gamma);
    const float wnum = prob * ez;
    const float delta = fmaxf(expf((eps - z_max) / P.gamma), eps);
    const float den = wnum + delta;
    float c[3], outc[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        c[d] = (P.amb[d] + P.dif[d] * diff) * tex[d] + P.spec[d] * sp;
        outc[d] = (wnum * c[d] + delta * P.bg[d]) / den;
    }
    // ---- blend
    float gc[3], g_wnum = 0.f;
#pragma unroll
    for (int d = 0; d < 3; ++d) { gc[d] = g[d] * wnum / den; g_wnum += g[d] * (c[d] - outc[d]) / den; }
    const float g_prob = g_wnum * ez;
    const float g_dist = g_prob * (-prob * (1.f - prob) / P.sigma);
    // ---- shading
    float g_tex[3], g_diff = 0.f, g_sp = 0.f;
#pragma unroll
    for (int d = 0; d < 3; ++d) { g_tex[d] = gc[d] * (P.amb[d] + P.dif[d] * diff); g_diff += gc[d] * P.dif[d] * tex[d]; g_sp += gc[d] * P.spec[d]; }
    float g_cos = (cosang > 0.f) ? g_diff : 0.f;
    float g_vr = 0.f;
    if (cosang > 0.f && vr > 0.f && al > 0.f) g_vr = g_sp * P.shininess * powf(al, P.shininess - 1.f);
    float g_v[3], g_r[3], g_l[3], g_n[3];
#pragma unroll
    for (int d = 0; d < 3; ++d) { g_v[d] = g_vr * r[d]; g_r[d] = g_vr * v[d]; }
    g_cos += 2.f * (g_r[0] * n[0] + g_r[1] * n[1] + g_r[2] * n[2]);
#pragma unroll
    for (int d = 0; d < 3; ++d) { g_l[d] = -g_r[d] + g_cos * n[d]; g_n[d] = 2.f * cosang * g_r[d] + g_cos * l[d]; }
    float g_vraw[3], g_lraw[3], g_nraw[3], g_p[3];
    normalize_bwd(vraw, g_v, g_vraw);
    normalize_bwd(lraw, g_l, g_lraw);
    normalize_bwd(nraw, g_n, g_nraw);
#pragma unroll
    for (int d = 0; d < 3; ++d) g_p[d] = -g_vraw[d] - g_lraw[d];
    // ---- interpolation: attribute gradients to the three vertices, and d/d(barycentrics)
    float g_w[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        g_w[k] = 0.f;
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            g_w[k] += g_p[d] * Pv[k][d] + g_nraw[d] * Nv[k][d] + g_tex[d] * Cv[k][d];
            atomicAdd(B.g_verts + o + 3 * vid[k] + d, w[k] * g_p[d]);
            atomicAdd(B.g_normals + o + 3 * vid[k] + d, w[k] * g_nraw[d]);
            atomicAdd(B.g_colours + o + 3 * vid[k] + d, w[k] * g_tex[d]);
        }
    }
    // ---- rasteriser backward on the face's NDC vertices
    const float* nd = B.verts_ndc + (size_t)img * P.V * 3;
    float X[3], Y[3], Z[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) { X[k] = nd[3 * vid[k]]; Y[k] = nd[3 * vid[k] + 1]; Z[k] = nd[3 * vid[k] + 2]; }
    const int Wd = P.HW / P.H_;
    const int row = pix / Wd, col = pix - row * Wd;
    const float px = 1.f - __fdiv_rn((float)(2 * col + 1), (float)Wd), py = 1.f - __fdiv_rn((float)(2 * row + 1), (float)P.H_);
    float gX[3] = {0.f, 0.f, 0.f}, gY[3] = {0.f, 0.f, 0.f}, gZ[3] = {0.f, 0.f, 0.f};
    // perspective correction: b_k = t_k / dn, t0 = w0 z1 z2, t1 = z0 w1 z2, t2 = z0 z1 w2 (w = screen-space barycentrics)
    const float area = edge_fn(X[2], Y[2], X[0], Y[0], X[1], Y[1]);
    const float dena = area + kEps;
    const float e0 = edge_fn(px, py, X[1], Y[1], X[2], Y[2]), e1 = edge_fn(px, py, X[2], Y[2], X[0], Y[0]), e2 = edge_fn(px, py, X[0], Y[0], X[1], Y[1]);
    const float ws[3] = {e0 / dena, e1 / dena, e2 / dena};
    const float t0 = ws[0] * Z[1] * Z[2], t1 = Z[0] * ws[1] * Z[2], t2 = Z[0] * Z[1] * ws[2];
    const float dn_raw = t0 + t1 + t2;
    float g_ws[3];
    if (dn_raw > kEps) {
        const float dn = dn_raw;
        const float bsum = g_w[0] * (t0 / dn) + g_w[1] * (t1 / dn) + g_w[2] * (t2 / dn);
        const float g_t0 = (g_w[0] - bsum) / dn, g_t1 = (g_w[1] - bsum) / dn, g_t2 = (g_w[2] - bsum) / dn;
        g_ws[0] = g_t0 * Z[1] * Z[2]; g_ws[1] = g_t1 * Z[0] * Z[2]; g_ws[2] = g_t2 * Z[0] * Z[1];
        gZ[0] += g_t1 * ws[1] * Z[2] + g_t2 * Z[1] * ws[2];
        gZ[1] += g_t0 * ws[0] * Z[2] + g_t2 * Z[0] * ws[2];
        gZ[2] += g_t0 * ws[0] * Z[1] + g_t1 * Z[0] * ws[1];
    } else {
        const float dn = kEps;                     // clamped denominator: a constant
        g_ws[0] = g_w[0] / dn * Z[1] * Z[2]; g_ws[1] = g_w[1] / dn * Z[0] * Z[2]; g_ws[2] = g_w[2] / dn * Z[0] * Z[1];
        gZ[0] += g_w[1] / dn * ws[1] * Z[2] + g_w[2] / dn * Z[1] * ws[2];
        gZ[1] += g_w[0] / dn * ws[0] * Z[2] + g_w[2] / dn * Z[0] * ws[2];
        gZ[2] += g_w[0] / dn * ws[0] * Z[1] + g_w[1] / dn * Z[0] * ws[1];
    }
    // screen-space barycentrics: ws_k = e_k / (area + eps);  edge(p; a, b) = (px - ax)(by - ay) - (py - ay)(bx - ax)
    const float g_e[3] = {g_ws[0] / dena, g_ws[1] / dena, g_ws[2] / dena};
    const float g_area = -(g_ws[0] * ws[0] + g_ws[1] * ws[1] + g_ws[2] * ws[2]) / dena;
    auto edge_bwd = [&](float ge, int a, int b) {       // p = the pixel (no gradient)
        gX[a] += ge * (py - Y[b]); gY[a] += ge * (X[b] - px);
        gX[b] += ge * -(py - Y[a]); gY[b] += ge * (px - X[a]);
    };
    edge_bwd(g_e[0], 1, 2);
    edge_bwd(g_e[1], 2, 0);
    edge_bwd(g_e[2], 0, 1);
    // area = edge(v2; v0, v1): also a function of its "point" v2
    gX[0] += g_area * (Y[2] - Y[1]); gY[0] += g_area * (X[1] - X[2]);
    gX[1] += g_area * -(Y[2] - Y[0]); gY[1] += g_area * (X[2] - X[0]);
    gX[2] += g_area * (Y[1] - Y[0]); gY[2] += g_area * -(X[1] - X[0]);
    // signed squared distance to the nearest edge (negative inside): d2 = |p - q|^2, q = a + t (b - a), t clamped
    if (g_dist != 0.f) {
        float tt[3];
        const float d01 = seg_dist2(px, py, X[0], Y[0], X[1], Y[1], tt[0]);
        const float d02 = seg_dist2(px, py, X[0], Y[0], X[2], Y[2], tt[1]);
        const float d12 = seg_dist2(px, py, X[1], Y[1], X[2], Y[2], tt[2]);
        int ed = 0;
        float dmin = d01;
        if (d02 < dmin) { dmin = d02; ed = 1; }
        if (d12 < dmin) { dmin = d12; ed = 2; }
        const int ia = ed == 2 ? 1 : 0, ib = ed == 0 ? 1 : 2;
        const bool inside = (ws[0] > 0.f) && (ws[1] > 0.f) && (ws[2] > 0.f);
        const float gs = inside ? -g_dist : g_dist;                  // d(signed)/d(d2)
        const float t = tt[ed];
        if (t == 2.f) {                                              // degenerate segment: distance to b
            gX[ib] += gs * -2.f * (px - X[ib]); gY[ib] += gs * -2.f * (py - Y[ib]);
        } else {
            const float qx = X[ia] + t * (X[ib] - X[ia]), qy = Y[ia] + t * (Y[ib] - Y[ia]);
            const float cx = px - qx, cy = py - qy;
            gX[ia] += gs * -2.f * (1.f - t) * cx; gY[ia] += gs * -2.f * (1.f - t) * cy;
            gX[ib] += gs * -2.f * t * cx; gY[ib] += gs * -2.f * t * cy;
        }
    }
    float* gn = B.g_ndc + (size_t)img * P.V * 3;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        atomicAdd(gn + 3 * vid[k], gX[k]); atomicAdd(gn + 3 * vid[k] + 1, gY[k]); atomicAdd(gn + 3 * vid[k] + 2, gZ[k]);
    }
}

// backward of vert_normals_kernel.  Pass 1 (per vertex): gradient w.r.t. the un-normalised normal.  Pass 2 (per vertex, gather over
// its incident faces): every face adds three cross products T_c = (P_next - P_c) x (P_prev - P_c), one per corner c, to the raw
// normal of that corner's vertex; with G = d/d(raw normal of corner c): d/dP_next = w x G, d/dP_prev = G x u, d/dP_c = -(both).
__global__ void vert_normals_bwd_raw_kernel(const float* __restrict__ verts, const int* __restrict__ faces, const int* __restrict__ vf_ptr,
                                            const int* __restrict__ vf_face, const int* __restrict__ vf_corner, int N, int V,
                                            const float* __restrict__ g_normals, float* __restrict__ g_raw) {
    pdl_wait_trigger();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)N * V) return;
    const int n = (int)(i / V), v = (int)(i - (long long)n * V);
    const float* vb = verts + (size_t)n * V * 3;
    float x[3] = {0.f, 0.f, 0.f};
    for (int k = vf_ptr[v]; k < vf_ptr[v + 1]; ++k) {
        const int f = vf_face[k], c = vf_corner[k];
        const int ia = faces[3 * f + c], ib = faces[3 * f + (c + 1) % 3], ic = faces[3 * f + (c + 2) % 3];
        const float ax = vb[3 * ia], ay = vb[3 * ia + 1], az = vb[3 * ia + 2];
        const float ux = vb[3 * ib] - ax, uy = vb[3 * ib + 1] - ay, uz = vb[3 * ib + 2] - az;
        const float wx = vb[3 * ic] - ax, wy = vb[3 * ic + 1] - ay, wz = vb[3 * ic + 2] - az;
        x[0] += uy * wz - uz * wy; x[1] += uz * wx - ux * wz; x[2] += ux * wy - uy * wx;
    }
    const float g[3] = {g_normals[3 * i], g_normals[3 * i + 1], g_normals[3 * i + 2]};
    float out[3];
    normalize_bwd(x, g, out);
    g_raw[3 * i] = out[0]; g_raw[3 * i + 1] = out[1]; g_raw[3 * i + 2] = out[2];
}

__global__ void vert_normals_bwd_gather_kernel(const float* __restrict__ verts, const int* __restrict__ faces, const int* __restrict__ vf_ptr,
                                               const int* __restrict__ vf_face, const int* __restrict__ vf_corner, int N, int V,
                                               const float* __restrict__ g_raw, float* __restrict__ g_verts) {
    pdl_wait_trigger();
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)N * V) return;
    const int n = (int)(i / V), v = (int)(i - (long long)n * V);
    const float* vb = verts + (size_t)n * V * 3;
    const float* gr = g_raw + (size_t)n * V * 3;
    float acc[3] = {0.f, 0.f, 0.f};
    for (int k = vf_ptr[v]; k < vf_ptr[v + 1]; ++k) {
        const int f = vf_face[k], c = vf_corner[k];               // this vertex is corner c of face f
        int id[3];
        float Pc[3][3];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            id[q] = faces[3 * f + q];
            Pc[q][0] = vb[3 * id[q]]; Pc[q][1] = vb[3 * id[q] + 1]; Pc[q][2] = vb[3 * id[q] + 2];
        }
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) {                           // the term that belongs to corner cc
            const int nx = (cc + 1) % 3, pv = (cc + 2) % 3;
            const float G[3] = {gr[3 * id[cc]], gr[3 * id[cc] + 1], gr[3 * id[cc] + 2]};
            const float u[3] = {Pc[nx][0] - Pc[cc][0], Pc[nx][1] - Pc[cc][1], Pc[nx][2] - Pc[cc][2]};
            const float w[3] = {Pc[pv][0] - Pc[cc][0], Pc[pv][1] - Pc[cc][1], Pc[pv][2] - Pc[cc][2]};
            const float gu[3] = {w[1] * G[2] - w[2] * G[1], w[2] * G[0] - w[0] * G[2], w[0] * G[1] - w[1] * G[0]};     // w x G
            const float gw[3] = {G[1] * u[2] - G[2] * u[1], G[2] * u[0] - G[0] * u[2], G[0] * u[1] - G[1] * u[0]};     // G x u
            if (c == nx) { acc[0] += gu[0]; acc[1] += gu[1]; acc[2] += gu[2]; }
            else if (c == pv) { acc[0] += gw[0]; acc[1] += gw[1]; acc[2] += gw[2]; }
            else { acc[0] -= gu[0] + gw[0]; acc[1] -= gu[1] + gw[1]; acc[2] -= gu[2] + gw[2]; }
        }
    }
    g_verts[3 * i] += acc[0]; g_verts[3 * i + 1] += acc[1]; g_verts[3 * i + 2] += acc[2];
}

}  // namespace raster
}  // namespace fb

extern "C" int findb200_vertex_normals(const float* verts, int N, int V, const int32_t* faces, const int32_t* vf_ptr,
                                       const int32_t* vf_face, const int32_t* vf_corner, float* normals, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(N >= 0 && V >= 1, FINDB200_ERR_SHAPE, "vertex_normals: bad shape N=%d V=%d", N, V);
    FB_REQUIRE(verts && faces && vf_ptr && vf_face && vf_corner && normals, FINDB200_ERR_ARG, "vertex_normals: null argument");
    if (N == 0) return 0;
    FB_MARK("vert_normals", st);
    const long long tot = (long long)N * V;
    FB_CUDA(launch_pdl(fb::raster::vert_normals_kernel, dim3((unsigned)((tot + 255) / 256)), dim3(256), 0, st, verts, faces, vf_ptr, vf_face,
                       vf_corner, N, V, normals));
    FB_MARK_END(st);
    return 0;
}

// shading[20] (host): light xyz, ambient rgb, diffuse rgb, specular rgb, background rgb, shininess, sigma, gamma, znear, zfar
extern "C" int findb200_phong_shade(const int32_t* pix_to_face, const float* bary, const float* zbuf, const float* dists,
                                    int n_images, int H, int W, int num_views, const float* verts_world, const float* vert_normals,
                                    const float* vert_colours, const int32_t* faces, int V, const float* cam_centres,
                                    const float* shading, float* image, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(n_images >= 0 && H >= 1 && W >= 1 && num_views >= 1 && V >= 1 && n_images % num_views == 0, FINDB200_ERR_SHAPE,
               "phong_shade: bad shape images=%d views=%d", n_images, num_views);
    FB_REQUIRE(pix_to_face && bary && zbuf && dists && verts_world && vert_normals && vert_colours && faces && cam_centres && shading && image,
               FINDB200_ERR_ARG, "phong_shade: null argument");
    if (n_images == 0) return 0;
    fb::raster::PhongParams P;
    P.p2f = pix_to_face; P.bary = bary; P.zbuf = zbuf; P.dists = dists; P.verts = verts_world; P.normals = vert_normals;
    P.colours = vert_colours; P.faces = faces; P.cam = cam_centres; P.HW = H * W; P.H_ = H; P.n_pix = (long long)n_images * H * W;
    P.V = V; P.views = num_views; P.out = image;
    for (int d = 0; d < 3; ++d) { P.light[d] = shading[d]; P.amb[d] = shading[3 + d]; P.dif[d] = shading[6 + d]; P.spec[d] = shading[9 + d]; P.bg[d] = shading[12 + d]; }
    P.shininess = shading[15]; P.sigma = shading[16]; P.gamma = shading[17]; P.znear = shading[18]; P.zfar = shading[19];
    FB_REQUIRE(P.sigma > 0.f && P.gamma > 0.f && P.zfar > P.znear, FINDB200_ERR_ARG, "phong_shade: sigma, gamma > 0 and zfar > znear required");
    FB_MARK("phong_shade", st);
    FB_CUDA(launch_pdl(fb::raster::phong_shade_kernel, dim3((unsigned)((P.n_pix + 255) / 256)), dim3(256), 0, st, P));
    FB_MARK_END(st);
    return 0;
}

// grad buffers are ACCUMULATED into (the caller zeroes them): grad_verts_world / grad_normals / grad_colours [M, V, 3],
// grad_verts_ndc [n_images, V, 3]
extern "C" int findb200_phong_shade_backward(const float* grad_image, const int32_t* pix_to_face, const float* bary, const float* zbuf,
                                             const float* dists, int n_images, int H, int W, int num_views, const float* verts_ndc,
                                             const float* verts_world, const float* vert_normals, const float* vert_colours,
                                             const int32_t* faces, int V, const float* cam_centres, const float* shading,
                                             float* grad_verts_world, float* grad_normals, float* grad_colours, float* grad_verts_ndc,
                                             void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(n_images >= 0 && H >= 1 && W >= 1 && num_views >= 1 && V >= 1 && n_images % num_views == 0, FINDB200_ERR_SHAPE,
               "phong_shade_backward: bad shape images=%d views=%d", n_images, num_views);
    FB_REQUIRE(grad_image && pix_to_face && bary && zbuf && dists && verts_ndc && verts_world && vert_normals && vert_colours && faces &&
               cam_centres && shading && grad_verts_world && grad_normals && grad_colours && grad_verts_ndc, FINDB200_ERR_ARG,
               "phong_shade_backward: null argument");
    if (n_images == 0) return 0;
    fb::raster::PhongBwdParams B;
    fb::raster::PhongParams& P = B.f;
    P.p2f = pix_to_face; P.bary = bary; P.zbuf = zbuf; P.dists = dists; P.verts = verts_world; P.normals = vert_normals;
    P.colours = vert_colours; P.faces = faces; P.cam = cam_centres; P.HW = H * W; P.H_ = H; P.n_pix = (long long)n_images * H * W;
    P.V = V; P.views = num_views; P.out = nullptr;
    for (int d = 0; d < 3; ++d) { P.light[d] = shading[d]; P.amb[d] = shading[3 + d]; P.dif[d] = shading[6 + d]; P.spec[d] = shading[9 + d]; P.bg[d] = shading[12 + d]; }
    P.shininess = shading[15]; P.sigma = shading[16]; P.gamma = shading[17]; P.znear = shading[18]; P.zfar = shading[19];
    FB_REQUIRE(P.sigma > 0.f && P.gamma > 0.f && P.zfar > P.znear, FINDB200_ERR_ARG, "phong_shade_backward: sigma, gamma > 0 and zfar > znear required");
    B.grad_image = grad_image; B.verts_ndc = verts_ndc; B.g_verts = grad_verts_world; B.g_normals = grad_normals; B.g_colours = grad_colours;
    B.g_ndc = grad_verts_ndc;
    FB_MARK("phong_shade_bwd", st);
    FB_CUDA(launch_pdl(fb::raster::phong_shade_bwd_kernel, dim3((unsigned)((P.n_pix + 127) / 128)), dim3(128), 0, st, B));
    FB_MARK_END(st);
    return 0;
}

// grad_verts [N, V, 3] += d/d(verts) through Meshes.verts_normals_padded(), given grad_normals [N, V, 3]; scratch: [N, V, 3] floats
extern "C" int findb200_vertex_normals_backward(const float* verts, int N, int V, const int32_t* faces, const int32_t* vf_ptr,
                                                const int32_t* vf_face, const int32_t* vf_corner, const float* grad_normals, float* scratch,
                                                float* grad_verts, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(N >= 0 && V >= 1, FINDB200_ERR_SHAPE, "vertex_normals_backward: bad shape N=%d V=%d", N, V);
    FB_REQUIRE(verts && faces && vf_ptr && vf_face && vf_corner && grad_normals && scratch && grad_verts, FINDB200_ERR_ARG,
               "vertex_normals_backward: null argument");
    if (N == 0) return 0;
    const long long tot = (long long)N * V;
    FB_MARK("vert_normals_bwd", st);
    FB_CUDA(launch_pdl(fb::raster::vert_normals_bwd_raw_kernel, dim3((unsigned)((tot + 255) / 256)), dim3(256), 0, st, verts, faces, vf_ptr,
                       vf_face, vf_corner, N, V, grad_normals, scratch));
    FB_CUDA(launch_pdl(fb::raster::vert_normals_bwd_gather_kernel, dim3((unsigned)((tot + 255) / 256)), dim3(256), 0, st, verts, faces, vf_ptr,
                       vf_face, vf_corner, N, V, (const float*)scratch, grad_verts));
    FB_MARK_END(st);
    return 0;
}

// ------------------------------------------------------------------------------------------------------------------------
// Keypoint overlay: PyTorch3D's PointsRenderer (PointsRasterizer radius r, K points per pixel + AlphaCompositor) as used by
// /root/reference/src/model/renderer.py:139-142, 365-373 (red discs at the keypoints, eval_3d.py:97).  Restated in
// oracle/raster_oracle.py::render_points.  One thread per pixel walks the cloud's points (staged through shared memory, a few
// dozen keypoints per cloud) and keeps the K candidates of smallest depth in a sorted register list (the earlier point on
// ties), then composites them front to back.  Forward only: the overlay is a visualisation of given keypoints.
namespace fb {
namespace raster {
constexpr int kPtsMaxK = 16, kPtsMaxC = 8, kPtsStage = 256;

__global__ void __launch_bounds__(256) points_render_kernel(const float* __restrict__ pts, const float* __restrict__ feats, int P, int C,
                                                            int H, int W, float r2, int K, float* __restrict__ image) {
    fb::pdl_wait_trigger();
    __shared__ float s_p[kPtsStage][3];
    const int n = blockIdx.y;
    const int pix = blockIdx.x * 256 + threadIdx.x;
    const bool live = pix < H * W;
    const int row = live ? pix / W : 0, col = live ? pix % W : 0;
    const float cx = 1.f - __fdiv_rn((float)(2 * col + 1), (float)W), cy = 1.f - __fdiv_rn((float)(2 * row + 1), (float)H);
    float kz[kPtsMaxK], kd[kPtsMaxK];
    int ki[kPtsMaxK];
    int cnt = 0;
    for (int p0 = 0; p0 < P; p0 += kPtsStage) {
        const int np = min(kPtsStage, P - p0);
        __syncthreads();
        for (int i = threadIdx.x; i < np * 3; i += 256) s_p[i / 3][i % 3] = pts[((size_t)n * P + p0) * 3 + i];
        __syncthreads();
        if (!live) continue;
        for (int j = 0; j < np; ++j) {
            const float z = s_p[j][2];
            if (!(z >= 0.f)) continue;
            const float dx = __fsub_rn(cx, s_p[j][0]), dy = __fsub_rn(cy, s_p[j][1]);
            const float d2 = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
            if (!(d2 < r2)) continue;
            if (cnt == K && !(z < kz[K - 1])) continue;          // full and not nearer than the farthest kept (ties: the earlier stays)
            int at = cnt < K ? cnt : K - 1;
#pragma unroll
            for (int s = kPtsMaxK - 1; s >= 1; --s)              // shift the entries behind the insertion point (z strictly larger)
                if (s <= at && kz[s - 1] > z) { kz[s] = kz[s - 1]; kd[s] = kd[s - 1]; ki[s] = ki[s - 1]; at = s - 1; }
            kz[at] = z; kd[at] = d2; ki[at] = p0 + j;
            if (cnt < K) ++cnt;
        }
    }
    if (!live) return;
    float out[kPtsMaxC];
#pragma unroll
    for (int c = 0; c < kPtsMaxC; ++c) out[c] = 0.f;
    float trans = 1.f;
#pragma unroll
    for (int s = 0; s < kPtsMaxK; ++s) {
        if (s >= cnt) break;
        const float w = __fsub_rn(1.f, __fdiv_rn(kd[s], r2));
        const float tw = __fmul_rn(trans, w);
        const float* f = feats + ((size_t)n * P + ki[s]) * C;
#pragma unroll
        for (int c = 0; c < kPtsMaxC; ++c)
            if (c < C) out[c] = __fadd_rn(out[c], __fmul_rn(tw, f[c]));
        trans = __fmul_rn(trans, __fsub_rn(1.f, w));
    }
    float* o = image + ((size_t)n * H * W + pix) * C;
#pragma unroll
    for (int c = 0; c < kPtsMaxC; ++c)
        if (c < C) o[c] = out[c];
}
}  // namespace raster
}  // namespace fb

// image [N, H, W, C] <- points_ndc [N, P, 3] (NDC x, y, view-space depth), features [N, P, C]; radius in NDC units
extern "C" int findb200_points_render(const float* points_ndc, const float* features, int N, int P, int C, int H, int W, float radius,
                                      int points_per_pixel, float* image, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(N >= 0 && P >= 0 && H >= 1 && W >= 1, FINDB200_ERR_SHAPE, "points_render: bad shape N=%d P=%d H=%d W=%d", N, P, H, W);
    FB_REQUIRE(C >= 1 && C <= fb::raster::kPtsMaxC, FINDB200_ERR_UNSUPPORTED, "points_render: 1..%d feature channels, got %d",
               fb::raster::kPtsMaxC, C);
    FB_REQUIRE(points_per_pixel >= 1 && points_per_pixel <= fb::raster::kPtsMaxK, FINDB200_ERR_UNSUPPORTED,
               "points_render: 1..%d points per pixel, got %d", fb::raster::kPtsMaxK, points_per_pixel);
    FB_REQUIRE(radius > 0.f, FINDB200_ERR_ARG, "points_render: radius must be positive");
    FB_REQUIRE(image && (P == 0 || (points_ndc && features)), FINDB200_ERR_ARG, "points_render: null argument");
    if (N == 0) return 0;
    const float r2 = (float)((double)radius * (double)radius);
    FB_MARK("points_render", st);
    FB_CUDA(launch_pdl(fb::raster::points_render_kernel, dim3((unsigned)((H * W + 255) / 256), (unsigned)N), dim3(256), 0, st, points_ndc,
                       features, P, C, H, W, r2, points_per_pixel, image));
    FB_CUDA(cudaGetLastError());
    FB_MARK_END(st);
    return 0;
}
