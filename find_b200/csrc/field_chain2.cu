// find_b200/csrc/field_chain2.cu — the chain kernels of the displacement field on CTA PAIRS (tcgen05 cta_group::2).
//
// Same four chains as field_chain.cu (kept as the single-CTA A/B build, FINDB200_CHAIN=1):
//   TRUNK_FWD : encode xyz (Fourier features) -> base.0 .. base.8 (+ReLU)                           model.py:421-426
//   HEAD_FWD  : trunk feature tile -> mlp_disp / mlp_col (latents folded into a per-foot bias)      model.py:428-451
//               -> tanh epilogue, registration transform                                           model.py:481-488
//   HEAD_BWD  : d(pre-tanh) -> dZ3 -> dZ2 -> dZ1 of one head (ReLU masks saved by the forward)
//   TRUNK_BWD : sum over feet/heads of dZ1 -> x W1h^T -> dZ of base.8 .. base.0
//
// Why pairs.  The single-CTA kernels wait for WEIGHT BYTES, not for the tensor pipe (profiles/r1_chain_trace.md): a [256 n x
// 64 k] fp16 weight chunk (32 KB) feeds 0.27 us of MMA for one 128-row tile, its L2 -> shared-memory round trip is ~1.4 us,
// and shared memory (2 x 64 KB activation tiles + ring) has room for three chunks in flight.  With cta_group::2 the two CTAs
// of a cluster compute ONE M = 256 MMA: each CTA keeps its own 128-row activation tile (A operand, its half of D in its own
// TMEM) and stages only HALF of every weight chunk (N/2 rows of the K-major image, 16 KB); the tensor core reads the B halves
// from both shared memories.  The same 96 KB ring now holds SIX chunk-halves = 1.6 us of MMA per SM, more than the round trip,
// and every weight byte leaves L2 once per 256 rows instead of once per 128.
//
// Roles per CTA (320 threads): warp 0 = producer (cp.async.bulk of this CTA's weight halves / input tiles), warp 1 = MMA issuer
// in the leader CTA (rank 0) and RELAY in the peer (forwards "my half has landed" to the leader's mbarriers: a plain bulk copy
// can only signal an mbarrier of its own CTA), warps 2..9 = epilogue: 8 warps share ONE tile at a time (two warps per TMEM lane
// quadrant, 128 accumulator columns each), so a tile's epilogue takes half as long as with 4 warps and the two tiles of a CTA
// alternate: while the warps convert tile t, the tensor core runs tile t^1.
//
// Synchronisation (every mbarrier exists at the same offset in both CTAs):
//   bar_full[s]      local      weight half of ring slot s has landed (tx bytes)
//   bar_peer_full[s] leader     the peer's half of slot s has landed (remote arrive by the peer's relay warp)
//   bar_empty[s]     both       MMAs reading slot s are complete (tcgen05.commit multicast to both CTAs)
//   bar_in_full[t]   local      bulk-loaded input tile t has landed;  bar_peer_in[t]: the leader's view of the peer's
//   bar_act_ready[t] leader     both CTAs' epilogues have written activation tile t (count 2, one elected arrive per CTA)
//   bar_acc_ready[t] both       accumulator t is complete (commit multicast)
//   bar_act_free[t]  local      activation buffer / accumulator t may be reused by the next item
#include <type_traits>

#include "field.cuh"
#include "field_chain_params.cuh"

namespace fb {
namespace field {

namespace c2 {

constexpr int kRing2 = 6;                 // ring slots of one chunk HALF each
constexpr int kSlotBytes = 16384;         // [128 n x 64 k] fp16
constexpr int kEpiWarps = 8;
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kThreads2 = 64 + kEpiThreads + 32;   // producer, scout / relay, 8 epilogue warps, issuer

FB_DEVINL void named_bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// diagnostics (findb200_chain_trace, tools/chain2_trace.py): %globaltimer stamps of cluster 0's first items into P.dbg[512]
FB_DEVINL void stamp(const ChainParams& P, bool on, int slot) {
    if (P.dbg && on && slot < 512) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        P.dbg[slot] = t;
    }
}

FB_DEVINL float grad_scale_from(uint32_t maxabs_bits) {      // same rule as field_chain.cu / field.cu::inv_grad_scale
    const float m = __uint_as_float(maxabs_bits);
    if (!(m > 0.f) || !(m < 3.0e38f)) return 1.0f;
    int e;
    frexpf(m, &e);
    return ldexpf(1.0f, 4 - e);
}

FB_DEVINL uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}
// (lo, hi) fp32 -> fp16x2 with ReLU folded into the conversion (F2FP.RELU): negative results (and -0) become +0
FB_DEVINL uint32_t cvt_relu_h2(float lo, float hi) {
    uint32_t r;
    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
    return r;
}
// per-half "is > 0" as 0xffff / 0x0000 (HSET2)
FB_DEVINL uint32_t gt0_mask_h2(uint32_t h2) {
    uint32_t r;
    asm("set.gt.u32.f16x2 %0, %1, %2;" : "=r"(r) : "r"(h2), "r"(0u));
    return r;
}
FB_DEVINL uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
    return r;
}

FB_DEVINL uint32_t img_row_off(int chunk32, int row) {   // byte offset of 16-B chunk `chunk32` (0..31) of `row` in a 4-image tile
    return (uint32_t)(chunk32 >> 3) * kImgBytes + (uint32_t)row * 128u + (uint32_t)(((chunk32 & 7) ^ (row & 7)) << 4);
}

// ReLU sign words.  One 32-bit word covers 32 consecutive columns = 16 packed fp16x2 words j = 0..15:
//   bit j      = column 2j   is > 0        bit 16 + j = column 2j + 1 is > 0
// so the forward packs a word with HSET2 + one LOP3 per fp16x2 word, and the backward expands it with PRMT's byte-sign
// replication: (bits << k) puts the two bits of word j = 7 - k (and of word 15 - k) on byte MSBs.

// 32 accumulator columns of one row: + bias (shared memory), ReLU, fp16, into the tile image; returns the sign word.
FB_DEVINL uint32_t relu_chunk(const uint32_t (&v)[32], int c32, uint8_t* act_tile, int row, const float* bias_s) {
    const float4* b4 = reinterpret_cast<const float4*>(bias_s + c32 * 32);
    uint32_t bits = 0;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const float4 b0 = b4[g * 2], b1 = b4[g * 2 + 1];
        float l0, h0, l1, h1, l2, h2, l3, h3;
        unpack2(add2(pack2(__uint_as_float(v[g * 8 + 0]), __uint_as_float(v[g * 8 + 1])), pack2(b0.x, b0.y)), l0, h0);
        unpack2(add2(pack2(__uint_as_float(v[g * 8 + 2]), __uint_as_float(v[g * 8 + 3])), pack2(b0.z, b0.w)), l1, h1);
        unpack2(add2(pack2(__uint_as_float(v[g * 8 + 4]), __uint_as_float(v[g * 8 + 5])), pack2(b1.x, b1.y)), l2, h2);
        unpack2(add2(pack2(__uint_as_float(v[g * 8 + 6]), __uint_as_float(v[g * 8 + 7])), pack2(b1.z, b1.w)), l3, h3);
        uint4 pk;
        pk.x = cvt_relu_h2(l0, h0); pk.y = cvt_relu_h2(l1, h1); pk.z = cvt_relu_h2(l2, h2); pk.w = cvt_relu_h2(l3, h3);
        *reinterpret_cast<uint4*>(act_tile + img_row_off(c32 * 4 + g, row)) = pk;
        bits |= gt0_mask_h2(pk.x) & (0x00010001u << (g * 4 + 0));
        bits |= gt0_mask_h2(pk.y) & (0x00010001u << (g * 4 + 1));
        bits |= gt0_mask_h2(pk.z) & (0x00010001u << (g * 4 + 2));
        bits |= gt0_mask_h2(pk.w) & (0x00010001u << (g * 4 + 3));
    }
    return bits;
}

// dZ = dA * [a > 0] for 32 columns: fp16x2 pack, then AND with the byte masks PRMT makes from the sign word
FB_DEVINL void mask_chunk(const uint32_t (&v)[32], int c32, uint8_t* act_tile, int row, uint32_t bits) {
    uint32_t pk[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) pk[j] = pack_h2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t s = bits << k;                  // bit p -> p + k: words 7 - k (bytes 0, 2) and 15 - k (bytes 1, 3)
        pk[7 - k] &= prmt(s, 0u, 0xAA88u);
        pk[15 - k] &= prmt(s, 0u, 0xBB99u);
    }
#pragma unroll
    for (int g = 0; g < 4; ++g)
        *reinterpret_cast<uint4*>(act_tile + img_row_off(c32 * 4 + g, row)) = make_uint4(pk[g * 4], pk[g * 4 + 1], pk[g * 4 + 2], pk[g * 4 + 3]);
}

// Fourier features of one row for encoded-input chunk c (64 internal columns), groups g0..g0+1 of the 4 (two threads share a
// row).  chunk c < 8: [sin(2 pi r_f) f = 32c..32c+31 | cos(2 pi r_f) same f],  r_f = x.B[:,f]; chunk 8: [x, y, z, 0...]
FB_DEVINL void encode_half_chunk(int c, int h, float x, float y, float z, const float* __restrict__ Bm, uint8_t* slot_img, int row) {
    if (c < 8) {
#pragma unroll 1
        for (int g = 2 * h; g < 2 * h + 2; ++g) {
            float s[8], co[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int f = c * 32 + g * 8 + j;
                const float r = fmaf(z, __ldg(Bm + 512 + f), fmaf(y, __ldg(Bm + 256 + f), x * __ldg(Bm + f)));
                const float fr = r - rintf(r);                    // exact; |fr| <= 0.5
                const float a = fr * 6.28318530717958647692f;
                s[j] = __sinf(a);
                co[j] = __cosf(a);
            }
            uint4 ps, pc;
            ps.x = pack_h2(s[0], s[1]); ps.y = pack_h2(s[2], s[3]); ps.z = pack_h2(s[4], s[5]); ps.w = pack_h2(s[6], s[7]);
            pc.x = pack_h2(co[0], co[1]); pc.y = pack_h2(co[2], co[3]); pc.z = pack_h2(co[4], co[5]); pc.w = pack_h2(co[6], co[7]);
            *reinterpret_cast<uint4*>(slot_img + (uint32_t)row * 128u + (uint32_t)((g ^ (row & 7)) << 4)) = ps;
            *reinterpret_cast<uint4*>(slot_img + (uint32_t)row * 128u + (uint32_t)(((g + 4) ^ (row & 7)) << 4)) = pc;
        }
    } else {
#pragma unroll
        for (int g = 4 * h; g < 4 * h + 4; ++g) {
            uint4 p = make_uint4(0, 0, 0, 0);
            if (g == 0) { p.x = pack_h2(x, y); p.y = pack_h2(z, 0.f); }
            *reinterpret_cast<uint4*>(slot_img + (uint32_t)row * 128u + (uint32_t)((g ^ (row & 7)) << 4)) = p;
        }
    }
}

FB_DEVINL void euler_xyz(const float* e, float (&R)[3][3]) {   // R = Rx(e0) Ry(e1) Rz(e2) (PyTorch3D 'XYZ')
    float sa, ca, sb, cb, sc, cc;
    sincosf(e[0], &sa, &ca); sincosf(e[1], &sb, &cb); sincosf(e[2], &sc, &cc);
    R[0][0] = cb * cc;                 R[0][1] = -cb * sc;                R[0][2] = sb;
    R[1][0] = sa * sb * cc + ca * sc;  R[1][1] = -sa * sb * sc + ca * cc; R[1][2] = -sa * cb;
    R[2][0] = -ca * sb * cc + sa * sc; R[2][1] = ca * sb * sc + sa * cc;  R[2][2] = ca * cb;
}

FB_DEVINL void bulk_wait_read(int pending) {       // this thread's bulk stores: all but the newest `pending` have read their source
    if (pending == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    else asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
}

// DIAG = true compiles the %globaltimer stamps and the timing-experiment switches in (findb200_chain_trace); the production
// instantiation carries none of them.
template <int MODE, int NT, bool DIAG>
__global__ void __launch_bounds__(kThreads2, 1) chain2_kernel(const __grid_constant__ ChainParams P) {
    auto stamp = [](const ChainParams& Pp, bool on, int slot) { if constexpr (DIAG) c2::stamp(Pp, on, slot); };
    const int dflags = DIAG ? P.dbg_flags : 0;
    if (threadIdx.x == 0 && blockIdx.x == 0) stamp(P, true, 500 + MODE * 2);
    // dynamic shared memory only, identical carve-up in both CTAs of the pair:
    // [NT activation tiles 64 KB][ring 6 x 16 KB][NT x 256 fp32 bias][mbarriers][tmem base]
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* act = smem;
    uint8_t* ring = smem + NT * kTileBytes;
    float* bias_sm = reinterpret_cast<float*>(ring + kRing2 * kSlotBytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(bias_sm + NT * kW);
    uint64_t* bar_full = bars;                          // [kRing2]
    uint64_t* bar_peer_full = bar_full + kRing2;        // [kRing2] (used in the leader)
    uint64_t* bar_empty = bar_peer_full + kRing2;       // [kRing2]
    uint64_t* bar_in_full = bar_empty + kRing2;         // [NT]
    uint64_t* bar_peer_in = bar_in_full + NT;           // [NT] (leader)
    uint64_t* bar_act_ready = bar_peer_in + NT;         // [NT] (leader)
    uint64_t* bar_acc_ready = bar_act_ready + NT;       // [NT]
    uint64_t* bar_act_free = bar_acc_ready + NT;        // [NT]
    uint32_t& tmem_base_s = *reinterpret_cast<uint32_t*>(bar_act_free + NT);
    uint32_t& ready_cnt = *(&tmem_base_s + 1);          // events (landings of both CTAs) the scout has seen complete
    if ((smem_u32(smem) & 1023u) != 0u) __trap();

    const int tid = threadIdx.x, lane = tid & 31;
    // warp index as a warp-UNIFORM value: with it and elect.sync as the single-thread guard ptxas keeps the issue path (descriptors,
    // loop counters) in uniform registers — UTCHMMA back to back — instead of wrapping every MMA in an R2UR waterfall loop
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const uint32_t rank = cluster_ctarank();
    const bool leader_cta = rank == 0;
    constexpr bool kTrunk = (MODE == TRUNK_FWD || MODE == TRUNK_BWD);
    // act_free arrivals per phase: HEAD_FWD = commit + elected store-drain + 128 final-layer readers; TRUNK_BWD = one
    constexpr uint32_t kActFreeCount = (MODE == HEAD_FWD) ? 130u : 1u;
    if (tid == 0) {
        for (int s = 0; s < kRing2; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_peer_full[s], 1); mbar_init(&bar_empty[s], 1); }
        for (int t = 0; t < NT; ++t) {
            mbar_init(&bar_in_full[t], 1);
            mbar_init(&bar_peer_in[t], 1);
            mbar_init(&bar_act_ready[t], 2);
            mbar_init(&bar_acc_ready[t], 1);
            mbar_init(&bar_act_free[t], kActFreeCount);
        }
        ready_cnt = 0;
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc2(&tmem_base_s, NT * 256);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    // Everything above touches only this CTA pair's shared memory and TMEM: with a programmatic launch it runs while the previous
    // kernel of the stream is still finishing (where an SM is free for it); global memory is first touched below.
    fb::pdl_wait_trigger();

    // ---- work decomposition: flat tile list, 2 tiles per pair-tile (one per CTA), NT pair-tiles per item --------------------
    const int n_tiles = (kTrunk ? P.trunk_feet : P.B) * P.tpf;                  // per head for the head kernels
    const int items_per_head = (n_tiles + 2 * NT - 1) / (2 * NT);
    const int n_items = kTrunk ? items_per_head : P.n_heads * items_per_head;
    const int n_clusters = gridDim.x >> 1, cluster_id = blockIdx.x >> 1;
    // item, slot j -> head, foot b, tile-in-foot t of THIS CTA (indices past the end repeat the last tile: same bytes rewritten)
    auto decode = [&](int it, int j, int& head, int& b, int& t) {
        int r = it;
        head = 0;
        if (!kTrunk) {
            const int hsel = it / items_per_head;
            head = (P.heads_mask == 2) ? 1 : hsel;
            r = it % items_per_head;
        }
        int g = (r * NT + j) * 2 + (int)rank;
        if (g > n_tiles - 1) g = n_tiles - 1;
        b = g / P.tpf;
        t = g % P.tpf;
    };

    if (warp == 0) {
        // ===================================== producer (both CTAs: own halves / own tiles) ========================
        if (elect_one()) {
            uint32_t pcount = 0, icount = 0;
            uint32_t n_free[NT];
            for (int t = 0; t < NT; ++t) n_free[t] = 0;
            auto produce = [&](const GemmW& g, int c0, int c1) {
                const int grp = g.group ? g.group : 1;
                const uint32_t half = g.chunk_bytes >> 1;
                for (int c = c0; c < c1; c += grp) {
                    const uint32_t s = pcount % kRing2, use = pcount / kRing2;
                    const int n = min(grp, c1 - c);
                    if (use > 0) mbar_wait_cluster(&bar_empty[s], (use - 1) & 1);
                    stamp(P, MODE == HEAD_FWD && cluster_id == 0 && pcount < 48 && (dflags & 1), (leader_cta ? 64 : 320) + (int)pcount);
                    if (dflags & 256) { mbar_arrive(&bar_full[s]); ++pcount; continue; }     // timing experiment: no weight traffic
                    mbar_expect_tx(&bar_full[s], (uint32_t)n * half);
                    for (int q = 0; q < n; ++q)
                        bulk_g2s(ring + s * kSlotBytes + q * half, P.wpack + g.off + (size_t)(c + q) * g.chunk_bytes + rank * half, half,
                                 &bar_full[s]);
                    ++pcount;
                }
            };
            for (int it = cluster_id; it < n_items; it += n_clusters, ++icount) {
                int head, b, t;
                if (MODE == TRUNK_FWD) {
                    // base.0 is consumed in three passes over the 9 encoded-input chunks, tile by tile within a pass
                    for (int p = 0; p < 3; ++p)
                        for (int j = 0; j < NT; ++j) produce(P.lay.fwd[P_T0], 4 * p, min(4 * p + 4, kEncChunks));
                    for (int l = 1; l < 5; ++l)
                        for (int j = 0; j < NT; ++j) produce(P.lay.fwd[P_T0 + l], 0, P.lay.fwd[P_T0 + l].n_chunks);
                } else if (MODE == HEAD_FWD) {
                    for (int j = 0; j < NT; ++j) {
                        decode(it, j, head, b, t);
                        if (icount > 0) mbar_wait_cluster(&bar_act_free[j], (icount - 1) & 1);
                        mbar_expect_tx(&bar_in_full[j], kTileBytes);
                        const size_t tt = (size_t)(P.shared ? 0 : b) * P.tpf2 + t;
                        bulk_g2s(act + j * kTileBytes, P.tact + ((size_t)4 * P.trunk_tiles + tt) * kTileBytes, kTileBytes, &bar_in_full[j]);
                    }
                    const int base = head ? P_C1 : P_D1;
                    for (int l = 0; l < 4; ++l)
                        for (int j = 0; j < NT; ++j) produce(P.lay.fwd[base + l], 0, P.lay.fwd[base + l].n_chunks);
                } else if (MODE == HEAD_BWD) {
                    decode(it, 0, head, b, t);
                    const int base = head ? P_C1 : P_D1;
                    for (int l = 3; l >= 1; --l)
                        for (int j = 0; j < NT; ++j) produce(P.lay.bwd[base + l], 0, P.lay.bwd[base + l].n_chunks);
                } else {
                    for (int hd = 0; hd < 2; ++hd) {
                        if (!(P.heads_mask & (1 << hd))) continue;
                        for (int j = 0; j < NT; ++j) {
                            decode(it, j, head, b, t);
                            if (n_free[j] > 0) mbar_wait_cluster(&bar_act_free[j], (n_free[j] - 1) & 1);   // every load but the first
                            ++n_free[j];
                            mbar_expect_tx(&bar_in_full[j], kTileBytes);
                            bulk_g2s(act + j * kTileBytes, P.tsrc[hd] + ((size_t)b * P.tpf2 + t) * kTileBytes, kTileBytes, &bar_in_full[j]);
                            produce(P.lay.bwd[hd ? P_C1 : P_D1], 0, 4);
                        }
                    }
                    for (int l = 4; l >= 1; --l)
                        for (int j = 0; j < NT; ++j) produce(P.lay.bwd[P_T0 + l], 0, 4);
                }
            }
        }
    } else if (warp == 1 || warp == 2 + kEpiWarps) {
        // ============ the issue path.  Three single-thread roles walk the SAME sequence of events (walk() below):
        //   scout   (leader CTA, warp 1): waits for "input tile landed" / "weight half landed" of BOTH CTAs and publishes the number
        //           of events that are complete in `ready_cnt` (shared memory)
        //   relay   (peer CTA, warp 1):   waits for its own CTA's landings and forwards them to the leader's mbarriers
        //   issuer  (leader CTA, warp 10): tcgen05.mma.cta_group::2 + commits; it never waits on a landing mbarrier itself, it only
        //           polls `ready_cnt`.
        // The issuer is the kernel's critical path, and one warp alone pays ~20 cycles for every taken branch and ~40 for every
        // dependent constant-bank load, so its code is straight-line: layer shapes are compile-time (they mirror make_layout() in
        // field.cu: [256 x 64] chunks, the N = 16 output layer as four 2 KB sub-chunks in one slot), every loop is unrolled, the
        // descriptors are a base plus an immediate, the role is a template argument, and diagnostics exist only in the DIAG build.
        const bool is_issuer = warp == 2 + kEpiWarps;
        if ((leader_cta || !is_issuer) && elect_one()) {
            const uint32_t act_s = smem_u32(act), ring_s = smem_u32(ring);
            const uint32_t ready_s = smem_u32(&ready_cnt);
            const uint32_t full_s = smem_u32(bar_full), pfull_s = smem_u32(bar_peer_full), empty_s = smem_u32(bar_empty);
            const uint32_t in_s = smem_u32(bar_in_full), pin_s = smem_u32(bar_peer_in), actr_s = smem_u32(bar_act_ready);
            const uint32_t accr_s = smem_u32(bar_acc_ready), actf_s = smem_u32(bar_act_free);
            // leader's mbarriers as seen from the peer
            const uint32_t r_peer_full = mapa_u32(pfull_s, 0), r_peer_in = mapa_u32(pin_s, 0);
            auto wait_bar = [&](uint32_t bar_addr, uint32_t parity, int role, uint32_t ev_now) {   // bounded spin, no clock reads
                uint32_t ok, spins = 0;
                do {
                    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                                 : "=r"(ok) : "r"(bar_addr), "r"(parity) : "memory");
                    if (!ok && ++spins > (1u << 22)) {
                        printf("find_b200: chain2 issue-path wait timeout mode %d block %d role %d event %u bar+%u parity %u\n", MODE, (int)blockIdx.x, role, ev_now,
                               bar_addr - full_s, parity);
                        __trap();
                    }
                } while (!ok);
            };
            auto commit_to = [&](uint32_t bar_addr) {
                asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                             ::"r"(bar_addr), "h"((uint16_t)3) : "memory");
            };
            auto walk = [&](auto role_tag) {
                constexpr int ROLE = decltype(role_tag)::value;            // 0 issuer, 1 scout, 2 relay
                uint32_t slot = 0, sphase = 0, ev = 0;
                uint32_t act_ph[NT], in_ph[NT];
#pragma unroll
                for (int t = 0; t < NT; ++t) act_ph[t] = in_ph[t] = 0;
                const uint64_t adesc0 = umma_desc(act_s, 16, 1024), bdesc0 = umma_desc(ring_s, 16, 1024);
                constexpr uint32_t idesc256 = umma_idesc_f16(256, 256, false, false), idesc16 = umma_idesc_f16(256, 16, false, false);
                long long cy_total0 = 0;
                if constexpr (DIAG) cy_total0 = clock64();
                // one landing event (a ring slot, or an input tile); after it `slot` still names the slot
                auto event_seen = [&]() {                                  // issuer: the scout has published event `ev`
                    ++ev;
                    uint32_t v, spins = 0;
                    for (;;) {
                        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(v) : "r"(ready_s) : "memory");
                        if ((int32_t)(v - ev) >= 0) break;
                        if (++spins > (1u << 23)) { printf("find_b200: chain2 issuer timeout mode %d block %d event %u (ready %u)\n", MODE, (int)blockIdx.x, ev, v); __trap(); }
                    }
                    tc_fence_after();
                };
                auto publish_event = [&]() {                               // scout
                    ++ev;
                    asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(ready_s), "r"(ev) : "memory");
                };
                auto slot_landed = [&]() {
                    if constexpr (ROLE == 0) event_seen();
                    else {
                        wait_bar(full_s + slot * 8u, sphase, ROLE, ev);
                        if constexpr (ROLE == 1) { wait_bar(pfull_s + slot * 8u, sphase, ROLE, ev); publish_event(); }
                        else mbar_arrive_cluster(r_peer_full + slot * 8u);
                    }
                };
                auto next_slot = [&]() { if (++slot == (uint32_t)kRing2) { slot = 0; sphase ^= 1u; } };
                auto wait_in = [&](int t) {
                    if constexpr (ROLE == 0) event_seen();
                    else {
                        wait_bar(in_s + (uint32_t)t * 8u, in_ph[t], ROLE, ev);
                        if constexpr (ROLE == 1) { wait_bar(pin_s + (uint32_t)t * 8u, in_ph[t], ROLE, ev); publish_event(); }
                        else mbar_arrive_cluster(r_peer_in + (uint32_t)t * 8u);
                    }
                    in_ph[t] ^= 1u;
                };
                auto wait_act = [&](int t) {
                    if constexpr (ROLE == 0) { wait_bar(actr_s + (uint32_t)t * 8u, act_ph[t], ROLE, ev); tc_fence_after(); }
                    act_ph[t] ^= 1u;
                };
                auto commit = [&](uint32_t bar_addr) { if constexpr (ROLE == 0) commit_to(bar_addr); };
                // D[tile t] (+)= A[tile t, images 0..NCH-1] * W chunks: NCH ring slots of one [256 n x 64 k] chunk (halves), NKLAST k-steps
                // in the last one
                auto block = [&](int t, auto nch_c, auto nklast_c, bool acc) {
                    constexpr int NCH = decltype(nch_c)::value, NKLAST = decltype(nklast_c)::value;
#pragma unroll
                    for (int c = 0; c < NCH; ++c) {
                        slot_landed();
                        if constexpr (ROLE == 0) {
                            const uint64_t ad = adesc0 + (uint64_t)(((uint32_t)t * kTileBytes + (uint32_t)c * kImgBytes) >> 4);
                            const uint64_t bd = bdesc0 + (uint64_t)((slot * (uint32_t)kSlotBytes) >> 4);
#pragma unroll
                            for (int k = 0; k < ((c == NCH - 1) ? NKLAST : 4); ++k)
                                umma2_f16(tmem + (uint32_t)t * 256u, ad + (uint64_t)(k * 2), bd + (uint64_t)(k * 2), idesc256, (acc || c > 0 || k > 0) ? 1u : 0u);
                            commit_to(empty_s + slot * 8u);
                        }
                        next_slot();
                    }
                };
                // the 256 -> 3 output layer: four [16 n x 64 k] sub-chunks (1 KB halves) in ONE slot, N = 16
                auto block_out = [&](int t) {
                    slot_landed();
                    if constexpr (ROLE == 0) {
                        const uint64_t ad = adesc0 + (uint64_t)(((uint32_t)t * kTileBytes) >> 4);
                        const uint64_t bd = bdesc0 + (uint64_t)((slot * (uint32_t)kSlotBytes) >> 4);
#pragma unroll
                        for (int sub = 0; sub < 4; ++sub)
#pragma unroll
                            for (int k = 0; k < 4; ++k)
                                umma2_f16(tmem + (uint32_t)t * 256u, ad + (uint64_t)(sub * (kImgBytes >> 4) + k * 2),
                                          bd + (uint64_t)(sub * ((kWSmallChunkBytes / 2) >> 4) + k * 2), idesc16, (sub > 0 || k > 0) ? 1u : 0u);
                        commit_to(empty_s + slot * 8u);
                    }
                    next_slot();
                };
                using I1 = std::integral_constant<int, 1>;
                using I4 = std::integral_constant<int, 4>;
                uint32_t icount = 0;
                for (int it = cluster_id; it < n_items; it += n_clusters, ++icount) {
                    if constexpr (MODE == TRUNK_FWD) {
#pragma unroll
                        for (int p = 0; p < 3; ++p)
#pragma unroll
                            for (int j = 0; j < NT; ++j) {
                                wait_act(j);
                                if (p < 2) block(j, I4{}, I4{}, p > 0); else block(j, I1{}, I1{}, true);
                                commit(accr_s + j * 8u);
                            }
#pragma unroll
                        for (int l = 1; l < 5; ++l)
#pragma unroll
                            for (int j = 0; j < NT; ++j) { wait_act(j); block(j, I4{}, I4{}, false); commit(accr_s + j * 8u); }
                    } else if constexpr (MODE == HEAD_FWD) {
#pragma unroll
                        for (int l = 0; l < 3; ++l)
#pragma unroll
                            for (int j = 0; j < NT; ++j) {
                                if (l == 0) wait_in(j); else wait_act(j);
                                if constexpr (DIAG) stamp(P, ROLE == 0 && cluster_id == 0 && icount == 0, (l * 2 + j) * 2);
                                block(j, I4{}, I4{}, false);
                                commit(accr_s + j * 8u);
                                if constexpr (DIAG) stamp(P, ROLE == 0 && cluster_id == 0 && icount == 0, (l * 2 + j) * 2 + 1);
                            }
#pragma unroll
                        for (int j = 0; j < NT; ++j) {
                            wait_act(j);
                            if constexpr (DIAG) stamp(P, ROLE == 0 && cluster_id == 0 && icount == 0, (6 + j) * 2);
                            block_out(j);
                            commit(accr_s + j * 8u);
                            commit(actf_s + j * 8u);
                            if constexpr (DIAG) stamp(P, ROLE == 0 && cluster_id == 0 && icount == 0, (6 + j) * 2 + 1);
                        }
                    } else if constexpr (MODE == HEAD_BWD) {
#pragma unroll
                        for (int j = 0; j < NT; ++j) { wait_act(j); block(j, I1{}, I1{}, false); commit(accr_s + j * 8u); }
#pragma unroll
                        for (int l = 2; l >= 1; --l)
#pragma unroll
                            for (int j = 0; j < NT; ++j) { wait_act(j); block(j, I4{}, I4{}, false); commit(accr_s + j * 8u); }
                    } else {
                        const int last_hd = (P.heads_mask & 2) ? 1 : 0;
#pragma unroll
                        for (int hd = 0; hd < 2; ++hd) {
                            if (!(P.heads_mask & (1 << hd))) continue;
#pragma unroll
                            for (int j = 0; j < NT; ++j) {
                                wait_in(j);
                                block(j, I4{}, I4{}, hd == 1 && (P.heads_mask & 1));
                                // not the last head: the buffer goes back to the producers; last: the accumulator to the epilogues
                                commit((hd == last_hd ? accr_s : actf_s) + j * 8u);
                            }
                        }
#pragma unroll
                        for (int l = 4; l >= 1; --l)
#pragma unroll
                            for (int j = 0; j < NT; ++j) { wait_act(j); block(j, I4{}, I4{}, false); commit(accr_s + j * 8u); }
                    }
                }
                if constexpr (DIAG && ROLE == 0) {
                    if (P.dbg != nullptr && cluster_id == 0) {
                        unsigned long long* o = P.dbg + 448 + MODE * 8;
                        o[0] = (unsigned long long)(clock64() - cy_total0); o[1] = 0ull; o[2] = 0ull; o[3] = 0ull; o[4] = 0ull;
                        o[5] = ev; o[6] = icount;
                    }
                }
            };
            if (is_issuer) walk(std::integral_constant<int, 0>{});
            else if (leader_cta) walk(std::integral_constant<int, 1>{});
            else walk(std::integral_constant<int, 2>{});
        }
    } else {
        // ===================================== epilogue warps (8 per CTA, one tile at a time) ==============================
        const int ew = warp - 2;
        const int h = ew >> 2;                       // column half of the accumulator this warp converts
        const int q = warp & 3;                      // TMEM lane quadrant this warp may touch
        const int row = q * 32 + lane;
        const int gt = ew * 32 + lane;               // 0..255 inside the epilogue group
        const bool elected = gt == 0;
        const uint32_t r_act_ready = mapa_u32(smem_u32(bar_act_ready), 0);
        uint32_t n_acc[NT], n_enc[NT];
        for (int t = 0; t < NT; ++t) n_acc[t] = n_enc[t] = 0;
        auto wait_acc = [&](int j) { mbar_wait_cluster(&bar_acc_ready[j], n_acc[j] & 1); ++n_acc[j]; tc_fence_after(); };
        // the elected thread owns every bulk store of this CTA; before tile j's buffer is overwritten its previous store must
        // have read it (the other tile's store, issued since, may still be pending)
        int last_store = -1;                         // (elected thread) tile of the newest bulk-store group
        auto before_overwrite = [&](int j) {
            // if the newest group belongs to the OTHER tile it may stay pending: every group of tile j is older, hence done
            if (elected) bulk_wait_read((NT == 2 && last_store >= 0 && last_store != j) ? 1 : 0);
            named_bar_sync(1, kEpiThreads);
        };
        // tile image complete: hand to the MMA (async proxy), store to HBM, tell the leader
        auto publish = [&](int j, uint8_t* gdst, bool signal) {
            fence_proxy_async();
            tc_fence_before();
            named_bar_sync(1, kEpiThreads);          // also: every thread is done with bias_s of this layer
            if (elected) {
                if (gdst && !(dflags & 2)) { bulk_s2g(gdst, act + j * kTileBytes, kTileBytes); bulk_commit(); last_store = j; }
                if (signal) mbar_arrive_cluster(r_act_ready + (uint32_t)j * 8u);
            }
        };
        auto bias_fetch = [&](const float* gb) -> float { return __ldg(gb + gt); };

        // a = relu(acc + bias): this thread's 128 columns of its row; TMEM loads run one 32-column chunk ahead
        auto epi_relu = [&](int j, uint32_t* mask_row) {
            const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + j * 256 + h * 128;
            uint8_t* act_tile = act + j * kTileBytes;
            const float* bias_s = bias_sm + j * kW;
            uint32_t va[32], vb[32], m[4] = {0u, 0u, 0u, 0u};
            // timing experiments only (findb200_chain_trace + FINDB200_CHAIN_DBG; results are garbage): 16 = no TMEM loads,
            // 8 = no conversion / shared-memory stores, 4 = no sign-word stores, 2 = no bulk stores of the tiles
            const bool no_ld = dflags & 16, no_cv = dflags & 8;
            if (!no_ld) tmem_ld32(taddr, va);
#pragma unroll
            for (int cc = 0; cc < 4; cc += 2) {
                if (!no_ld) { tmem_ld_wait(); tmem_ld32(taddr + (cc + 1) * 32, vb); }
                if (!no_cv) m[cc] = relu_chunk(va, h * 4 + cc, act_tile, row, bias_s);
                if (!no_ld) { tmem_ld_wait(); if (cc + 2 < 4) tmem_ld32(taddr + (cc + 2) * 32, va); }
                if (!no_cv) m[cc + 1] = relu_chunk(vb, h * 4 + cc + 1, act_tile, row, bias_s);
            }
            if (mask_row && !(dflags & 4)) reinterpret_cast<uint4*>(mask_row)[h] = make_uint4(m[0], m[1], m[2], m[3]);
        };
        auto epi_mask = [&](int j, const uint4& mw) {
            const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + j * 256 + h * 128;
            uint8_t* act_tile = act + j * kTileBytes;
            const uint32_t m[4] = {mw.x, mw.y, mw.z, mw.w};
            uint32_t va[32], vb[32];
            tmem_ld32(taddr, va);
#pragma unroll
            for (int cc = 0; cc < 4; cc += 2) {
                tmem_ld_wait();
                tmem_ld32(taddr + (cc + 1) * 32, vb);
                mask_chunk(va, h * 4 + cc, act_tile, row, m[cc]);
                tmem_ld_wait();
                if (cc + 2 < 4) tmem_ld32(taddr + (cc + 2) * 32, va);
                mask_chunk(vb, h * 4 + cc + 1, act_tile, row, m[cc + 1]);
            }
        };

        for (int it = cluster_id; it < n_items; it += n_clusters) {
            int head[NT], b[NT], t[NT];
            for (int j = 0; j < NT; ++j) decode(it, j, head[j], b[j], t[j]);

            if (MODE == TRUNK_FWD) {
                float x[NT], y[NT], z[NT];
                for (int j = 0; j < NT; ++j) {
                    const int vrow = t[j] * kTile + row;
                    x[j] = y[j] = z[j] = 0.f;
                    if (vrow < P.V) {
                        const float* pp = P.pos + ((size_t)b[j] * P.V + vrow) * 3;
                        x[j] = pp[0]; y[j] = pp[1]; z[j] = pp[2];
                    }
                }
                for (int p = 0; p < 3; ++p)
                    for (int j = 0; j < NT; ++j) {
                        const size_t tt = (size_t)b[j] * P.tpf2 + t[j];
                        uint8_t* act_tile = act + j * kTileBytes;
                        if (p > 0) wait_acc(j);              // the MMAs of the previous pass have read the buffer
                        before_overwrite(j);
                        const int c_lo = 4 * p, c_hi = min(4 * p + 4, kEncChunks);
                        if (P.enc_cached) {
                            // the template is constant: an earlier forward left these images in the workspace
                            if (elected) {
                                const uint32_t bytes = (uint32_t)(c_hi - c_lo) * kImgBytes;
                                mbar_expect_tx(&bar_in_full[j], bytes);
                                bulk_g2s(act_tile, P.enc + tt * kEncTileBytes + (size_t)c_lo * kImgBytes, bytes, &bar_in_full[j]);
                                mbar_wait(&bar_in_full[j], n_enc[j] & 1);
                                mbar_arrive_cluster(r_act_ready + (uint32_t)j * 8u);
                            }
                            ++n_enc[j];
                            continue;
                        }
                        for (int c = c_lo; c < c_hi; ++c)
                            encode_half_chunk(c, h, x[j], y[j], z[j], P.Bmat, act_tile + (c - c_lo) * kImgBytes, row);
                        fence_proxy_async();
                        named_bar_sync(1, kEpiThreads);
                        if (elected) {
                            if (P.save) {
                                bulk_s2g(P.enc + tt * kEncTileBytes + (size_t)c_lo * kImgBytes, act_tile, (uint32_t)(c_hi - c_lo) * kImgBytes);
                                bulk_commit();
                                last_store = j;
                            }
                            mbar_arrive_cluster(r_act_ready + (uint32_t)j * 8u);
                        }
                    }
                for (int l = 0; l < 5; ++l)
                    for (int j = 0; j < NT; ++j) {
                        const size_t tt = (size_t)b[j] * P.tpf2 + t[j];
                        const float bv = bias_fetch(P.bias[P_T0 + l]);
                        wait_acc(j);
                        bias_sm[j * kW + gt] = bv;
                        before_overwrite(j);
                        epi_relu(j, P.save ? P.tmask + (((size_t)l * P.trunk_tiles + tt) * kTile + row) * 8 : nullptr);
                        publish(j, (P.save || l == 4) ? P.tact + ((size_t)l * P.trunk_tiles + tt) * kTileBytes : nullptr, l < 4);
                    }
            } else if (MODE == HEAD_FWD) {
                for (int l = 0; l < 3; ++l)
                    for (int j = 0; j < NT; ++j) {
                        const int base = head[j] ? P_C1 : P_D1;
                        const size_t ht = (size_t)b[j] * P.tpf2 + t[j];
                        const float bv = bias_fetch(l == 0 ? P.biasfoot + ((size_t)head[j] * P.B + b[j]) * kW : P.bias[base + l]);
                        const bool tr = elected && cluster_id == 0 && it == cluster_id;
                        const int ts = (leader_cta ? 16 : 272) + (j * 3 + l) * 4;
                        wait_acc(j);
                        stamp(P, tr, ts);
                        bias_sm[j * kW + gt] = bv;
                        before_overwrite(j);
                        stamp(P, tr, ts + 1);
                        const size_t slot = ((size_t)head[j] * 3 + l) * P.B * P.tpf2 + ht;
                        epi_relu(j, P.save ? P.hmask + (slot * kTile + row) * 8 : nullptr);
                        stamp(P, tr, ts + 2);
                        publish(j, P.save ? P.hact + slot * kTileBytes : nullptr, true);
                        stamp(P, tr, ts + 3);
                    }
                // output layer: 3 of the 16 accumulator columns; warps of column half j read tile j (half 0 when NT = 1)
                if (elected) { bulk_wait_read(0); for (int j = 0; j < NT; ++j) mbar_arrive(&bar_act_free[j]); }
                for (int j = 0; j < NT; ++j) {
                    // Only the warps that READ accumulator j wait for it.  Nothing gates the next item's first MMA on tile j on the
                    // other 128 threads (bar_act_free counts the readers), so if they waited here too, a fast issuer could complete
                    // the NEXT phase of bar_acc_ready[j] before a slow one of them had looked — and it would wait for ever (seen
                    // once the issue path went straight-line).  They stay in step through the named barrier of the next item's
                    // first epilogue, which the readers only reach after this phase.
                    if (h != (j & 1)) { ++n_acc[j]; continue; }
                    wait_acc(j);
                    const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + j * 256;
                    uint32_t v[32];
                    tmem_ld32(taddr, v);
                    tmem_ld_wait();
                    tc_fence_before();
                    mbar_arrive(&bar_act_free[j]);       // accumulator of this item is free for the next one
                    const int vrow = t[j] * kTile + row;
                    if (vrow < P.V) {
                        const int base = head[j] ? P_C1 : P_D1;
                        const float* bo = P.bias[base + 3];
                        const float t0v = tanhf(__uint_as_float(v[0]) + __ldg(bo + 0));
                        const float t1v = tanhf(__uint_as_float(v[1]) + __ldg(bo + 1));
                        const float t2v = tanhf(__uint_as_float(v[2]) + __ldg(bo + 2));
                        const size_t o = ((size_t)b[j] * P.V + vrow) * 3;
                        if (head[j] == 0) {
                            const float d0 = 0.1f * t0v, d1 = 0.1f * t1v, d2 = 0.1f * t2v;     // model.py:444
                            P.disp[o] = d0; P.disp[o + 1] = d1; P.disp[o + 2] = d2;
                            if (P.verts) {
                                const float* pp = P.pos + ((size_t)(P.shared ? 0 : b[j]) * P.V + vrow) * 3;
                                float u0 = pp[0] + d0, u1 = pp[1] + d1, u2 = pp[2] + d2;
                                if (P.reg) {                                                   // model.py:481-488
                                    const float* rg = P.reg + (size_t)b[j] * 9;
                                    float R[3][3];
                                    euler_xyz(rg + 3, R);
                                    u0 *= rg[6]; u1 *= rg[7]; u2 *= rg[8];
                                    const float X0 = u0 * R[0][0] + u1 * R[1][0] + u2 * R[2][0] + rg[0];
                                    const float X1 = u0 * R[0][1] + u1 * R[1][1] + u2 * R[2][1] + rg[1];
                                    const float X2 = u0 * R[0][2] + u1 * R[1][2] + u2 * R[2][2] + rg[2];
                                    u0 = X0; u1 = X1; u2 = X2;
                                }
                                P.verts[o] = u0; P.verts[o + 1] = u1; P.verts[o + 2] = u2;
                            }
                        } else {
                            const float a0 = P.avg_col ? P.avg_col[0] : 0.f, a1 = P.avg_col ? P.avg_col[1] : 0.f,
                                        a2 = P.avg_col ? P.avg_col[2] : 0.f;
                            P.col[o] = a0 + 0.5f * (1.f + t0v);                                // model.py:446,449-451
                            P.col[o + 1] = a1 + 0.5f * (1.f + t1v);
                            P.col[o + 2] = a2 + 0.5f * (1.f + t2v);
                        }
                    }
                }
            } else if (MODE == HEAD_BWD) {
                // prologue: scaled d(pre-tanh) -> columns 0..2 of image 0 (K = 16 GEMM) and the dW operand image
                const float sc = grad_scale_from(*P.maxabs);
                for (int j = 0; j < NT; ++j) {
                    const size_t ht = (size_t)b[j] * P.tpf2 + t[j];
                    uint8_t* act_tile = act + j * kTileBytes;
                    before_overwrite(j);
                    if (h == 0) {
                        const int vrow = t[j] * kTile + row;
                        float d0 = 0.f, d1 = 0.f, d2 = 0.f;
                        if (vrow < P.V) {
                            const float4 dv = *reinterpret_cast<const float4*>(P.dovals + (((size_t)head[j] * P.B + b[j]) * P.V + vrow) * 4);
                            d0 = dv.x * sc; d1 = dv.y * sc; d2 = dv.z * sc;
                        }
                        const uint4 c0 = make_uint4(pack_h2(d0, d1), pack_h2(d2, 0.f), 0, 0), c1 = make_uint4(0, 0, 0, 0);
                        const uint32_t o0 = (uint32_t)row * 128u + (uint32_t)((0 ^ (row & 7)) << 4);
                        const uint32_t o1 = (uint32_t)row * 128u + (uint32_t)((1 ^ (row & 7)) << 4);
                        *reinterpret_cast<uint4*>(act_tile + o0) = c0;
                        *reinterpret_cast<uint4*>(act_tile + o1) = c1;
                        uint8_t* gi = P.doimg + ((size_t)head[j] * P.B * P.tpf2 + ht) * kImgBytes;
                        *reinterpret_cast<uint4*>(gi + o0) = c0;
                        *reinterpret_cast<uint4*>(gi + o1) = c1;
                    }
                    publish(j, nullptr, true);
                }
                for (int l = 2; l >= 0; --l)         // l indexes a_{l+1} / dZ_{l+1}
                    for (int j = 0; j < NT; ++j) {
                        const size_t ht = (size_t)b[j] * P.tpf2 + t[j];
                        const size_t slot = ((size_t)head[j] * 3 + l) * P.B * P.tpf2 + ht;
                        const uint4 mw = __ldg(reinterpret_cast<const uint4*>(P.hmask + (slot * kTile + row) * 8) + h);
                        wait_acc(j);
                        before_overwrite(j);
                        epi_mask(j, mw);
                        publish(j, P.hdz + slot * kTileBytes, l > 0);
                    }
            } else {   // TRUNK_BWD
                for (int l = 4; l >= 0; --l)
                    for (int j = 0; j < NT; ++j) {
                        const size_t tt = (size_t)b[j] * P.tpf2 + t[j];
                        const size_t slot = (size_t)l * P.trunk_tiles + tt;
                        const uint4 mw = __ldg(reinterpret_cast<const uint4*>(P.tmask + (slot * kTile + row) * 8) + h);
                        wait_acc(j);
                        before_overwrite(j);
                        epi_mask(j, mw);
                        publish(j, P.tdz + slot * kTileBytes, l > 0);
                    }
                if (elected) {                       // buffers free for the next item's loads
                    bulk_wait_read(0);
                    for (int j = 0; j < NT; ++j) mbar_arrive(&bar_act_free[j]);
                }
            }
        }
        if (elected) bulk_wait0();
    }
    tc_fence_before();
    cluster_sync_all();                              // the peer may still be read by the leader's MMAs / signalled remotely
    if (warp == 0) tmem_dealloc2(tmem, NT * 256);
    if (threadIdx.x == 0 && blockIdx.x == 0) stamp(P, true, 500 + MODE * 2 + 1);
}

template <int MODE, int NT, bool DIAG>
static int launch_t(const ChainParams& P, int n_items, int sms, cudaStream_t st) {
    const int smem = NT * kTileBytes + kRing2 * kSlotBytes + NT * kW * 4 + (3 * kRing2 + 5 * NT + 1) * 8;
    static bool attr_set = false;      // one process drives one GPU (one rank per GPU)
    if (!attr_set) {
        FB_CUDA(cudaFuncSetAttribute(chain2_kernel<MODE, NT, DIAG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    int clusters = sms / 2;
    if (n_items < clusters) clusters = n_items;
    if (clusters <= 0) return 0;
    static const char* const kNames[4] = {"trunk_fwd", "head_fwd", "head_bwd", "trunk_bwd"};
    FB_MARK(kNames[MODE], st);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * clusters); cfg.blockDim = dim3(kThreads2); cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = fb::pdl_enabled() ? 2 : 1;
    FB_CUDA(cudaLaunchKernelEx(&cfg, chain2_kernel<MODE, NT, DIAG>, P));
    FB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace c2

// mode: TRUNK_FWD .. TRUNK_BWD; picks NT (pair-tiles per item) and the item count from the shapes in P
int launch_chain2(int mode, const ChainParams& P, int sms, cudaStream_t st) {
    using namespace c2;
    const bool trunk = (mode == TRUNK_FWD || mode == TRUNK_BWD);
    const int n_tiles = (trunk ? P.trunk_feet : P.B) * P.tpf;
    const int pairs = (n_tiles + 1) / 2;
    const int clusters = sms / 2;
    // trunk: one pair-tile per item while that still leaves clusters idle or evenly loaded (latency), two when there are many
    const int nt = trunk ? ((pairs <= 2 * clusters) ? 1 : 2) : 2;
    const int per_head = (n_tiles + 2 * nt - 1) / (2 * nt);
    const int n_items = trunk ? per_head : P.n_heads * per_head;
    switch (mode * 2 + (nt - 1)) {
        case 0: return launch_t<TRUNK_FWD, 1, false>(P, n_items, sms, st);
        case 1: return launch_t<TRUNK_FWD, 2, false>(P, n_items, sms, st);
        case 3: return P.dbg ? launch_t<HEAD_FWD, 2, true>(P, n_items, sms, st) : launch_t<HEAD_FWD, 2, false>(P, n_items, sms, st);
        case 5: return launch_t<HEAD_BWD, 2, false>(P, n_items, sms, st);
        case 6: return launch_t<TRUNK_BWD, 1, false>(P, n_items, sms, st);
        case 7: return launch_t<TRUNK_BWD, 2, false>(P, n_items, sms, st);
    }
    set_error("launch_chain2: bad mode %d nt %d", mode, nt);
    return FINDB200_ERR_ARG;
}

}  // namespace field
}  // namespace fb
