// find_b200/csrc/field_chain.cu — the "chain" kernels of the displacement field: a tile of 128 (or 2 x 128) rows
// is pushed through a sequence of 256-wide GEMMs without leaving the SM.
//
//   TRUNK_FWD : encode xyz (Fourier features, fp32 sincos -> fp16) -> base.0 .. base.8 (+ReLU)      model.py:421-426
//   HEAD_FWD  : trunk feature tile -> mlp_disp / mlp_col (latents folded into a per-foot bias)     model.py:428-451
//               -> tanh epilogue, registration transform                                          model.py:481-488
//   HEAD_BWD  : d(pre-tanh) -> dZ3 -> dZ2 -> dZ1 of one head (ReLU masks re-read from the saved activations)
//   TRUNK_BWD : sum over feet/heads of dZ1 -> x W1h^T -> dZ of base.8 .. base.0
//
// Per CTA: warp 0 = weight/input producer (cp.async.bulk into a 3 x 32 KB ring), warp 1 = tcgen05.mma issuer,
// warps 2.. = epilogue (4 warps per tile: TMEM -> registers -> bias/ReLU/mask -> fp16 tile image in shared memory,
// which is the A operand of the next GEMM and is also bulk-stored to HBM for the weight-gradient kernel).
// Activations of a tile live in shared memory (64 KB per tile, overwritten in place layer by layer) and the
// accumulator [128 x 256] fp32 in TMEM (256 columns per tile).
#include "field.cuh"
#include "field_chain_params.cuh"

namespace fb {
namespace field {

unsigned long long* g_chain_dbg = nullptr;
int g_chain_dbg_flags = 1;

FB_DEVINL void dbg_stamp(const ChainParams& P, bool first, int slot) {
    if (P.dbg && first && blockIdx.x == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        P.dbg[slot] = t;
    }
}

FB_DEVINL void named_bar_sync(int id, int count) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory"); }

// power-of-two gradient scale that puts max |d(pre-tanh)| into [8, 16): fp16 operands keep full precision
FB_DEVINL float grad_scale_from(uint32_t maxabs_bits) {
    const float m = __uint_as_float(maxabs_bits);
    if (!(m > 0.f) || !(m < 3.0e38f)) return 1.0f;
    int e;
    frexpf(m, &e);                       // m = f * 2^e, f in [0.5, 1)
    return ldexpf(1.0f, 4 - e);          // m * scale in [8, 16)
}

FB_DEVINL uint32_t pack_h2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

// ---- producer: stream one weight operand through the ring ------------------------------------------------
FB_DEVINL void produce_gemm(const ChainParams& P, const GemmW& g, uint8_t* ring, uint64_t* bar_full, uint64_t* bar_empty,
                            uint32_t& pcount, int c0, int c1) {
    const int grp = g.group ? g.group : 1;
    for (int c = c0; c < c1; c += grp) {
        const uint32_t s = pcount % kRing, use = pcount / kRing;
        const uint32_t bytes = (uint32_t)min(grp, c1 - c) * g.chunk_bytes;
        if (use > 0) mbar_wait(&bar_empty[s], (use - 1) & 1);
        mbar_expect_tx(&bar_full[s], bytes);
        bulk_g2s(ring + s * kWChunkBytes, P.wpack + g.off + (size_t)c * g.chunk_bytes, bytes, &bar_full[s]);
        dbg_stamp(P, pcount < 32, 64 + (int)pcount);
        ++pcount;
    }
}

// ---- MMA issuer: D[tile] (+)= A[tile](smem images) * W(ring) for chunks [c0, c1) ------------------------------
FB_DEVINL void mma_gemm(const GemmW& g, uint32_t act_s, uint32_t ring_s, uint32_t tmem, uint64_t* bar_full,
                        uint64_t* bar_empty, uint32_t& ccount, int c0, int c1, bool acc_init, const ChainParams* dbgP = nullptr) {
    const uint32_t idesc = umma_idesc_f16(128, g.N, false, false);
    const int grp = g.group ? g.group : 1;
    for (int c = c0; c < c1; c += grp) {
        const uint32_t s = ccount % kRing, use = ccount / kRing;
        mbar_wait(&bar_full[s], use & 1);
        tc_fence_after();
        if (dbgP) dbg_stamp(*dbgP, ccount < 32, 96 + (int)ccount);
        for (int sub = 0; sub < grp && c + sub < c1; ++sub) {
            const int cc = c + sub;
            const int nk = (cc == g.n_chunks - 1) ? g.k16_last : 4;
            for (int k = 0; k < nk; ++k) {
                const uint64_t ad = umma_desc(act_s + (cc - c0) * kImgBytes + k * 32, 16, 1024);
                const uint64_t bd = umma_desc(ring_s + s * kWChunkBytes + sub * g.chunk_bytes + k * 32, 16, 1024);
                umma_f16(tmem, ad, bd, idesc, (acc_init || cc > c0 || k > 0) ? 1u : 0u);
            }
        }
        umma_commit(&bar_empty[s]);
        ++ccount;
    }
}

// ---- two tiles, ONE pass over the weight chunks: every chunk that lands in the ring feeds both tiles' accumulators ------
// The chain kernels are bound by the arrival rate of weight chunks (one 32 KB chunk per ~0.5 us per SM with 3 ring slots
// and ~0.8 us L2 latency, measured with findb200_chain_trace), not by the tensor pipe (0.28 us of MMA per chunk and tile),
// so halving the number of chunks per layer matters more than overlapping one tile's epilogue with the other tile's MMAs.
// Measured (tools/chain_trace.py, profiles/r1_chain_trace.md): sharing is SLOWER than the per-tile ping-pong (head_fwd 110 -> 128 us):
// a slot is held for two tiles' MMAs, so chunks arrive every 0.75-0.9 us instead of 0.5 us and the two epilogues no longer
// hide behind the other tile's MMAs.  Kept behind kShareChunks = false as the record of that experiment.
constexpr bool kShareChunks = false;
template <typename WaitT1>
FB_DEVINL void mma_gemm2(const GemmW& g, uint32_t act_s, uint32_t ring_s, uint32_t tmem, uint64_t* bar_full, uint64_t* bar_empty,
                         uint32_t& ccount, WaitT1&& wait_tile1, const ChainParams* dbgP = nullptr) {
    const uint32_t idesc = umma_idesc_f16(128, g.N, false, false);
    for (int c = 0; c < g.n_chunks; ++c) {
        const uint32_t s = ccount % kRing, use = ccount / kRing;
        mbar_wait(&bar_full[s], use & 1);
        tc_fence_after();
        if (dbgP) dbg_stamp(*dbgP, ccount < 32, 96 + (int)ccount);
        const int nk = (c == g.n_chunks - 1) ? g.k16_last : 4;
#pragma unroll
        for (int t = 0; t < 2; ++t) {
            if (t == 1 && c == 0) wait_tile1();          // tile 1's operand is only needed now: tile 0 starts earlier
            for (int k = 0; k < nk; ++k) {
                const uint64_t ad = umma_desc(act_s + t * kTileBytes + c * kImgBytes + k * 32, 16, 1024);
                const uint64_t bd = umma_desc(ring_s + s * kWChunkBytes + k * 32, 16, 1024);
                umma_f16(tmem + t * 256, ad, bd, idesc, (c > 0 || k > 0) ? 1u : 0u);
            }
        }
        umma_commit(&bar_empty[s]);
        ++ccount;
    }
}

// ---- epilogues (one thread = one row of the tile) ---------------------------------------------------------
FB_DEVINL uint32_t img_row_off(int chunk32, int row) {   // byte offset of 16-B chunk `chunk32` (0..31) of `row`
    return (uint32_t)(chunk32 >> 3) * kImgBytes + (uint32_t)row * 128u + (uint32_t)(((chunk32 & 7) ^ (row & 7)) << 4);
}

// 32 accumulator columns of one row: + bias (shared memory), ReLU, fp16, into the tile image; returns the ReLU sign
// bits of the 32 columns: column 2j -> bit 15-j, column 2j+1 -> bit 31-j (j = 0..15), i.e. of packed word j.
FB_DEVINL uint32_t relu_chunk(const uint32_t (&v)[32], int cc, uint8_t* act_tile, int row, const float* bias_s) {
    const float4* b4 = reinterpret_cast<const float4*>(bias_s + cc * 32);
    uint32_t bits = 0;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        const float4 b0 = b4[g * 2], b1 = b4[g * 2 + 1];
        uint4 pk;
        pk.x = pack_h2(fmaxf(__uint_as_float(v[g * 8 + 0]) + b0.x, 0.f), fmaxf(__uint_as_float(v[g * 8 + 1]) + b0.y, 0.f));
        pk.y = pack_h2(fmaxf(__uint_as_float(v[g * 8 + 2]) + b0.z, 0.f), fmaxf(__uint_as_float(v[g * 8 + 3]) + b0.w, 0.f));
        pk.z = pack_h2(fmaxf(__uint_as_float(v[g * 8 + 4]) + b1.x, 0.f), fmaxf(__uint_as_float(v[g * 8 + 5]) + b1.y, 0.f));
        pk.w = pack_h2(fmaxf(__uint_as_float(v[g * 8 + 6]) + b1.z, 0.f), fmaxf(__uint_as_float(v[g * 8 + 7]) + b1.w, 0.f));
        *reinterpret_cast<uint4*>(act_tile + img_row_off(cc * 4 + g, row)) = pk;
        // a packed half >= +0 is non-zero iff adding 0x7fff carries into its top bit (no carry crosses the halves)
        bits |= ((pk.x + 0x7fff7fffu) & 0x80008000u) >> (g * 4 + 0);
        bits |= ((pk.y + 0x7fff7fffu) & 0x80008000u) >> (g * 4 + 1);
        bits |= ((pk.z + 0x7fff7fffu) & 0x80008000u) >> (g * 4 + 2);
        bits |= ((pk.w + 0x7fff7fffu) & 0x80008000u) >> (g * 4 + 3);
    }
    return bits;
}

// a = relu(acc + bias) for one row of the tile; TMEM loads run one 32-column chunk ahead of the arithmetic.
// mask_row: global [8] words for this row's sign bits (saved for the backward pass), or null.
FB_DEVINL void epi_relu(uint32_t taddr, uint8_t* act_tile, int row, const float* bias_s, uint32_t* mask_row) {
    uint32_t va[32], vb[32], m[8];
    tmem_ld32(taddr, va);
#pragma unroll
    for (int cc = 0; cc < 8; cc += 2) {
        tmem_ld_wait();
        tmem_ld32(taddr + (cc + 1) * 32, vb);
        m[cc] = relu_chunk(va, cc, act_tile, row, bias_s);
        tmem_ld_wait();
        if (cc + 2 < 8) tmem_ld32(taddr + (cc + 2) * 32, va);
        m[cc + 1] = relu_chunk(vb, cc + 1, act_tile, row, bias_s);
    }
    if (mask_row) {
        reinterpret_cast<uint4*>(mask_row)[0] = make_uint4(m[0], m[1], m[2], m[3]);
        reinterpret_cast<uint4*>(mask_row)[1] = make_uint4(m[4], m[5], m[6], m[7]);
    }
}

FB_DEVINL uint32_t mask_pack(uint32_t bits, int j, float lo, float hi) {   // sign bits of packed word j, see relu_chunk
    const float l = (bits & (0x00008000u >> j)) ? lo : 0.f;
    const float h = (bits & (0x80000000u >> j)) ? hi : 0.f;
    return pack_h2(l, h);
}

FB_DEVINL void mask_chunk(const uint32_t (&v)[32], int cc, uint8_t* act_tile, int row, uint32_t bits) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
        uint4 pk;
        pk.x = mask_pack(bits, g * 4 + 0, __uint_as_float(v[g * 8 + 0]), __uint_as_float(v[g * 8 + 1]));
        pk.y = mask_pack(bits, g * 4 + 1, __uint_as_float(v[g * 8 + 2]), __uint_as_float(v[g * 8 + 3]));
        pk.z = mask_pack(bits, g * 4 + 2, __uint_as_float(v[g * 8 + 4]), __uint_as_float(v[g * 8 + 5]));
        pk.w = mask_pack(bits, g * 4 + 3, __uint_as_float(v[g * 8 + 6]), __uint_as_float(v[g * 8 + 7]));
        *reinterpret_cast<uint4*>(act_tile + img_row_off(cc * 4 + g, row)) = pk;
    }
}

// dZ = dA * [a > 0]; m = the row's 8 sign words saved by the forward epilogue (loaded by the caller before it waits)
FB_DEVINL void epi_mask(uint32_t taddr, uint8_t* act_tile, int row, const uint32_t (&m)[8]) {
    uint32_t va[32], vb[32];
    tmem_ld32(taddr, va);
#pragma unroll
    for (int cc = 0; cc < 8; cc += 2) {
        tmem_ld_wait();
        tmem_ld32(taddr + (cc + 1) * 32, vb);
        mask_chunk(va, cc, act_tile, row, m[cc]);
        tmem_ld_wait();
        if (cc + 2 < 8) tmem_ld32(taddr + (cc + 2) * 32, va);
        mask_chunk(vb, cc + 1, act_tile, row, m[cc + 1]);
    }
}

FB_DEVINL void load_mask_row(const uint32_t* mask_row, uint32_t (&m)[8]) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(mask_row)), b = __ldg(reinterpret_cast<const uint4*>(mask_row) + 1);
    m[0] = a.x; m[1] = a.y; m[2] = a.z; m[3] = a.w; m[4] = b.x; m[5] = b.y; m[6] = b.z; m[7] = b.w;
}

// Fourier features of one row for encoded-input chunk c (64 internal columns), written to image slot `slot`.
// chunk c < 8: [sin(2 pi r_f) f = 32c..32c+31 | cos(2 pi r_f) same f],  r_f = x.B[:,f]; chunk 8: [x, y, z, 0...]
FB_DEVINL void encode_chunk(int c, float x, float y, float z, const float* __restrict__ Bm, uint8_t* slot_img, int row,
                            uint8_t* g_img) {
    if (c < 8) {
#pragma unroll 1
        for (int g = 0; g < 4; ++g) {
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
            const uint32_t os = (uint32_t)row * 128u + (uint32_t)((g ^ (row & 7)) << 4);
            const uint32_t oc = (uint32_t)row * 128u + (uint32_t)(((g + 4) ^ (row & 7)) << 4);
            *reinterpret_cast<uint4*>(slot_img + os) = ps;
            *reinterpret_cast<uint4*>(slot_img + oc) = pc;
            if (g_img) {
                *reinterpret_cast<uint4*>(g_img + os) = ps;
                *reinterpret_cast<uint4*>(g_img + oc) = pc;
            }
        }
    } else {
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            uint4 p = make_uint4(0, 0, 0, 0);
            if (g == 0) { p.x = pack_h2(x, y); p.y = pack_h2(z, 0.f); }
            const uint32_t o = (uint32_t)row * 128u + (uint32_t)((g ^ (row & 7)) << 4);
            *reinterpret_cast<uint4*>(slot_img + o) = p;
            if (g_img) *reinterpret_cast<uint4*>(g_img + o) = p;
        }
    }
}

// rotation of the registration transform: R = Rx(e0) Ry(e1) Rz(e2) (PyTorch3D euler_angles_to_matrix 'XYZ')
FB_DEVINL void euler_xyz(const float* e, float (&R)[3][3]) {
    float sa, ca, sb, cb, sc, cc;
    sincosf(e[0], &sa, &ca); sincosf(e[1], &sb, &cb); sincosf(e[2], &sc, &cc);
    R[0][0] = cb * cc;                 R[0][1] = -cb * sc;                R[0][2] = sb;
    R[1][0] = sa * sb * cc + ca * sc;  R[1][1] = -sa * sb * sc + ca * cc; R[1][2] = -sa * cb;
    R[2][0] = -ca * sb * cc + sa * sc; R[2][1] = ca * sb * sc + sa * cc;  R[2][2] = ca * cb;
}

template <int MODE, int NT>
__global__ void __launch_bounds__(64 + 128 * NT, 1) chain_kernel(const __grid_constant__ ChainParams P) {
    fb::pdl_wait_trigger();
    if (threadIdx.x == 0) dbg_stamp(P, true, 200 + MODE * 2);
    // dynamic shared memory only (no static __shared__), so the window starts 1024-aligned as SWIZZLE_128B needs:
    // [NT activation tiles 64 KB][weight ring 3 x 32 KB][NT x 256 fp32 bias of the layer in flight][mbarriers]
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* act = smem;
    uint8_t* ring = smem + NT * kTileBytes;
    float* bias_sm = reinterpret_cast<float*>(ring + kRing * kWChunkBytes);
    uint64_t* bars = reinterpret_cast<uint64_t*>(bias_sm + NT * kW);
    uint64_t* bar_full = bars;                    // [kRing]
    uint64_t* bar_empty = bars + kRing;           // [kRing]
    // per tile slot: the NT tiles of a CTA run half a layer out of phase (tile t's epilogue overlaps tile t^1's MMAs)
    uint64_t* bar_act_ready = bars + 2 * kRing;   // [NT]
    uint64_t* bar_acc_ready = bar_act_ready + NT; // [NT]
    uint64_t* bar_in_full = bar_acc_ready + NT;   // [NT]
    uint64_t* bar_act_free = bar_in_full + NT;    // [NT]
    uint32_t& tmem_base_s = *reinterpret_cast<uint32_t*>(bar_act_free + NT);
    if ((smem_u32(smem) & 1023u) != 0u) __trap();

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);    // warp-uniform: keeps the issue path in uniform registers (see field_chain2.cu)
    if (tid == 0) {
        for (int s = 0; s < kRing; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1); }
        for (int t = 0; t < NT; ++t) {
            mbar_init(&bar_act_ready[t], 128);
            mbar_init(&bar_acc_ready[t], 1);
            mbar_init(&bar_in_full[t], 1);
            mbar_init(&bar_act_free[t], MODE == TRUNK_BWD ? 1 : 1 + 128);
        }
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base_s, NT * 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const int groups_per_foot = P.tpf2 / NT;
    const int items_per_head = P.B * groups_per_foot;

    // decode an item -> head, foot, first tile (heads) or first trunk tile (trunk)
    auto decode = [&](int it, int& head, int& b, int& t0) {
        if (MODE == TRUNK_FWD || MODE == TRUNK_BWD) {
            head = 0; b = (it * NT) / P.tpf2; t0 = (it * NT) % P.tpf2;
        } else {
            const int hsel = it / items_per_head;
            head = (P.heads_mask == 2) ? 1 : hsel;       // mask 1: disp only, 2: col only, 3: both (hsel = head)
            const int r = it % items_per_head;
            b = r / groups_per_foot; t0 = (r % groups_per_foot) * NT;
        }
    };

    if (warp == 0) {
        // ===================================== producer ===========================================
        if (elect_one()) {
            uint32_t pcount = 0, icount = 0;
            uint32_t n_free[NT];
            for (int t = 0; t < NT; ++t) n_free[t] = 0;
            for (int it = blockIdx.x; it < P.n_items; it += gridDim.x, ++icount) {
                int head, b, t0;
                decode(it, head, b, t0);
                // weights are streamed once per (layer, tile), in the order the MMA warp consumes them
                if (MODE == TRUNK_FWD) {
                    // base.0 is consumed in three passes over the 9 encoded-input chunks, tile by tile within a pass: the ring
                    // must carry the chunks in exactly that order (with NT = 2 a per-tile order would hand tile 1 the wrong ones)
                    for (int p = 0; p < 3; ++p)
                        for (int t = 0; t < NT; ++t)
                            produce_gemm(P, P.lay.fwd[P_T0], ring, bar_full, bar_empty, pcount, 4 * p, min(4 * p + 4, kEncChunks));
                    for (int l = 1; l < 5; ++l)
                        for (int t = 0; t < NT; ++t)
                            produce_gemm(P, P.lay.fwd[P_T0 + l], ring, bar_full, bar_empty, pcount, 0, P.lay.fwd[P_T0 + l].n_chunks);
                } else if (MODE == HEAD_FWD) {
                    for (int t = 0; t < NT; ++t) {
                        if (icount > 0) mbar_wait(&bar_act_free[t], (icount - 1) & 1);
                        mbar_expect_tx(&bar_in_full[t], kTileBytes);
                        const size_t tt = (size_t)(P.shared ? 0 : b) * P.tpf2 + t0 + t;
                        bulk_g2s(act + t * kTileBytes, P.tact + ((size_t)4 * P.trunk_tiles + tt) * kTileBytes, kTileBytes, &bar_in_full[t]);
                    }
                    const int base = head ? P_C1 : P_D1;
                    for (int l = 0; l < 4; ++l)
                        for (int t = 0; t < ((NT == 2 && kShareChunks) ? 1 : NT); ++t)
                            produce_gemm(P, P.lay.fwd[base + l], ring, bar_full, bar_empty, pcount, 0, P.lay.fwd[base + l].n_chunks);
                } else if (MODE == HEAD_BWD) {
                    const int base = head ? P_C1 : P_D1;
                    for (int l = 3; l >= 1; --l)
                        for (int t = 0; t < ((NT == 2 && kShareChunks) ? 1 : NT); ++t)
                            produce_gemm(P, P.lay.bwd[base + l], ring, bar_full, bar_empty, pcount, 0, P.lay.bwd[base + l].n_chunks);
                } else {
                    // A operand of the first GEMM(s): sum over feet of dZ1 of each head (P.tsrc, made by sum_feet_kernel), bulk
                    // loaded into the tile's activation buffer once the buffer's previous reader is done
                    for (int hd = 0; hd < 2; ++hd) {
                        if (!(P.heads_mask & (1 << hd))) continue;
                        for (int t = 0; t < NT; ++t) {
                            if (n_free[t] > 0) mbar_wait(&bar_act_free[t], (n_free[t] - 1) & 1);   // every load but the first
                            ++n_free[t];
                            mbar_expect_tx(&bar_in_full[t], kTileBytes);
                            bulk_g2s(act + t * kTileBytes, P.tsrc[hd] + ((size_t)b * P.tpf2 + t0 + t) * kTileBytes, kTileBytes, &bar_in_full[t]);
                            produce_gemm(P, P.lay.bwd[hd ? P_C1 : P_D1], ring, bar_full, bar_empty, pcount, 0, 4);
                        }
                    }
                    for (int l = 4; l >= 1; --l)
                        for (int t = 0; t < NT; ++t) produce_gemm(P, P.lay.bwd[P_T0 + l], ring, bar_full, bar_empty, pcount, 0, 4);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================== MMA issuer =========================================
        if (elect_one()) {
            uint32_t ccount = 0, icount = 0;
            uint32_t n_act[NT], n_in[NT];
            for (int t = 0; t < NT; ++t) n_act[t] = n_in[t] = 0;
            const uint32_t act_s = smem_u32(act), ring_s = smem_u32(ring);
            auto wait_act = [&](int t) { mbar_wait(&bar_act_ready[t], n_act[t] & 1); ++n_act[t]; tc_fence_after(); };
            // one layer of one tile; then straight on to the other tile while this one's epilogue runs
            auto gemm = [&](int t, const GemmW& g, int c0, int c1, bool acc_init) {
                mma_gemm(g, act_s + t * kTileBytes, ring_s, tmem + t * 256, bar_full, bar_empty, ccount, c0, c1, acc_init, MODE == HEAD_FWD ? &P : nullptr);
                umma_commit(&bar_acc_ready[t]);
            };
            for (int it = blockIdx.x; it < P.n_items; it += gridDim.x, ++icount) {
                int head, b, t0;
                decode(it, head, b, t0);
                if (MODE == TRUNK_FWD) {
                    for (int p = 0; p < 3; ++p)
                        for (int t = 0; t < NT; ++t) {
                            wait_act(t);
                            dbg_stamp(P, icount == 0 && t == 0, 128 + p * 2);
                            gemm(t, P.lay.fwd[P_T0], 4 * p, min(4 * p + 4, kEncChunks), p > 0);
                            dbg_stamp(P, icount == 0 && t == 0, 128 + p * 2 + 1);
                        }
                    for (int l = 1; l < 5; ++l)
                        for (int t = 0; t < NT; ++t) {
                            wait_act(t);
                            dbg_stamp(P, icount == 0 && t == 0, 128 + (2 + l) * 2);
                            gemm(t, P.lay.fwd[P_T0 + l], 0, 4, false);
                            dbg_stamp(P, icount == 0 && t == 0, 128 + (2 + l) * 2 + 1);
                        }
                } else if (MODE == HEAD_FWD) {
                    const int base = head ? P_C1 : P_D1;
                    if (NT == 2 && kShareChunks) {
                        for (int l = 0; l < 4; ++l) {
                            auto wait_in = [&](int t) {
                                if (l == 0) { mbar_wait(&bar_in_full[t], icount & 1); tc_fence_after(); }
                                else wait_act(t);
                            };
                            wait_in(0);
                            dbg_stamp(P, icount == 0, l * 4);
                            mma_gemm2(P.lay.fwd[base + l], act_s, ring_s, tmem, bar_full, bar_empty, ccount, [&]() { wait_in(1); }, &P);
                            umma_commit(&bar_acc_ready[0]);
                            umma_commit(&bar_acc_ready[1]);
                            if (l == 3) { umma_commit(&bar_act_free[0]); umma_commit(&bar_act_free[1]); }
                            dbg_stamp(P, icount == 0, l * 4 + 1);
                        }
                    } else {
                        for (int l = 0; l < 4; ++l)
                            for (int t = 0; t < NT; ++t) {
                                if (l == 0) { mbar_wait(&bar_in_full[t], icount & 1); tc_fence_after(); }
                                else wait_act(t);
                                dbg_stamp(P, icount == 0, (l * 2 + t) * 2);
                                gemm(t, P.lay.fwd[base + l], 0, 4, false);
                                dbg_stamp(P, icount == 0, (l * 2 + t) * 2 + 1);
                                if (l == 3) umma_commit(&bar_act_free[t]);
                            }
                    }
                } else if (MODE == HEAD_BWD) {
                    const int base = head ? P_C1 : P_D1;
                    if (NT == 2 && kShareChunks) {
                        for (int l = 3; l >= 1; --l) {
                            wait_act(0);
                            mma_gemm2(P.lay.bwd[base + l], act_s, ring_s, tmem, bar_full, bar_empty, ccount, [&]() { wait_act(1); });
                            umma_commit(&bar_acc_ready[0]);
                            umma_commit(&bar_acc_ready[1]);
                        }
                    } else {
                        for (int l = 3; l >= 1; --l)
                            for (int t = 0; t < NT; ++t) {
                                wait_act(t);
                                gemm(t, P.lay.bwd[base + l], 0, P.lay.bwd[base + l].n_chunks, false);
                            }
                    }
                } else {
                    const int last_hd = (P.heads_mask & 2) ? 1 : 0;
                    for (int hd = 0; hd < 2; ++hd) {
                        if (!(P.heads_mask & (1 << hd))) continue;
                        for (int t = 0; t < NT; ++t) {
                            mbar_wait(&bar_in_full[t], n_in[t] & 1); ++n_in[t];
                            tc_fence_after();
                            mma_gemm(P.lay.bwd[hd ? P_C1 : P_D1], act_s + t * kTileBytes, ring_s, tmem + t * 256, bar_full, bar_empty,
                                     ccount, 0, 4, hd == 1 && (P.heads_mask & 1));
                            // not the last head: the buffer goes back to the producer; last: the accumulator to the epilogue
                            umma_commit(hd == last_hd ? &bar_acc_ready[t] : &bar_act_free[t]);
                        }
                    }
                    for (int l = 4; l >= 1; --l)
                        for (int t = 0; t < NT; ++t) { wait_act(t); gemm(t, P.lay.bwd[P_T0 + l], 0, 4, false); }
                }
            }
        }
    } else {
        // ===================================== epilogue warps =====================================
        const int ew = warp - 2;
        const int tl = ew >> 2;                      // tile of the group this warp serves
        const int q = warp & 3;                      // TMEM lane quadrant this warp may touch
        const int row = q * 32 + lane;
        const int gt = (ew & 3) * 32 + lane;         // thread index inside the tile's 128-thread group
        const bool leader = gt == 0;
        uint8_t* act_tile = act + tl * kTileBytes;
        const uint32_t taddr = tmem + ((uint32_t)(q * 32) << 16) + tl * 256;
        uint32_t n_acc = 0, n_enc = 0;
        auto wait_acc = [&]() { mbar_wait(&bar_acc_ready[tl], n_acc & 1); ++n_acc; tc_fence_after(); };
        auto before_overwrite = [&]() {              // pending bulk store of this tile must have read smem
            if (leader) bulk_wait_read0();
            named_bar_sync(1 + tl, 128);
        };
        auto publish = [&](uint8_t* gdst) {          // tile image complete: hand to the MMA proxy, store to HBM
            fence_proxy_async();
            named_bar_sync(1 + tl, 128);             // also: every thread is done with bias_s of this layer
            if (gdst && leader) { bulk_s2g(gdst, act_tile, kTileBytes); bulk_commit(); }
        };
        float* bias_s = bias_sm + tl * kW;
        // bias of the next layer: global -> registers before the accumulator wait, -> shared memory after it (the
        // barrier in before_overwrite() publishes it to the tile's 128 threads)
        auto bias_fetch = [&](const float* gb) -> float2 { return __ldg(reinterpret_cast<const float2*>(gb) + gt); };
        auto bias_put = [&](float2 v) { reinterpret_cast<float2*>(bias_s)[gt] = v; };
        auto signal_act = [&]() { tc_fence_before(); mbar_arrive(&bar_act_ready[tl]); };

        for (int it = blockIdx.x; it < P.n_items; it += gridDim.x) {
            int head, b, t0;
            decode(it, head, b, t0);
            const int t = t0 + tl;                   // tile inside the foot
            const int vrow = t * kTile + row;        // vertex index of this thread's row
            const bool valid = vrow < P.V;

            if (MODE == TRUNK_FWD) {
                const size_t tt = (size_t)b * P.tpf2 + t;
                float x = 0.f, y = 0.f, z = 0.f;
                if (valid) {
                    const float* pp = P.pos + ((size_t)b * P.V + vrow) * 3;
                    x = pp[0]; y = pp[1]; z = pp[2];
                }
                const bool first = leader && tl == 0 && it == (int)blockIdx.x;
                for (int p = 0; p < 3; ++p) {
                    if (p > 0) wait_acc();
                    before_overwrite();                  // the previous bulk store of this buffer has read it
                    dbg_stamp(P, first, 160 + p * 2);
                    const int c_lo = 4 * p, c_hi = min(4 * p + 4, kEncChunks);
                    if (P.enc_cached) {
                        // the template is constant: an earlier forward left these images in the workspace, one bulk load
                        // replaces 128 sin/cos pairs per row and pass
                        if (leader) {
                            const uint32_t bytes = (uint32_t)(c_hi - c_lo) * kImgBytes;
                            mbar_expect_tx(&bar_in_full[tl], bytes);
                            bulk_g2s(act_tile, P.enc + tt * kEncTileBytes + (size_t)c_lo * kImgBytes, bytes, &bar_in_full[tl]);
                        }
                        mbar_wait(&bar_in_full[tl], n_enc & 1);
                        ++n_enc;
                        signal_act();
                        dbg_stamp(P, first, 160 + p * 2 + 1);
                        continue;
                    }
                    for (int c = c_lo; c < c_hi; ++c)
                        encode_chunk(c, x, y, z, P.Bmat, act_tile + (c - c_lo) * kImgBytes, row, nullptr);
                    // the saved copy for the dW kernel leaves as ONE bulk store of the finished images (the per-thread
                    // 16-byte global stores it replaces touched 32 lines per warp instruction: 4 us per pass)
                    fence_proxy_async();
                    if (P.save) {
                        named_bar_sync(1 + tl, 128);
                        if (leader) {
                            bulk_s2g(P.enc + tt * kEncTileBytes + (size_t)c_lo * kImgBytes, act_tile, (uint32_t)(c_hi - c_lo) * kImgBytes);
                            bulk_commit();
                        }
                    }
                    signal_act();
                    dbg_stamp(P, first, 160 + p * 2 + 1);
                }
                for (int l = 0; l < 5; ++l) {
                    dbg_stamp(P, first, 170 + l * 3);
                    const float2 bv = bias_fetch(P.bias[P_T0 + l]);
                    wait_acc();
                    dbg_stamp(P, first, 170 + l * 3 + 1);
                    bias_put(bv);
                    before_overwrite();
                    epi_relu(taddr, act_tile, row, bias_s,
                             P.save ? P.tmask + (((size_t)l * P.trunk_tiles + tt) * kTile + row) * 8 : nullptr);
                    publish((P.save || l == 4) ? P.tact + ((size_t)l * P.trunk_tiles + tt) * kTileBytes : nullptr);
                    if (l < 4) signal_act();
                    dbg_stamp(P, first, 170 + l * 3 + 2);
                }
            } else if (MODE == HEAD_FWD) {
                const int base = head ? P_C1 : P_D1;
                const size_t ht = (size_t)b * P.tpf2 + t;
                for (int l = 0; l < 3; ++l) {
                    const float2 bv = bias_fetch(l == 0 ? P.biasfoot + ((size_t)head * P.B + b) * kW : P.bias[base + l]);
                    const bool first = leader && it == (int)blockIdx.x;
                    wait_acc();
                    dbg_stamp(P, first, 32 + (tl * 3 + l) * 4);
                    bias_put(bv);
                    before_overwrite();
                    dbg_stamp(P, first, 32 + (tl * 3 + l) * 4 + 1);
                    const size_t slot = ((size_t)head * 3 + l) * P.B * P.tpf2 + ht;
                    epi_relu(taddr, act_tile, row, bias_s, P.save ? P.hmask + (slot * kTile + row) * 8 : nullptr);
                    dbg_stamp(P, first, 32 + (tl * 3 + l) * 4 + 2);
                    publish(P.save ? P.hact + slot * kTileBytes : nullptr);
                    signal_act();
                    dbg_stamp(P, first, 32 + (tl * 3 + l) * 4 + 3);
                }
                wait_acc();
                if (leader) bulk_wait_read0();
                uint32_t v[32];
                tmem_ld32(taddr, v);
                tmem_ld_wait();
                tc_fence_before();
                mbar_arrive(&bar_act_free[tl]);      // act buffer and TMEM of this item are free for the next one
                if (valid) {
                    const float* bo = P.bias[base + 3];
                    const float t0v = tanhf(__uint_as_float(v[0]) + __ldg(bo + 0));
                    const float t1v = tanhf(__uint_as_float(v[1]) + __ldg(bo + 1));
                    const float t2v = tanhf(__uint_as_float(v[2]) + __ldg(bo + 2));
                    const size_t o = ((size_t)b * P.V + vrow) * 3;
                    if (head == 0) {
                        const float d0 = 0.1f * t0v, d1 = 0.1f * t1v, d2 = 0.1f * t2v;     // model.py:444
                        P.disp[o] = d0; P.disp[o + 1] = d1; P.disp[o + 2] = d2;
                        if (P.verts) {
                            const float* pp = P.pos + ((size_t)(P.shared ? 0 : b) * P.V + vrow) * 3;
                            float u0 = pp[0] + d0, u1 = pp[1] + d1, u2 = pp[2] + d2;
                            if (P.reg) {                                                   // model.py:481-488
                                const float* rg = P.reg + (size_t)b * 9;
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
            } else if (MODE == HEAD_BWD) {
                const size_t ht = (size_t)b * P.tpf2 + t;
                // prologue: scaled d(pre-tanh) -> columns 0..2 of image 0 (K = 16 GEMM) and the dW operand image
                before_overwrite();
                {
                    const float sc = grad_scale_from(*P.maxabs);
                    float d0 = 0.f, d1 = 0.f, d2 = 0.f;
                    if (valid) {
                        const float4 dv = *reinterpret_cast<const float4*>(P.dovals + (((size_t)head * P.B + b) * P.V + vrow) * 4);
                        d0 = dv.x * sc; d1 = dv.y * sc; d2 = dv.z * sc;
                    }
                    uint4 c0 = make_uint4(pack_h2(d0, d1), pack_h2(d2, 0.f), 0, 0), c1 = make_uint4(0, 0, 0, 0);
                    const uint32_t o0 = (uint32_t)row * 128u + (uint32_t)((0 ^ (row & 7)) << 4);
                    const uint32_t o1 = (uint32_t)row * 128u + (uint32_t)((1 ^ (row & 7)) << 4);
                    *reinterpret_cast<uint4*>(act_tile + o0) = c0;
                    *reinterpret_cast<uint4*>(act_tile + o1) = c1;
                    uint8_t* gi = P.doimg + ((size_t)head * P.B * P.tpf2 + ht) * kImgBytes;
                    *reinterpret_cast<uint4*>(gi + o0) = c0;
                    *reinterpret_cast<uint4*>(gi + o1) = c1;
                    fence_proxy_async();
                    signal_act();
                }
                for (int l = 2; l >= 0; --l) {       // l indexes a_{l+1} / dZ_{l+1}
                    const size_t slot = ((size_t)head * 3 + l) * P.B * P.tpf2 + ht;
                    uint32_t m[8];
                    load_mask_row(P.hmask + (slot * kTile + row) * 8, m);
                    wait_acc();
                    before_overwrite();
                    epi_mask(taddr, act_tile, row, m);
                    publish(P.hdz + slot * kTileBytes);
                    if (l > 0) signal_act();
                }
            } else {   // TRUNK_BWD
                const size_t tt = (size_t)b * P.tpf2 + t;
                for (int l = 4; l >= 0; --l) {
                    const size_t slot = (size_t)l * P.trunk_tiles + tt;
                    uint32_t m[8];
                    load_mask_row(P.tmask + (slot * kTile + row) * 8, m);
                    wait_acc();
                    before_overwrite();
                    epi_mask(taddr, act_tile, row, m);
                    publish(P.tdz + slot * kTileBytes);
                    if (l > 0) signal_act();
                }
                if (leader) { bulk_wait_read0(); mbar_arrive(&bar_act_free[tl]); }   // buffer free for the next item's load
            }
        }
        if (leader) bulk_wait0();
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, NT * 256);
    if (threadIdx.x == 0) dbg_stamp(P, true, 200 + MODE * 2 + 1);
}

// ---- launch ---------------------------------------------------------------------------------------------
template <int MODE, int NT>
static int launch_chain_t(const ChainParams& P, int sms, cudaStream_t st) {
    const int smem = NT * kTileBytes + kRing * kWChunkBytes + NT * kW * 4 + (2 * kRing + 4 * NT + 1) * 8;
    static bool attr_set = false;      // one process drives one GPU (one rank per GPU)
    if (!attr_set) {
        FB_CUDA(cudaFuncSetAttribute(chain_kernel<MODE, NT>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    int grid = P.n_items < sms ? P.n_items : sms;
    if (grid <= 0) return 0;
    static const char* const kNames[4] = {"trunk_fwd", "head_fwd", "head_bwd", "trunk_bwd"};
    FB_MARK(kNames[MODE], st);
    FB_CUDA(fb::launch_pdl(chain_kernel<MODE, NT>, grid, 64 + 128 * NT, smem, st, P));
    FB_CUDA(cudaGetLastError());
    return 0;
}

int launch_chain(int mode, int nt, const ChainParams& P, int sms, cudaStream_t st) {
    switch (mode * 2 + (nt - 1)) {
        case 0: return launch_chain_t<TRUNK_FWD, 1>(P, sms, st);
        case 1: return launch_chain_t<TRUNK_FWD, 2>(P, sms, st);
        case 2: return launch_chain_t<HEAD_FWD, 1>(P, sms, st);
        case 3: return launch_chain_t<HEAD_FWD, 2>(P, sms, st);
        case 4: return launch_chain_t<HEAD_BWD, 1>(P, sms, st);
        case 5: return launch_chain_t<HEAD_BWD, 2>(P, sms, st);
        case 6: return launch_chain_t<TRUNK_BWD, 1>(P, sms, st);
        case 7: return launch_chain_t<TRUNK_BWD, 2>(P, sms, st);
    }
    set_error("launch_chain: bad mode %d nt %d", mode, nt);
    return FINDB200_ERR_ARG;
}

}  // namespace field
}  // namespace fb
