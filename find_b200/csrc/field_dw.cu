// find_b200/csrc/field_dw.cu — weight-gradient GEMMs of the displacement field on tcgen05.
//
// dW_l[n][k] = sum over rows r of dZ_l[r][n] * a_{l-1}[r][k]   (autograd of the nn.Linear layers, model.py:250-257,
// 351-371).  Both operands are the fp16 tile images the chain kernels left in HBM; read as MN-major UMMA operands
// (contraction over the 128 rows of a tile), so no transposed copy is ever made.  One CTA owns one (GEMM, row-slice)
// pair and keeps the whole [256 x 256] fp32 partial sum in TMEM (2 x 256 columns) while it streams its tiles through a
// 3 x 64 KB shared-memory ring; four otherwise idle warps add up the columns of dZ from the same stages (bias
// gradients, and the per-foot sums the latent-code gradients are made of).  Partials go to HBM and are folded
// deterministically by reduce_dw_kernel (field.cu).
#include "field.cuh"
#include "field_dw.cuh"

namespace fb {
namespace field {

constexpr int kDwStages = 3;
constexpr int kDwStageBytes = 65536;   // A half-tile (<= 4 x 8 KB) | B half-tile (<= 4 x 8 KB)
constexpr int kDwThreads = 192;        // warp 0 producer, warp 1 MMA, warps 2..5 column sums + final epilogue

__global__ void __launch_bounds__(kDwThreads, 1) dw_kernel(const __grid_constant__ DwParams P) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bar_full[kDwStages], bar_empty[kDwStages], bar_done, bar_tmem_free;
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, lane = tid & 31;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);    // warp-uniform: the MMA issue path stays in uniform registers
    if (tid == 0) {
        for (int s = 0; s < kDwStages; ++s) { mbar_init(&bar_full[s], 1); mbar_init(&bar_empty[s], 1 + 128); }
        mbar_init(&bar_done, 1);
        mbar_init(&bar_tmem_free, 128);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base_s, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    fb::pdl_wait_trigger();            // the set-up above needs nothing from the previous kernel of the stream

    // this CTA's segments
    const int cta = (int)blockIdx.x + P.cta_base;          // the early / late launches of one plan share the CTA numbering
    const int s0 = P.cta_seg[cta], s1 = P.cta_seg[cta + 1];
    if (P.dbg && tid == 0 && cta < 240) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); P.dbg[2 * cta] = t; }

    if (warp == 0) {
        // ---- producer: 64-row halves of the A and B tiles, one 8 KB bulk copy per image -----------------------
        uint32_t cnt = 0;
        for (int si = s0; si < s1; ++si) {
            const DwSeg sg = P.seg[si];
            const DwGemm& G = P.g[sg.gemm];
            for (int u = sg.u0; u < sg.u1; ++u) {
                const int b = u / G.tpf, t = u % G.tpf;
                const uint8_t* At = G.A + ((size_t)b * G.a_foot_stride + t) * G.a_tile_bytes;
                const uint8_t* Bt = G.Bm + ((size_t)b * G.b_foot_stride + t) * G.b_tile_bytes + G.b_img_off;
                for (int hh = 0; hh < 2; ++hh, ++cnt) {
                    const uint32_t s = cnt % kDwStages, use = cnt / kDwStages;
                    uint8_t* st = ring + s * kDwStageBytes;
                    if (lane == 0) {
                        if (use > 0) mbar_wait(&bar_empty[s], (use - 1) & 1);
                        mbar_expect_tx(&bar_full[s], (uint32_t)(G.a_imgs + G.b_imgs) * 8192u);
                    }
                    __syncwarp();
                    if (lane < G.a_imgs) bulk_g2s(st + lane * 8192, At + (size_t)lane * kImgBytes + hh * 8192, 8192u, &bar_full[s]);
                    else if (lane >= 8 && lane < 8 + G.b_imgs)
                        bulk_g2s(st + 32768 + (lane - 8) * 8192, Bt + (size_t)(lane - 8) * kImgBytes + hh * 8192, 8192u, &bar_full[s]);
                    __syncwarp();
                }
            }
        }
    } else if (warp == 1) {
        // ---- MMA issuer ------------------------------------------------------------------------------------------
        if (elect_one()) {
            uint32_t cnt = 0;
            for (int si = s0; si < s1; ++si) {
                const DwSeg sg = P.seg[si];
                const DwGemm& G = P.g[sg.gemm];
                const uint32_t idesc = umma_idesc_f16(128, G.N, true, true);
                const uint32_t a_lbo = G.a_imgs > 1 ? 8192u : 0u;
                const int n_st = (sg.u1 - sg.u0) * 2;
                if (si > s0) { mbar_wait(&bar_tmem_free, (uint32_t)(si - s0 - 1) & 1u); tc_fence_after(); }   // the previous partial has left TMEM
                for (int i = 0; i < n_st; ++i, ++cnt) {
                    const uint32_t s = cnt % kDwStages, use = cnt / kDwStages;
                    mbar_wait(&bar_full[s], use & 1);
                    tc_fence_after();
                    const uint32_t sa = smem_u32(ring + s * kDwStageBytes), sb = sa + 32768;
                    for (int mh = 0; mh < G.mhalves; ++mh)
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t ad = umma_desc(sa + mh * 2 * 8192 + ks * 2048, a_lbo, 1024);
                            const uint64_t bd = umma_desc(sb + ks * 2048, 8192, 1024);
                            umma_f16(tmem + mh * 256, ad, bd, idesc, (i > 0 || ks > 0) ? 1u : 0u);
                        }
                    umma_commit(&bar_empty[s]);
                }
                umma_commit(&bar_done);
            }
        }
    } else {
        // ---- column sums of the A operand (dZ), then the TMEM -> HBM epilogue, per segment ------------------------
        const int ct = (warp - 2) * 32 + lane;           // column pair 2ct, 2ct+1
        const int img = ct >> 5, w = ct & 31;
        const int q = warp & 3, row = q * 32 + lane;
        uint32_t cnt = 0;
        for (int si = s0; si < s1; ++si) {
            const DwSeg sg = P.seg[si];
            const DwGemm& G = P.g[sg.gemm];
            const bool active = G.colsum_mode != 0 && (2 * ct < G.colsum_cols) && img < G.a_imgs;
            float c0 = 0.f, c1 = 0.f;
            int cur_foot = -1;
            auto flush = [&](int foot) {
                if (active && foot >= 0) {
                    long long* dst = G.colsum + (G.colsum_mode == 2 ? (size_t)foot * kW : 0) + 2 * ct;
                    fx_add(dst, c0, kColsumFx);
                    fx_add(dst + 1, c1, kColsumFx);
                }
                c0 = c1 = 0.f;
            };
            for (int u = sg.u0; u < sg.u1; ++u) {
                const int b = u / G.tpf;
                if (G.colsum_mode == 2 && b != cur_foot) { flush(cur_foot); }
                cur_foot = b;
                for (int hh = 0; hh < 2; ++hh, ++cnt) {
                    const uint32_t s = cnt % kDwStages, use = cnt / kDwStages;
                    mbar_wait(&bar_full[s], use & 1);
                    if (active) {
                        const uint8_t* base = ring + s * kDwStageBytes + img * 8192 + (w & 3) * 4;
#pragma unroll 8
                        for (int r = 0; r < 64; ++r) {
                            const uint32_t x = *reinterpret_cast<const uint32_t*>(base + r * 128 + (((w >> 2) ^ (r & 7)) << 4));
                            const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&x));
                            c0 += f.x;
                            c1 += f.y;
                        }
                    }
                    mbar_arrive(&bar_empty[s]);
                }
            }
            flush(cur_foot);

            mbar_wait(&bar_done, (uint32_t)(si - s0) & 1u);
            tc_fence_after();
            float* out = G.partial + (size_t)sg.slot * kW * kW;
            for (int mh = 0; mh < G.mhalves; ++mh)
                for (int cc = 0; cc < G.N / 32; ++cc) {
                    uint32_t v[32];
                    tmem_ld32(tmem + ((uint32_t)(q * 32) << 16) + mh * 256 + cc * 32, v);
                    tmem_ld_wait();
                    float4* o = reinterpret_cast<float4*>(out + (size_t)(mh * 128 + row) * kW + cc * 32);
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        o[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                           __uint_as_float(v[4 * j + 3]));
                }
            tc_fence_before();
            mbar_arrive(&bar_tmem_free);                 // the accumulators may be overwritten by the next segment
        }
    }
    tc_fence_before();
    __syncthreads();
    if (P.dbg && tid == 0 && cta < 240) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); P.dbg[2 * cta + 1] = t; }
    if (warp == 0) tmem_dealloc(tmem, 512);
}

int launch_dw(const DwParams& P, int n_ctas, cudaStream_t st, bool tpc_pairs) {
    if (n_ctas <= 0) return 0;
    const int smem = kDwStages * kDwStageBytes + 1024;
    static bool attr_set = false;
    if (!attr_set) {
        FB_CUDA(cudaFuncSetAttribute(dw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set = true;
    }
    // launch profiler: the early launch (side stream) and the late one are named apart; the dW plan's span runs from the fork
    // (the sum_feet mark of the main stream) to the end of the late launch, see bench.py
    FB_MARK(tpc_pairs ? "dw_gemms_early" : (P.cta_base > 0 ? "dw_gemms_late" : "dw_gemms"), st);
    if (tpc_pairs) {
        // the early launch runs beside the trunk's CTA-pair kernel: launched as 2-CTA clusters its CTAs fill whole TPCs and leave
        // whole TPCs free for the pairs (the kernel itself makes no use of the cluster)
        FB_REQUIRE(n_ctas % 2 == 0, FINDB200_ERR_UNSUPPORTED, "launch_dw: a TPC-paired launch needs an even number of CTAs");
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(n_ctas); cfg.blockDim = dim3(kDwThreads); cfg.dynamicSmemBytes = smem; cfg.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        FB_CUDA(cudaLaunchKernelEx(&cfg, dw_kernel, P));
    } else {
        FB_CUDA(fb::launch_pdl(dw_kernel, n_ctas, kDwThreads, smem, st, P));
    }
    FB_CUDA(cudaGetLastError());
    return 0;
}

}  // namespace field
}  // namespace fb
