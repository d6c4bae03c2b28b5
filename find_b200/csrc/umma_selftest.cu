// find_b200/csrc/umma_selftest.cu — one-CTA tcgen05 self-test of the conventions the field kernels use.
//
//   mode 0 (forward / dX shape):  D[128 x 256] = A[128 x 256] * B[256 x 256]^T
//        A : 4 column-block images [128 rows x 64 k]           (K-major, SWIZZLE_128B)
//        B : 8 chunk images, index nh*4+kc, [128 n x 64 k]     (K-major, SWIZZLE_128B)
//        two accumulators of 128 columns (N halves), 16 MMAs of M=128,N=128,K=16 each
//   mode 1 (dW shape):            D[128 x 256] = sum_r  G[r, m] * H[r, n],  r = 0..127
//        G : 2 column-block images [128 rows x 64 m]  read as MN-major A (M=128, K=rows)
//        H : 4 column-block images [128 rows x 64 n]  read as MN-major B (N=256, K=rows)
//        one accumulator of 256 columns, 8 MMAs of M=128,N=256,K=16
// Output D is fp32 [128][256] row-major (TMEM lane = row, column = column).
#include "common.cuh"
#include "../../include/find_b200.h"

namespace fb {

__global__ void __launch_bounds__(128) umma_selftest_kernel(int mode, const __half* __restrict__ a_img,
                                                            const __half* __restrict__ b_img, float* __restrict__ d_out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                 // up to 64 KB
    uint8_t* sB = smem + 65536;         // up to 128 KB
    __shared__ __align__(8) uint64_t bar_load, bar_mma;
    __shared__ uint32_t tmem_base_s;

    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t a_bytes = mode == 0 ? 65536u : 32768u;
    const uint32_t b_bytes = mode == 0 ? 131072u : 65536u;

    if (tid == 0) {
        mbar_init(&bar_load, 1);
        mbar_init(&bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc(&tmem_base_s, 256);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;

    if (tid == 0) {
        mbar_expect_tx(&bar_load, a_bytes + b_bytes);
        bulk_g2s(sA, a_img, a_bytes, &bar_load);
        for (uint32_t off = 0; off < b_bytes; off += 65536u) bulk_g2s(sB + off, (const uint8_t*)b_img + off, 65536u, &bar_load);
        mbar_wait(&bar_load, 0);
        tc_fence_after();
        if (mode == 0) {
            const uint32_t idesc = umma_idesc_f16(128, 128, false, false);
            for (int nh = 0; nh < 2; ++nh)
                for (int kc = 0; kc < 4; ++kc)
                    for (int k16 = 0; k16 < 4; ++k16) {
                        const uint64_t ad = umma_desc(smem_u32(sA) + kc * 16384 + k16 * 32, 16, 1024);
                        const uint64_t bd = umma_desc(smem_u32(sB) + (nh * 4 + kc) * 16384 + k16 * 32, 16, 1024);
                        umma_f16(tmem + nh * 128, ad, bd, idesc, (kc | k16) != 0);
                    }
        } else {
            const uint32_t idesc = umma_idesc_f16(128, 256, true, true);
            for (int ks = 0; ks < 8; ++ks) {      // 16 rows (K) per MMA = two 8-row groups = 2048 B
                const uint64_t ad = umma_desc(smem_u32(sA) + ks * 2048, 16384, 1024);
                const uint64_t bd = umma_desc(smem_u32(sB) + ks * 2048, 16384, 1024);
                umma_f16(tmem, ad, bd, idesc, ks != 0);
            }
        }
        umma_commit(&bar_mma);
    }
    __syncwarp();
    mbar_wait(&bar_mma, 0);
    tc_fence_after();

    // warp w may only touch TMEM lanes [32*(w%4), +32)
    const int row = warp * 32 + (tid & 31);
    for (int g = 0; g < 8; ++g) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + g * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) d_out[row * 256 + g * 32 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, 256);
}


// ---- CTA-pair variant (the conventions of field_chain2.cu) -----------------------------------------------------------
//   mode 2 / 3:  D[256 x 256] = A[256 x 256] * W[256 x 256]^T with ONE tcgen05.mma.cta_group::2 stream (M = 256, N = 256):
//        A : 2 tiles (CTA r owns rows 128r..128r+127) x 4 images [128 x 64 k]
//        W : 4 chunk images [256 n x 64 k]; CTA r stages rows n = 128r..128r+127 of every chunk (16 KB halves)
//   mode 2: the peer forwards "my half has landed" to the leader with a remote mbarrier arrive (relay);
//   mode 3: the peer's bulk copies complete_tx DIRECTLY on the leader's mbarrier (mapa address) — probes whether the
//           plain cp.async.bulk accepts an mbarrier of another CTA.
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128)
umma_selftest2_kernel(int mode, const uint8_t* __restrict__ a_img, const uint8_t* __restrict__ b_img, float* __restrict__ d_out) {
    extern __shared__ uint8_t smem_raw[];
    // identical carve-up in both CTAs (same dynamic-smem base offset), so "same offset" addressing works
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                 // 64 KB: this CTA's 128 rows
    uint8_t* sB = smem + 65536;         // 64 KB: 4 x 16 KB halves
    __shared__ __align__(8) uint64_t bar_load, bar_peer, bar_mma;
    __shared__ uint32_t tmem_base_s;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t rank = cluster_ctarank();
    if (tid == 0) {
        mbar_init(&bar_load, 1);
        mbar_init(&bar_peer, 1);
        mbar_init(&bar_mma, 1);
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc2(&tmem_base_s, 256);
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const uint32_t own_bytes = 65536u + 4u * 16384u;

    if (tid == 0) {
        if (mode == 2 || rank == 0) {
            // leader in mode 3 expects BOTH CTAs' bytes on its own barrier
            mbar_expect_tx(&bar_load, (mode == 3) ? 2u * own_bytes : own_bytes);
        }
        uint64_t* lbar = &bar_load;
        uint32_t bar_addr = smem_u32(lbar);
        if (mode == 3 && rank == 1) bar_addr = mapa_u32(bar_addr, 0);
        auto copy = [&](void* dst, const void* src, uint32_t bytes) {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(bar_addr) : "memory");
        };
        copy(sA, a_img + (size_t)rank * 65536, 65536u);
        for (int c = 0; c < 4; ++c) copy(sB + c * 16384, b_img + (size_t)c * 32768 + rank * 16384, 16384u);
        if (mode == 2 && rank == 1) {
            mbar_wait(&bar_load, 0);                                   // my half has landed ...
            mbar_arrive_cluster(mapa_u32(smem_u32(&bar_peer), 0));     // ... tell the leader
        }
        if (rank == 0) {
            mbar_wait_cluster(&bar_load, 0);
            if (mode == 2) mbar_wait_cluster(&bar_peer, 0);
            tc_fence_after();
            const uint32_t idesc = umma_idesc_f16(256, 256, false, false);
            for (int kc = 0; kc < 4; ++kc)
                for (int k16 = 0; k16 < 4; ++k16) {
                    const uint64_t ad = umma_desc(smem_u32(sA) + kc * 16384 + k16 * 32, 16, 1024);
                    const uint64_t bd = umma_desc(smem_u32(sB) + kc * 16384 + k16 * 32, 16, 1024);
                    umma2_f16(tmem, ad, bd, idesc, (kc | k16) != 0);
                }
            umma2_commit_mc(&bar_mma, 3);
        }
    }
    __syncwarp();
    mbar_wait_cluster(&bar_mma, 0);
    tc_fence_after();
    const int row = warp * 32 + (tid & 31);
    for (int g = 0; g < 8; ++g) {
        uint32_t v[32];
        tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + g * 32, v);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) d_out[(size_t)(rank * 128 + row) * 256 + g * 32 + j] = __uint_as_float(v[j]);
    }
    tc_fence_before();
    cluster_sync_all();
    if (warp == 0) tmem_dealloc2(tmem, 256);
}

// ---- tensor-pipe rate probe -------------------------------------------------------------------------------------------
// How many SM cycles does ONE K = 16 MMA of the shapes the chain kernels issue really take, with nothing else running on
// the SM?  Operands are whatever bytes sit in shared memory / TMEM (values are irrelevant for timing).
//   mode 0: cta_group::1, M = 128, N = 256, A and B from shared memory (SS)      — field_chain.cu
//   mode 1: cta_group::1, M = 128, N = 128, SS
//   mode 2: cta_group::2, M = 256, N = 256, SS (each CTA: 128 rows of A, 128 rows of B) — field_chain2.cu
//   mode 3: cta_group::1, M = 128, N = 256, A from TMEM (TS), B from shared memory
//   mode 4: cta_group::2, M = 256, N = 256, TS
// out[0] = cycles (clock64) from the first issue to the commit arriving, out[1] = MMAs issued.
template <int CG>
__global__ void __launch_bounds__(128) umma_rate_kernel(int mode, int iters, int flags, const uint8_t* __restrict__ gsrc,
                                                        unsigned long long* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* sA = smem;                 // 64 KB
    uint8_t* sB = smem + 65536;         // 128 KB (CG = 1) / 64 KB (CG = 2)
    // flags (what the chain kernels do around the MMAs): 1 = tcgen05.commit to a ring barrier after every 4 MMAs, 2 = tcgen05.fence
    // ::after_thread_sync before every 4, 4 = an (already satisfied) mbarrier wait before every 4, 8 = a second warp streams 16 KB
    // bulk copies from global memory into a spare shared-memory ring all the while, 16 = two warps keep writing shared memory
    __shared__ __align__(8) uint64_t bar_mma, bar_ring[8], bar_done_dummy, bar_bg[4];
    __shared__ uint32_t tmem_base_s;
    __shared__ volatile int stop_flag;
    const int tid = threadIdx.x;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);     // warp-uniform (uniform-register issue path, as in the chain kernels)
    const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
    for (int i = tid; i < (65536 + (CG == 2 ? 65536 : 131072)) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (tid == 0) {
        mbar_init(&bar_mma, 1); mbar_init(&bar_done_dummy, 1);
        for (int i = 0; i < 8; ++i) mbar_init(&bar_ring[i], 1);
        for (int i = 0; i < 4; ++i) mbar_init(&bar_bg[i], 1);
        stop_flag = 0;
        fence_barrier_init();
    }
    if (warp == 0) { if (CG == 2) tmem_alloc2(&tmem_base_s, 512); else tmem_alloc(&tmem_base_s, 512); }
    fence_proxy_async();
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    const bool ts = mode >= 3;
    const int N = mode == 1 ? 128 : 256;
    if (warp == 0 && rank == 0) {
      if (elect_one()) {
        const uint32_t idesc = umma_idesc_f16(CG == 2 ? 256 : 128, N, false, false);
        const long long t0 = clock64();
        unsigned long long n = 0;
        for (int it = 0; it < iters; ++it)
            for (int kc = 0; kc < 4; ++kc) {
                if (flags & 4) mbar_wait(&bar_done_dummy, 1);          // fresh barrier, parity 1: satisfied at once
                if (flags & 2) tc_fence_after();
                for (int k16 = 0; k16 < 4; ++k16, ++n) {
                    const uint64_t ad = umma_desc(smem_u32(sA) + kc * 16384 + k16 * 32, 16, 1024);
                    const uint64_t bd = umma_desc(smem_u32(sB) + kc * (CG == 2 ? 16384 : 32768) + k16 * 32, 16, 1024);
                    if (!ts) {
                        if (CG == 2) umma2_f16(tmem, ad, bd, idesc, 1u); else umma_f16(tmem, ad, bd, idesc, 1u);
                    } else {
                        const uint32_t a_t = tmem + 256 + (uint32_t)(kc * 4 + k16) * 8;     // [128 lanes x 16 fp16] = 8 columns
                        if (CG == 2)
                            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                         "tcgen05.mma.cta_group::2.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                                         ::"r"(tmem), "r"(a_t), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                        else
                            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                         "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                                         ::"r"(tmem), "r"(a_t), "l"(bd), "r"(idesc), "r"(1u) : "memory");
                    }
                }
                if (flags & 1) { if (CG == 2) umma2_commit_mc(&bar_ring[(it * 4 + kc) & 7], 3); else umma_commit(&bar_ring[(it * 4 + kc) & 7]); }
            }
        if (CG == 2) umma2_commit_mc(&bar_mma, 3); else umma_commit(&bar_mma);
        mbar_wait(&bar_mma, 0);
        out[0] = (unsigned long long)(clock64() - t0);
        out[1] = n;
        stop_flag = 1;
      }
    } else if (warp == 1 && (tid & 31) == 0 && (flags & 8) && gsrc && CG == 2) {
        // background weight stream: 16 KB bulk copies into the upper 64 KB of the B region (unused by the pair's MMAs)
        uint8_t* bg = sB + 65536;
        uint32_t i = 0;
        while (!stop_flag && i < (1u << 20)) {          // bounded: a protocol slip must not hang the GPU
            const uint32_t sl = i & 3, use = i >> 2;
            if (use > 0) mbar_wait(&bar_bg[sl], (use - 1) & 1);
            mbar_expect_tx(&bar_bg[sl], 16384u);
            bulk_g2s(bg + sl * 16384, gsrc + (size_t)((i * 2 + rank) & 63) * 16384, 16384u, &bar_bg[sl]);
            ++i;
        }
        for (uint32_t k = (i >= 4 ? i - 4 : 0); k < i; ++k) mbar_wait(&bar_bg[k & 3], (k >> 2) & 1);
    } else if (warp >= 2 && (flags & 16) && CG == 2) {
        uint8_t* bg = sB + 65536;
        uint32_t i = 0;
        while (!stop_flag && i < (1u << 24)) {
            *reinterpret_cast<uint4*>(bg + (((tid - 64) * 16 + i * 1024) & 65535)) = make_uint4(i, i, i, i);
            ++i;
        }
    }
    __syncwarp();
    if (CG == 2 && !(tid == 0 && rank == 0)) { if (tid == 0) { mbar_wait(&bar_mma, 0); stop_flag = 1; } }   // the peer's own flag
    tc_fence_before();
    if (CG == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 0) { if (CG == 2) tmem_dealloc2(tmem, 512); else tmem_dealloc(tmem, 512); }
}

// ---- synchronisation-primitive cost probe --------------------------------------------------------------------------------
// What does ONE elected thread of the leader CTA of a pair pay for each primitive of the chain kernels' issue path?  op:
//   0 tcgen05.commit cta_group::1 (nothing pending)        1 tcgen05.commit cta_group::2 multicast to both CTAs (nothing pending)
//   2 mbarrier.try_wait, phase already complete            3 ld.volatile.shared                  4 mbarrier.arrive (own CTA)
//   5 mbarrier.arrive on the PEER's barrier                6 fence.proxy.async                   7 tcgen05.fence::after_thread_sync
//   8 4 x MMA (cg2 M256 N256) + multicast commit + wait for that commit (round trip, pipe empty at the start)
//   9 4 x MMA (cg2) + multicast commit, no wait (issue-side cost with the pipe kept busy)
//  10 S2R of the cluster CTA id (what cvta.shared costs inside a cluster kernel)    11 clock64 read
// out[0] = cycles for `iters` repetitions, out[1] = iters.
__global__ void __launch_bounds__(64) sync_probe_kernel(int op, int iters, unsigned long long* __restrict__ out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bar[4];
    __shared__ uint32_t tmem_base_s;
    __shared__ volatile uint32_t word;
    const int tid = threadIdx.x;
    const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
    const uint32_t rank = cluster_ctarank();
    for (int i = tid; i < 131072 / 4; i += 64) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (tid == 0) {
        for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1);
        word = 1;
        fence_barrier_init();
    }
    if (warp == 0) tmem_alloc2(&tmem_base_s, 512);
    fence_proxy_async();
    tc_fence_before();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem = tmem_base_s;
    if (warp == 1 && rank == 0) {
        if (elect_one()) {
            const uint32_t idesc = umma_idesc_f16(256, 256, false, false);
            const uint32_t sa = smem_u32(smem), sb = sa + 65536;
            const uint32_t peer_bar = mapa_u32(smem_u32(&bar[1]), 1);
            uint32_t acc = 0;
            const long long t0 = clock64();
            for (int it = 0; it < iters; ++it) {
                switch (op) {
                    case 0: umma_commit(&bar[0]); break;
                    case 1: umma2_commit_mc(&bar[0], 3); break;
                    case 2: acc += mbar_try_wait(&bar[2], 1) ? 1u : 0u; break;          // fresh barrier: parity 1 is "complete"
                    case 3: acc += word; break;
                    case 4: mbar_arrive(&bar[3]); break;
                    case 5: mbar_arrive_cluster(peer_bar); break;
                    case 6: fence_proxy_async(); break;
                    case 7: tc_fence_after(); break;
                    case 8:
                    case 9:
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            umma2_f16(tmem, umma_desc(sa + (it & 3) * 16384 + k * 32, 16, 1024), umma_desc(sb + (it & 3) * 16384 + k * 32, 16, 1024),
                                      idesc, 1u);
                        umma2_commit_mc(&bar[0], 3);
                        if (op == 8) mbar_wait(&bar[0], it & 1);
                        break;
                    case 10: { uint32_t r; asm volatile("mov.u32 %0, %%cluster_ctaid.x;" : "=r"(r)); acc += r; } break;
                    default: acc += (uint32_t)clock64(); break;
                }
            }
            const long long t1 = clock64();
            if (op == 9 || op == 1 || op == 0) { umma2_commit_mc(&bar[1], 3); mbar_wait(&bar[1], 0); }
            out[0] = (unsigned long long)(t1 - t0);
            out[1] = (unsigned long long)iters + (acc == 0xdeadbeefu ? 1ull : 0ull);
        }
    }
    __syncwarp();
    tc_fence_before();
    cluster_sync_all();
    if (warp == 0) tmem_dealloc2(tmem, 512);
}

}  // namespace fb

extern "C" int findb200_umma_selftest(int mode, const void* a_img, const void* b_img, float* d_out, void* stream) {
    using namespace fb;
    FB_REQUIRE(mode >= 0 && mode <= 3, FINDB200_ERR_ARG, "umma_selftest: mode must be 0..3");
    if (mode >= 2) {       // CTA pair: d_out is [256][256]
        const int smem2 = 65536 + 65536 + 1024;
        FB_CUDA(cudaFuncSetAttribute(umma_selftest2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem2));
        umma_selftest2_kernel<<<2, 128, smem2, (cudaStream_t)stream>>>(mode, (const uint8_t*)a_img, (const uint8_t*)b_img, d_out);
        FB_CUDA(cudaGetLastError());
        return 0;
    }
    const int smem = 65536 + 131072 + 1024;
    FB_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    umma_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(mode, (const __half*)a_img, (const __half*)b_img, d_out);
    FB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int findb200_umma_rate(int mode, int iters, int flags, const void* gsrc_1mb, unsigned long long* out2_dev, void* stream) {
    using namespace fb;
    FB_REQUIRE(mode >= 0 && mode <= 4 && iters >= 1 && out2_dev, FINDB200_ERR_ARG, "umma_rate: mode 0..4, iters >= 1");
    const bool pair = (mode == 2 || mode == 4);
    const int smem = 65536 + 131072 + 1024;
    if (pair) {
        FB_CUDA(cudaFuncSetAttribute(umma_rate_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem; cfg.stream = (cudaStream_t)stream;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        FB_CUDA(cudaLaunchKernelEx(&cfg, umma_rate_kernel<2>, mode, iters, flags, (const uint8_t*)gsrc_1mb, out2_dev));
    } else {
        FB_CUDA(cudaFuncSetAttribute(umma_rate_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        umma_rate_kernel<1><<<1, 128, smem, (cudaStream_t)stream>>>(mode, iters, flags, (const uint8_t*)gsrc_1mb, out2_dev);
    }
    FB_CUDA(cudaGetLastError());
    return 0;
}

extern "C" int findb200_sync_probe(int op, int iters, unsigned long long* out2_dev, void* stream) {
    using namespace fb;
    FB_REQUIRE(op >= 0 && op <= 11 && iters >= 1 && out2_dev, FINDB200_ERR_ARG, "sync_probe: op 0..11, iters >= 1");
    const int smem = 131072 + 1024;
    FB_CUDA(cudaFuncSetAttribute(sync_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2); cfg.blockDim = dim3(64); cfg.dynamicSmemBytes = smem; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    FB_CUDA(cudaLaunchKernelEx(&cfg, sync_probe_kernel, op, iters, out2_dev));
    FB_CUDA(cudaGetLastError());
    return 0;
}
