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
