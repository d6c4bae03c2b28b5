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

}  // namespace fb

extern "C" int findb200_umma_selftest(int mode, const void* a_img, const void* b_img, float* d_out, void* stream) {
    using namespace fb;
    FB_REQUIRE(mode == 0 || mode == 1, FINDB200_ERR_ARG, "umma_selftest: mode must be 0 or 1");
    const int smem = 65536 + 131072 + 1024;
    FB_CUDA(cudaFuncSetAttribute(umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    umma_selftest_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(mode, (const __half*)a_img, (const __half*)b_img, d_out);
    FB_CUDA(cudaGetLastError());
    return 0;
}
