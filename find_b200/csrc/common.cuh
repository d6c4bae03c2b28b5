// find_b200/csrc/common.cuh — sm_100a PTX wrappers shared by the Chamfer and field kernels.
// mbarrier / bulk-copy (TMA engine, UBLKCP) / tcgen05 (UTCHMMA, LDTM) / packed fp32x2 math.
// Every blocking wait is bounded: a protocol bug traps (launch failure) instead of hanging the GPU.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <cstdio>

#define FB_DEVINL __device__ __forceinline__

namespace fb {

// ------------------------------------------------------------------------------------------------
// error plumbing (C ABI: 0 = ok, negative = error, message via findb200_last_error)
void set_error(const char* fmt, ...);
int  check_cuda(cudaError_t e, const char* what);
#define FB_CUDA(call)                                                      \
    do {                                                                   \
        int _rc = fb::check_cuda((call), #call);                           \
        if (_rc) return _rc;                                               \
    } while (0)
#define FB_REQUIRE(cond, code, ...)                                        \
    do {                                                                   \
        if (!(cond)) { fb::set_error(__VA_ARGS__); return (code); }        \
    } while (0)

// Launch profiler (off by default): FB_MARK("kernel") before a launch records a CUDA event on the launching stream;
// the time to the next mark is attributed to that name.  FB_MARK_END closes a C-ABI call.  See capi.cu.
void prof_mark(const char* name, cudaStream_t st);
#define FB_MARK(name, st) fb::prof_mark((name), (st))
#define FB_MARK_END(st) fb::prof_mark(nullptr, (st))

// ------------------------------------------------------------------------------------------------
// Programmatic dependent launch: every kernel of the step starts with pdl_wait_trigger() and is launched through
// launch_pdl(), so kernel N+1 is scheduled (launch latency, CTA ramp) while kernel N drains.  griddepcontrol.wait blocks
// until the previous grid has COMPLETED and its memory is visible, so data hazards are exactly those of plain stream
// order; launch_dependents comes after the wait, which keeps at most two kernels of the chain resident.  The tensor-core kernels
// (chain2_kernel, dw_kernel) call it AFTER their set-up (mbarrier init, TMEM allocation, cluster sync: shared memory and TMEM only),
// so that part overlaps the previous kernel's tail wherever an SM is already free: 0.4518 -> 0.4461 ms per cfg2 step.
FB_DEVINL void pdl_wait_trigger() {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
bool pdl_enabled();     // capi.cu: FINDB200_PDL=0 turns the launch attribute off (A/B measurement)
template <typename... KP, typename... A>
inline cudaError_t launch_pdl(void (*kernel)(KP...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KP>(args)...);
}

// ------------------------------------------------------------------------------------------------
FB_DEVINL uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

FB_DEVINL bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}

// ---- mbarrier ------------------------------------------------------------------------------------
FB_DEVINL void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
FB_DEVINL void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
FB_DEVINL void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
FB_DEVINL void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
FB_DEVINL bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
#ifndef FB_WAIT_TIMEOUT_CYCLES
#define FB_WAIT_TIMEOUT_CYCLES (4000000000ll)   // ~2 s at 1.9 GHz: far beyond any legitimate wait
#endif
FB_DEVINL void mbar_wait(uint64_t* bar, uint32_t parity) {
    if (mbar_try_wait(bar, parity)) return;
    long long t0 = clock64();
    while (!mbar_try_wait(bar, parity)) {
        if (clock64() - t0 > FB_WAIT_TIMEOUT_CYCLES) {
            printf("find_b200: mbarrier wait timeout block %d thread %d bar@%u parity %u\n", (int)blockIdx.x,
                   (int)threadIdx.x, smem_u32(bar), parity);
            __trap();
        }
    }
}

// ---- proxies / fences ------------------------------------------------------------------------------
FB_DEVINL void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
FB_DEVINL void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
FB_DEVINL void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- bulk async copies (1-D TMA: no tensor map needed because tiles are stored as smem images) --------
FB_DEVINL void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
FB_DEVINL void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
                 "r"(smem_u32(smem_src)), "r"(bytes)
                 : "memory");
}
FB_DEVINL void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
FB_DEVINL void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
FB_DEVINL void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// ---- TMEM -----------------------------------------------------------------------------------------
FB_DEVINL void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {   // whole warp; ncols power of 2 >= 32
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
FB_DEVINL void tmem_dealloc(uint32_t taddr, uint32_t ncols) {     // same warp that allocated
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread t of the warp receives lane (lane_base + t).
FB_DEVINL void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
        "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
FB_DEVINL void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// ---- UMMA descriptors -------------------------------------------------------------------------------
// Shared-memory matrix descriptor, SWIZZLE_128B canonical layouts (8 rows x 128 B atoms, 16-B chunks
// XOR-swizzled by row&7; tile bases 1024-B aligned).  Fields: [0,14) addr>>4, [16,30) LBO>>4,
// [32,46) SBO>>4, [46,48) version=1 (sm_100), [61,64) layout (2 = SWIZZLE_128B).
//   K-major  : rows = M/N index, 128 B = 64 fp16 along K; SBO = stride between 8-row groups; LBO unused.
//   MN-major : rows = K index,   128 B = 64 fp16 along M/N; SBO = stride between 8-K-row groups,
//              LBO = stride between 64-element M/N atoms.
FB_DEVINL uint64_t umma_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
// Instruction descriptor for kind::f16 with fp16 A/B (format 0), fp32 accumulate (c_format 1).
FB_DEVINL constexpr uint32_t umma_idesc_f16(int M, int N, bool a_mn_major, bool b_mn_major) {
    return (1u << 4) | (0u << 7) | (0u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread.
FB_DEVINL void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// mbarrier arrive when all previously issued MMAs of this thread complete (implies fence::before_thread_sync)
FB_DEVINL void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

// ---- thread-block clusters / CTA pairs (tcgen05 cta_group::2) ---------------------------------------------
// A CTA pair = the two CTAs of a cluster of 2 (same TPC).  One tcgen05.mma.cta_group::2 issued by the leader (rank 0)
// computes M = 256 rows: rows 0..127 from the leader's A tile into the leader's TMEM, rows 128..255 from the peer's;
// each CTA stages HALF of the B operand (N/2 rows of the K-major weight image), so a weight byte is fetched from L2 once
// per 256 rows instead of once per 128.
FB_DEVINL uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
FB_DEVINL void cluster_sync_all() {      // every thread of every CTA of the cluster
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local_smem_addr` (a shared::cta address) in CTA `rank` of the cluster
FB_DEVINL uint32_t mapa_u32(uint32_t local_smem_addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local_smem_addr), "r"(rank));
    return r;
}
// arrive on an mbarrier anywhere in the cluster (address from mapa_u32).  Default semantics (.release at CTA scope), as
// CUTLASS's ClusterBarrier::arrive(cta_id) does: the explicit .release.cluster form makes ptxas emit MEMBAR.ALL.GPU before every
// arrive and the .acquire.cluster wait a CCTL.IVALL (L1 invalidate) after every wait — measured: the pair kernels ran 25 %
// SLOWER than the single-CTA ones with them.  What travels through these barriers is shared-memory data that is already in
// place (bulk-copy completions, or st.shared + fence.proxy.async + bar.sync by the writers), consumed by the async proxy.
FB_DEVINL void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
FB_DEVINL void mbar_wait_cluster(uint64_t* bar, uint32_t parity) { mbar_wait(bar, parity); }
FB_DEVINL void tmem_alloc2(uint32_t* smem_dst, uint32_t ncols) {   // one warp in EACH CTA of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
FB_DEVINL void tmem_dealloc2(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// D[tmem of both CTAs] (+)= A[smem of each CTA: its 128 rows] * B[smem: each CTA holds N/2 rows]; leader thread only
FB_DEVINL void umma2_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// arrive (when all MMAs issued so far by this thread have completed) on the mbarrier at the SAME offset in every CTA of
// `cta_mask`
FB_DEVINL void umma2_commit_mc(uint64_t* bar, uint16_t cta_mask) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"(cta_mask)
                 : "memory");
}

// ---- packed fp32x2 (FADD2 / FMUL2 / FFMA2), IEEE round-to-nearest, no flush-to-zero ----------------------
FB_DEVINL uint64_t pack2(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
FB_DEVINL void unpack2(uint64_t v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
FB_DEVINL uint64_t sub2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
FB_DEVINL uint64_t add2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
FB_DEVINL uint64_t mul2(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
FB_DEVINL uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
FB_DEVINL float min3(float a, float b, float c) {
    float r;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(r) : "f"(a), "f"(b), "f"(c));
    return r;
}

// ---- swizzled tile images ------------------------------------------------------------------------------
// An "image" is a [rows x 64] fp16 block stored exactly as UMMA SWIZZLE_128B wants it in shared memory:
// row r at byte r*128, its eight 16-B chunks permuted by (chunk ^ (r & 7)).  Images live unchanged in
// global memory too, so one bulk copy moves them either way.
FB_DEVINL uint32_t img_off(uint32_t row, uint32_t chunk) { return row * 128u + ((chunk ^ (row & 7u)) << 4); }

}  // namespace fb
