// find_b200/csrc/allreduce.cu — the one exchange step of the foot-batch-sharded training step: the sum of
// the flat MLP-gradient bucket (868 358 floats = 3.47 MB) over the ranks, as ONE kernel over NVLink / NVSwitch peer memory.
//
// Every rank holds the bucket in symmetric memory (same size on every GPU, peer-mapped; the host side gets the mappings from
// torch's symmetric-memory rendezvous, find_b200/parallel.py).  Two-shot, in place:
//   barrier A   every rank's bucket is complete (its backward kernels precede this launch in stream order)
//   reduce      rank r owns slice r: sums that slice over all W buckets in FIXED rank order (every element is reduced by
//               exactly one rank, so all ranks end with bit-identical values), scales, and stores the result into slice r of
//               EVERY rank's bucket (the all-gather is done by the remote stores)
//   barrier B   all ranks have finished reading and writing: the bucket is the reduced one everywhere
// Barriers are per CTA: CTA b of rank r exchanges epoch flags with CTA b of every other rank (flag[b][src_rank] in each rank's
// symmetric flag page, st.release.sys / ld.acquire.sys), so no grid-wide sync is needed; a kernel completes only when all of its
// CTAs have passed barrier B.  The epoch lives on the device and is bumped by the last CTA out: the launch is CUDA-graph safe.
// NCCL needs ~50 us for this message inside the step graph (launch + protocol latency); this kernel moves 2 x 3.5 MB x (W-1)/W
// per rank at NVLink speed and costs two flag round trips.
#include "common.cuh"
#include "../../include/find_b200.h"
#include <cstdlib>
#include <cstring>

namespace fb {
namespace comm {

constexpr int kMaxRanks = 16;
constexpr int kMaxCtas = 160;

struct P2PParams {
    float* buf[kMaxRanks];          // peer-mapped bucket of every rank (buf[rank] is the local one)
    uint32_t* flags[kMaxRanks];     // peer-mapped flag page of every rank: [kMaxCtas][kMaxRanks] u32, zero-initialised
    int rank, world;
    long long n4;                   // bucket length in float4 units (padded by the host)
    float scale;
    uint32_t* epoch;                // device: [0] number of completed calls, [1] CTAs of this launch that are done,
                                    //         [2] error code (0 ok, 1 a rank never arrived, 2 ranks out of lock-step), [3] culprit rank
    long long timeout_cycles;       // bound of one flag wait (FINDB200_P2P_TIMEOUT_S, default 600 s: rank skew of a checkpoint
                                    // save or a validation pass is ordinary; NCCL would simply wait)
    float* mc;                      // multicast mapping of the bucket (NVLS variant), else null
};

FB_DEVINL void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
FB_DEVINL uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
FB_DEVINL float4 ld_peer(const float4* p) {       // peer data: never serve it from a stale line
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// CTA b of this rank <-> CTA b of every rank.  Bounded: a rank that never arrives within `timeout_cycles` makes the kernel give
// up WITHOUT trapping — it records an error code in epoch[2] (host-visible through P2PAllReduce.check()) and every later wait of
// this and of following launches returns at once, so the CUDA context survives and the job can fall back to NCCL or stop cleanly.
// Ranks must enqueue the call in lock-step: barrier A also verifies that no peer is a whole call ahead (flag > epoch + 1 means
// that rank enqueued one call more than this one) and reports it as error 2 instead of silently pairing different steps.
// kPublish: this CTA's threads have written peer memory in this kernel and those writes must be visible system-wide before the
// flag (barrier B).  Barrier A publishes nothing of this kernel: the bucket was written by EARLIER kernels into local memory,
// which a peer reads through this GPU's L2 — the kernel boundary already ordered those writes — so it skips the system fences
// (each costs an NVLink round trip while remote traffic is outstanding; dropping the 256 per-thread fences of barrier A took
// 9.5 us off the N = 2 step).
template <bool kPublish>
FB_DEVINL void cross_rank_barrier(const P2PParams& p, uint32_t epoch) {
    if (kPublish) __syncthreads();   // every thread's remote stores happen-before the flag threads' release below: st.release.sys is
                                     // cumulative over them (the pattern of a grid barrier: bar.sync, then ONE thread fences + signals)
    const int t = threadIdx.x;
    if (t < p.world) {
        st_release_sys(p.flags[t] + blockIdx.x * kMaxRanks + p.rank, epoch);
        const uint32_t* mine = p.flags[p.rank] + blockIdx.x * kMaxRanks + t;
        const long long t0 = clock64();
        uint32_t seen;
        while ((int32_t)((seen = ld_acquire_sys(mine)) - epoch) < 0) {
            if (*(volatile uint32_t*)(p.epoch + 2) != 0u) break;                  // an earlier wait already failed
            if (clock64() - t0 > p.timeout_cycles) {
                if (atomicCAS(p.epoch + 2, 0u, 1u) == 0u) {
                    p.epoch[3] = (uint32_t)t;
                    printf("find_b200: allreduce_p2p rank %d cta %d: rank %d never reached epoch %u (gave up, error flag set)\n", p.rank,
                           (int)blockIdx.x, t, epoch);
                }
                break;
            }
        }
        if (!kPublish && (int32_t)(seen - epoch) >= 2 && atomicCAS(p.epoch + 2, 0u, 2u) == 0u) {
            p.epoch[3] = (uint32_t)t;
            printf("find_b200: allreduce_p2p rank %d: rank %d is a call ahead (flag %u, expected %u): ranks must enqueue in lock-step\n",
                   p.rank, t, seen, epoch);
        }
    }
    __syncthreads();
}

// NVLS variant: the bucket is also mapped through a multicast object.  multimem.ld_reduce returns the SUM of the addressed 16
// bytes over every rank's copy, added inside the NVSwitch (one request per rank slice instead of W - 1 peer loads), and
// multimem.st writes the scaled result into every rank's copy at once.  Same two barriers, same slice ownership: every element
// is reduced by exactly one rank and broadcast, so all ranks still end with bit-identical buckets.
FB_DEVINL float4 mc_ld_reduce(const float4* p) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
FB_DEVINL void mc_st(float4* p, const float4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

__global__ void __launch_bounds__(256) allreduce_mc_kernel(const __grid_constant__ P2PParams p) {
    const uint32_t e = *p.epoch;
    cross_rank_barrier<false>(p, 2u * e + 1u);
    const long long per = (p.n4 + p.world - 1) / p.world;
    const long long lo = per * p.rank, hi = min(lo + per, p.n4);
    constexpr int kUnroll = 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    float4* mc = reinterpret_cast<float4*>(p.mc);
    for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * kUnroll) {
        float4 v[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i < hi) v[u] = mc_ld_reduce(mc + i);
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i >= hi) continue;
            float4 a = v[u];
            a.x *= p.scale; a.y *= p.scale; a.z *= p.scale; a.w *= p.scale;
            mc_st(mc + i, a);
        }
    }
    cross_rank_barrier<true>(p, 2u * e + 2u);
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(p.epoch + 1, 1u) == gridDim.x - 1) { p.epoch[1] = 0u; p.epoch[0] = e + 1u; }
    }
}

__global__ void __launch_bounds__(256) allreduce_p2p_kernel(const __grid_constant__ P2PParams p) {
    const uint32_t e = *p.epoch;
    cross_rank_barrier<false>(p, 2u * e + 1u);
    // slice of this rank, in float4 units
    const long long per = (p.n4 + p.world - 1) / p.world;
    const long long lo = per * p.rank, hi = min(lo + per, p.n4);
    // kUnroll x world independent 16-byte peer loads are in flight per thread before the first add: the loop is bound by
    // NVLink round trips (~2 us), not by bytes
    constexpr int kUnroll = 4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * kUnroll) {
        float4 v[kUnroll][kMaxRanks > 8 ? 8 : kMaxRanks];
        float4 acc[kUnroll];
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int q0 = 0; q0 < p.world; q0 += 8) {
            const int qn = min(8, p.world - q0);
#pragma unroll
            for (int u = 0; u < kUnroll; ++u) {
                const long long i = i0 + u * stride;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q < qn && i < hi) v[u][q] = ld_peer(reinterpret_cast<const float4*>(p.buf[q0 + q]) + i);
            }
#pragma unroll
            for (int u = 0; u < kUnroll; ++u)
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (q < qn) { acc[u].x += v[u][q].x; acc[u].y += v[u][q].y; acc[u].z += v[u][q].z; acc[u].w += v[u][q].w; }
        }
#pragma unroll
        for (int u = 0; u < kUnroll; ++u) {
            const long long i = i0 + u * stride;
            if (i >= hi) continue;
            float4 a = acc[u];
            a.x *= p.scale; a.y *= p.scale; a.z *= p.scale; a.w *= p.scale;
            for (int q = 0; q < p.world; ++q) reinterpret_cast<float4*>(p.buf[q])[i] = a;
        }
    }
    cross_rank_barrier<true>(p, 2u * e + 2u);
    // the last CTA out bumps the epoch (every CTA read it before its first barrier; the next launch starts after this one ends)
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(p.epoch + 1, 1u) == gridDim.x - 1) { p.epoch[1] = 0u; p.epoch[0] = e + 1u; }
    }
}

}  // namespace comm
}  // namespace fb

using namespace fb;
using namespace fb::comm;

static long long p2p_timeout_cycles() {
    static long long cyc = 0;
    if (!cyc) {
        const char* e = getenv("FINDB200_P2P_TIMEOUT_S");
        double s = e ? atof(e) : 600.0;
        if (!(s > 0.0)) s = 600.0;
        cyc = (long long)(s * 2.0e9);           // clock64 ticks at <= ~2 GHz: the bound is generous, not exact
    }
    return cyc;
}

extern "C" int findb200_allreduce_p2p(void* const* peer_bufs, void* const* peer_flags, int rank, int world, long long n_floats,
                                      float scale, unsigned int* epoch_dev, int n_ctas, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(peer_bufs && peer_flags && epoch_dev, FINDB200_ERR_ARG, "allreduce_p2p: null argument");
    FB_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, FINDB200_ERR_ARG, "allreduce_p2p: bad rank/world %d/%d", rank, world);
    FB_REQUIRE(n_floats >= 0 && (n_floats & 3) == 0, FINDB200_ERR_SHAPE, "allreduce_p2p: the bucket length must be a multiple of 4 floats (pad it)");
    if (n_ctas < 1) n_ctas = 128;
    if (n_ctas > kMaxCtas) n_ctas = kMaxCtas;
    P2PParams p;
    memset(&p, 0, sizeof(p));
    for (int q = 0; q < world; ++q) {
        FB_REQUIRE(peer_bufs[q] && peer_flags[q], FINDB200_ERR_ARG, "allreduce_p2p: peer %d has no mapping", q);
        FB_REQUIRE(((uintptr_t)peer_bufs[q] & 15) == 0, FINDB200_ERR_ARG, "allreduce_p2p: peer buffer %d is not 16-byte aligned", q);
        p.buf[q] = (float*)peer_bufs[q];
        p.flags[q] = (uint32_t*)peer_flags[q];
    }
    p.rank = rank; p.world = world; p.n4 = n_floats / 4; p.scale = scale; p.epoch = epoch_dev;
    p.timeout_cycles = p2p_timeout_cycles();
    FB_MARK("allreduce_p2p", st);
    allreduce_p2p_kernel<<<n_ctas, 256, 0, st>>>(p);
    FB_CUDA(cudaGetLastError());
    FB_MARK_END(st);
    return 0;
}

// Same contract as findb200_allreduce_p2p, with the reduction done inside the NVSwitch: mc_buf = the multicast mapping of the
// bucket (torch symmetric memory: handle.multicast_ptr); peer_flags as above (the barriers are the same peer-memory flags).
extern "C" int findb200_allreduce_mc(void* mc_buf, void* const* peer_flags, int rank, int world, long long n_floats, float scale,
                                     unsigned int* epoch_dev, int n_ctas, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    FB_REQUIRE(mc_buf && peer_flags && epoch_dev, FINDB200_ERR_ARG, "allreduce_mc: null argument");
    FB_REQUIRE(world >= 1 && world <= kMaxRanks && rank >= 0 && rank < world, FINDB200_ERR_ARG, "allreduce_mc: bad rank/world %d/%d", rank, world);
    FB_REQUIRE(n_floats >= 0 && (n_floats & 3) == 0, FINDB200_ERR_SHAPE, "allreduce_mc: the bucket length must be a multiple of 4 floats (pad it)");
    FB_REQUIRE(((uintptr_t)mc_buf & 15) == 0, FINDB200_ERR_ARG, "allreduce_mc: the multicast pointer is not 16-byte aligned");
    if (n_ctas < 1) n_ctas = 128;
    if (n_ctas > kMaxCtas) n_ctas = kMaxCtas;
    P2PParams p;
    memset(&p, 0, sizeof(p));
    for (int q = 0; q < world; ++q) {
        FB_REQUIRE(peer_flags[q], FINDB200_ERR_ARG, "allreduce_mc: peer %d has no flag mapping", q);
        p.flags[q] = (uint32_t*)peer_flags[q];
    }
    p.rank = rank; p.world = world; p.n4 = n_floats / 4; p.scale = scale; p.epoch = epoch_dev; p.mc = (float*)mc_buf;
    p.timeout_cycles = p2p_timeout_cycles();
    FB_MARK("allreduce_mc", st);
    allreduce_mc_kernel<<<n_ctas, 256, 0, st>>>(p);
    FB_CUDA(cudaGetLastError());
    FB_MARK_END(st);
    return 0;
}

extern "C" size_t findb200_allreduce_p2p_flag_bytes(void) { return (size_t)kMaxCtas * kMaxRanks * sizeof(uint32_t); }
