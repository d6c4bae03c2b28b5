// find_b200/csrc/capi.cu — error plumbing of the C ABI (include/find_b200.h).
#include <cstdlib>
#include "common.cuh"
#include "../../include/find_b200.h"
#include <cstdarg>
#include <cstdio>
#include <cstring>

namespace fb {

bool pdl_enabled() {
    static int on = -1;
    if (on < 0) { const char* e = getenv("FINDB200_PDL"); on = (e && e[0] == '0') ? 0 : 1; }
    return on != 0;
}


static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_cuda(cudaError_t e, const char* what) {
    if (e == cudaSuccess) return 0;
    set_error("CUDA error %s (%d) at %s", cudaGetErrorString(e), (int)e, what);
    return FINDB200_ERR_CUDA;
}

// ---- launch profiler ----------------------------------------------------------------------------------------
constexpr int kProfMax = 8192;
static bool g_prof_on = false;
static int g_prof_n = 0;
static cudaEvent_t g_prof_ev[kProfMax];
static const char* g_prof_name[kProfMax];
static cudaStream_t g_prof_stream[kProfMax];      // a mark's time runs to the next mark on the SAME stream
static int g_prof_created = 0;
static bool g_prof_keep = false;

void prof_mark(const char* name, cudaStream_t st) {
    if (!g_prof_on || g_prof_n >= kProfMax) return;
    if (g_prof_n >= g_prof_created) {
        if (cudaEventCreate(&g_prof_ev[g_prof_created]) != cudaSuccess) return;
        ++g_prof_created;
    }
    // inside a stream capture a plain record is only a dependency edge; the External flag makes it an event-record NODE that
    // every replay re-records and that cudaEventElapsedTime accepts
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) cs = cudaStreamCaptureStatusNone;
    if (cs == cudaStreamCaptureStatusActive) cudaEventRecordWithFlags(g_prof_ev[g_prof_n], st, cudaEventRecordExternal);
    else cudaEventRecord(g_prof_ev[g_prof_n], st);
    g_prof_name[g_prof_n] = name;
    g_prof_stream[g_prof_n] = st;
    ++g_prof_n;
}

// A side stream per device for the two places where a training step runs kernels beside its main stream (chamfer_hooks.cuh):
// created on first use; inside a stream capture the side stream joins the capture through the fork event and leaves it through
// the join event.
int side_stream(int which, cudaStream_t* s, cudaEvent_t* fork, cudaEvent_t* join) {
    constexpr int kMaxDev = 32;
    static cudaStream_t streams[kMaxDev];
    static cudaEvent_t forks[kMaxDev][2], joins[kMaxDev][2];
    static bool made[kMaxDev];
    int dev = 0;
    FB_CUDA(cudaGetDevice(&dev));
    FB_REQUIRE(dev >= 0 && dev < kMaxDev && which >= 0 && which < 2, FINDB200_ERR_UNSUPPORTED, "side_stream: device ordinal out of range");
    if (!made[dev]) {
        FB_CUDA(cudaStreamCreateWithFlags(&streams[dev], cudaStreamNonBlocking));
        for (int k = 0; k < 2; ++k) {
            FB_CUDA(cudaEventCreateWithFlags(&forks[dev][k], cudaEventDisableTiming));
            FB_CUDA(cudaEventCreateWithFlags(&joins[dev][k], cudaEventDisableTiming));
        }
        made[dev] = true;
    }
    *s = streams[dev]; *fork = forks[dev][which]; *join = joins[dev][which];
    return 0;
}

}  // namespace fb

extern "C" {
int findb200_profile_enable(int on) {
    // 1: start recording (drops earlier marks); 0: off; 2: stop recording but KEEP the marks — for marks that were captured
    // into a CUDA graph as event-record nodes: every replay re-records them and findb200_profile_read can be called after each
    if (on == 2) { fb::g_prof_on = false; fb::g_prof_keep = true; return 0; }
    fb::g_prof_on = on != 0;
    fb::g_prof_keep = false;
    if (on) fb::g_prof_n = 0;
    return 0;
}

// Synchronises on the recorded events and writes one "name,launches,total_ms" line per kernel name (first-appearance
// order) into buf; returns the number of bytes written (0 if nothing was recorded), negative on error.
int findb200_profile_read(char* buf, size_t cap) {
    using namespace fb;
    if (!buf || cap < 2) return FINDB200_ERR_ARG;
    buf[0] = 0;
    const int n = g_prof_n;
    if (n < 2) return 0;
    if (cudaDeviceSynchronize() != cudaSuccess) return FINDB200_ERR_CUDA;
    const char* names[256];
    double tot[256];
    int cnt[256], nn = 0;
    bool side[256];
    // marks of a side stream (the presort of a training step, chamfer_hooks.cuh) are timed along their own stream and reported
    // as "side:<name>": they run beside the main stream's kernels
    const cudaStream_t main_stream = g_prof_stream[0];
    for (int i = 0; i + 1 < n; ++i) {
        if (!g_prof_name[i]) continue;
        int j = i + 1;
        while (j < n && g_prof_stream[j] != g_prof_stream[i]) ++j;
        if (j == n) continue;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, g_prof_ev[i], g_prof_ev[j]) != cudaSuccess) { (void)cudaGetLastError(); continue; }
        const bool sd = g_prof_stream[i] != main_stream;
        int k = 0;
        while (k < nn && (side[k] != sd || strcmp(names[k], g_prof_name[i]) != 0)) ++k;
        if (k == nn) { if (nn == 256) continue; names[nn] = g_prof_name[i]; tot[nn] = 0; cnt[nn] = 0; side[nn] = sd; ++nn; }
        tot[k] += ms; cnt[k] += 1;
    }
    size_t off = 0;
    for (int k = 0; k < nn; ++k) {
        int w = snprintf(buf + off, cap - off, "%s%s,%d,%.6f\n", side[k] ? "side:" : "", names[k], cnt[k], tot[k]);
        if (w < 0 || (size_t)w >= cap - off) break;
        off += (size_t)w;
    }
    if (!g_prof_keep) g_prof_n = 0;
    return (int)off;
}
const char* findb200_last_error(void) { return fb::g_err; }
const char* findb200_version(void) { return "find_b200 0.1 (sm_100a)"; }
}
