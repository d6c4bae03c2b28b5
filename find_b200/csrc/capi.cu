// find_b200/csrc/capi.cu — error plumbing of the C ABI (include/find_b200.h).
#include "common.cuh"
#include "../../include/find_b200.h"
#include <cstdarg>
#include <cstdio>

namespace fb {

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

}  // namespace fb

extern "C" {
const char* findb200_last_error(void) { return fb::g_err; }
const char* findb200_version(void) { return "find_b200 0.1 (sm_100a)"; }
}
