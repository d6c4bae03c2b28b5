// find_b200/csrc/optim.cu — the optimiser updates of FIND's training step as ONE launch per optimiser.
//
// The reference steps four torch optimisers after every backward (src/train/train.py:161-168, src/train/trainer.py:122-123):
//   Adam(main_params, lr_net), SGD(reg_params, lr_reg, momentum 0.9), Adam / SGD(latent_params, lr_latent).
// torch launches several kernels per parameter tensor (or a foreach chain); here one kernel walks a table of
// (param, grad, state) tensors, 4 floats per thread, so the whole update of the 868 358 MLP weights + the latent tables is
// one HBM pass: 16 B read + 12 B written per Adam element (p, g, m, v -> p, m, v), 12 + 8 for SGD momentum.
// Semantics follow torch.optim (single-tensor reference implementation, defaults: no amsgrad, no nesterov, dampening 0,
// maximize False), including DENSE updates of embedding-style tables: rows that were not in the batch have grad 0 and
// still move by their momentum, exactly as in the reference.  The step counter lives on the device so the launch can
// sit inside a captured CUDA graph.
#include "common.cuh"
#include "../../include/find_b200.h"

namespace fb {
namespace optim {

constexpr int kMaxTensors = 48;
struct Table {
    float* p[kMaxTensors];
    const float* g[kMaxTensors];
    float* m[kMaxTensors];
    float* v[kMaxTensors];
    long long start[kMaxTensors + 1];     // prefix sums of element counts, in units of 4-float groups
    long long numel[kMaxTensors];
    int n;
};

FB_DEVINL int find_tensor(const Table& t, long long grp) {
    int lo = 0, hi = t.n - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (t.start[mid] <= grp) lo = mid; else hi = mid - 1;
    }
    return lo;
}

// mode 0: Adam;  mode 1: SGD with momentum (m = momentum buffer; first step: buf = grad)
template <int MODE>
__global__ void __launch_bounds__(256) optim_kernel(const __grid_constant__ Table t, double lr_d, double b1_d, double b2_d, float eps, float wd,
                                                    const long long* __restrict__ step_dev) {
    // hyper-parameters arrive as doubles, as torch holds them: 1 - beta is formed in double and rounded once
    const float lr = (float)lr_d, b1 = (float)b1_d, b2 = (float)b2_d, omb1 = (float)(1.0 - b1_d), omb2 = (float)(1.0 - b2_d);
    const long long total = t.start[t.n];
    const long long step = *step_dev + 1;
    float step_size = lr, inv_sqrt_bc2 = 1.f;
    if (MODE == 0) {
        const double bc1 = 1.0 - pow(b1_d, (double)step), bc2 = 1.0 - pow(b2_d, (double)step);
        step_size = (float)(lr_d / bc1);
        inv_sqrt_bc2 = 1.f / (float)sqrt(bc2);        // torch divides by the scalar through its fp32 reciprocal
    }
    for (long long grp = (long long)blockIdx.x * blockDim.x + threadIdx.x; grp < total; grp += (long long)gridDim.x * blockDim.x) {
        const int k = find_tensor(t, grp);
        const long long e0 = (grp - t.start[k]) * 4;
        const long long n = t.numel[k];
        float* p = t.p[k] + e0;
        const float* g = t.g[k] + e0;
        float* m = t.m[k] + e0;
        float* v = MODE == 0 ? t.v[k] + e0 : nullptr;
        const int cnt = (int)min(4ll, n - e0);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (j >= cnt) break;
            float gj = g[j], pj = p[j];
            if (wd != 0.f) gj = gj + wd * pj;
            if (MODE == 0) {
                const float mj = m[j] + (gj - m[j]) * omb1;                    // exp_avg.lerp_(grad, 1 - beta1)
                const float vj = v[j] * b2 + omb2 * gj * gj;                   // exp_avg_sq.mul_(b2).addcmul_(g, g, 1 - b2)
                const float denom = sqrtf(vj) * inv_sqrt_bc2 + eps;
                p[j] = pj + (-step_size * mj) / denom;                          // addcdiv_(exp_avg, denom, value=-step_size)
                m[j] = mj; v[j] = vj;
            } else {
                const float mj = step == 1 ? gj : m[j] * b1 + gj;              // b1 = momentum
                p[j] = pj - lr * mj;
                m[j] = mj;
            }
        }
    }
}

__global__ void bump_step_kernel(long long* step) { *step += 1; }

}  // namespace optim
}  // namespace fb

using namespace fb;
using namespace fb::optim;

static int launch_optim(int mode, const findb200_optim_tensor* tensors, int n_tensors, double lr, double b1, double b2, double eps,
                        double wd, long long* step_dev, cudaStream_t st) {
    FB_REQUIRE(tensors && step_dev, FINDB200_ERR_ARG, "optim: null argument");
    FB_REQUIRE(n_tensors >= 1 && n_tensors <= kMaxTensors, FINDB200_ERR_SHAPE, "optim: 1..%d tensors per call (got %d)", kMaxTensors,
               n_tensors);
    Table t;
    memset(&t, 0, sizeof(t));
    long long groups = 0;
    for (int i = 0; i < n_tensors; ++i) {
        const findb200_optim_tensor& x = tensors[i];
        FB_REQUIRE(x.param && x.grad && x.m && (mode == 1 || x.v) && x.numel >= 0, FINDB200_ERR_ARG, "optim: tensor %d is incomplete", i);
        t.p[i] = x.param; t.g[i] = x.grad; t.m[i] = x.m; t.v[i] = x.v;
        t.start[i] = groups;
        groups += (x.numel + 3) / 4;
        t.numel[i] = x.numel;
    }
    t.start[n_tensors] = groups;
    t.n = n_tensors;
    if (groups == 0) return 0;
    long long blocks = (groups + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    FB_MARK(mode == 0 ? "adam" : "sgd_momentum", st);
    if (mode == 0)
        optim_kernel<0><<<(int)blocks, 256, 0, st>>>(t, lr, b1, b2, (float)eps, (float)wd, step_dev);
    else
        optim_kernel<1><<<(int)blocks, 256, 0, st>>>(t, lr, b1, 0.0, 0.f, (float)wd, step_dev);
    FB_CUDA(cudaGetLastError());
    bump_step_kernel<<<1, 1, 0, st>>>(step_dev);
    FB_CUDA(cudaGetLastError());
    FB_MARK_END(st);
    return 0;
}

extern "C" {

int findb200_adam_step(const findb200_optim_tensor* tensors, int n_tensors, double lr, double beta1, double beta2, double eps,
                       double weight_decay, long long* step_dev, void* stream) {
    return launch_optim(0, tensors, n_tensors, lr, beta1, beta2, eps, weight_decay, step_dev, (cudaStream_t)stream);
}

int findb200_sgd_momentum_step(const findb200_optim_tensor* tensors, int n_tensors, double lr, double momentum, double weight_decay,
                               long long* step_dev, void* stream) {
    return launch_optim(1, tensors, n_tensors, lr, momentum, 0.0, 0.0, weight_decay, step_dev, (cudaStream_t)stream);
}

}  // extern "C"
