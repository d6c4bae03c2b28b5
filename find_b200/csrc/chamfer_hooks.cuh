// find_b200/csrc/chamfer_hooks.cuh — internal hooks between the fused training step (step.cu) and the Chamfer code (chamfer.cu).
//
// Presorted training step.  The sort of the target clouds does not depend on the field's forward pass, and the deformed template
// can be laid out in the curve order of the TEMPLATE's vertices (it deforms smoothly; any order gives the same kNN bit for bit).
// So step.cu forks a side stream at the start of the step and runs train_presort there — memsets, the six sort kernels over
// {template (one cloud), targets (N clouds)}, the clearing of the step's accumulators — beside the forward kernels; after the
// join, findb200_chamfer_train with g_train_presorted set only gathers the deformed clouds in the template's order (one launch:
// SoA rows, permutation, chunk and super-chunk boxes) before the kNN.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace fb {
namespace chamfer {

bool presort_supported(int N, int P1, int P2);      // culled kNN modes, six-kernel sort, FINDB200_PRESORT != 0
int train_presort(const float* x_template, const float* y, const int64_t* y_len, int N, int P1, int P2, void* zero_a, size_t zero_a_bytes,
                  void* zero_b, size_t zero_b_bytes, void* ws, size_t ws_bytes, cudaStream_t st);
extern int g_presort;                                // -1 environment default, 0 off, 1 on (findb200_set_step_overlap)
extern thread_local bool g_train_presorted;          // consumed by the next findb200_chamfer_train on this thread

}  // namespace chamfer
namespace field { extern float g_dw_early_weight; }     // < 0 environment default (FINDB200_DW_SPLIT, else 1.65); 0 = one dW launch

// a side stream per device with fork / join event pairs (created on first use): which = 0 presort, 1 early dW launch
int side_stream(int which, cudaStream_t* s, cudaEvent_t* fork, cudaEvent_t* join);

}  // namespace fb
