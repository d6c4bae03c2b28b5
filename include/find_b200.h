/*
 * include/find_b200.h — C ABI of libfindb200.so (sm_100a): FIND's field + Chamfer hot path.
 *
 * The reference (OllieBoyne/FIND) is pure Python and has no FFI; the boundary a maintainer binds is the set
 * of torch / PyTorch3D calls on its training hot path.  Each entry point below names the reference call it
 * replaces (paths relative to the reference repo).  INTEGRATION.md shows the ctypes stub that goes with it.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless stated; row-major, contiguous, fp32 unless stated;
 *   - the caller owns every buffer, including `workspace` (size it with the *_workspace_bytes call);
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing synchronises;
 *   - return value 0 = success, negative = FINDB200_ERR_*; findb200_last_error() gives a thread-local message;
 *   - lengths (`*_len`) are int64 [N] as in PyTorch3D, or NULL meaning "all points valid".
 */
#ifndef FIND_B200_H
#define FIND_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FINDB200_OK 0
#define FINDB200_ERR_ARG (-1)        /* null / inconsistent argument            */
#define FINDB200_ERR_SHAPE (-2)      /* negative or unsupported size            */
#define FINDB200_ERR_WORKSPACE (-3)  /* workspace missing or too small          */
#define FINDB200_ERR_CUDA (-4)       /* a CUDA runtime call or launch failed    */
#define FINDB200_ERR_UNSUPPORTED (-5)

const char* findb200_last_error(void);
const char* findb200_version(void);

/* ------------------------------------------------------------------------------------------------------
 * Chamfer / 1-NN   (replaces pytorch3d.ops.knn_points(K=1) and pytorch3d.loss.chamfer_distance as called at
 *                   src/model/losses.py:77,85,88, src/eval/eval_3d.py:151,159,176-177)
 * dist = fmaf(dz,dz,fmaf(dy,dy,dx*dx)) in fp32, lowest index on ties; idx is int64 like PyTorch3D.
 * ------------------------------------------------------------------------------------------------------ */
size_t findb200_knn_workspace_bytes(int N, int P1, int P2);

/* knn_points(p1 [N,P1,3], p2 [N,P2,3], lengths1, lengths2, K=1) -> dists [N,P1], idx [N,P1] */
int findb200_knn1(const float* p1, const float* p2, const int64_t* len1, const int64_t* len2, int N, int P1, int P2,
                  float* dists, int64_t* idx, void* workspace, size_t workspace_bytes, void* stream);

/* autograd backward of the above: grad_p1 [N,P1,3], grad_p2 [N,P2,3] (both fully overwritten) */
int findb200_knn1_backward(const float* p1, const float* p2, const int64_t* len1, const int64_t* len2,
                           const int64_t* idx, const float* grad_dists, int N, int P1, int P2, float* grad_p1,
                           float* grad_p2, void* stream);

/* chamfer_distance(x, y, x_lengths, y_lengths, batch_reduction="mean", point_reduction="mean")[0]
 * loss: device scalar.  dist_x/idx_x [N,P1], dist_y/idx_y [N,P2]: optional (NULL to skip), idx_* are needed
 * by findb200_chamfer_backward. */
int findb200_chamfer_forward(const float* x, const float* y, const int64_t* x_len, const int64_t* y_len, int N, int P1,
                             int P2, float* loss, float* dist_x, int64_t* idx_x, float* dist_y, int64_t* idx_y,
                             void* workspace, size_t workspace_bytes, void* stream);

/* d loss / d x (and d y if grad_y != NULL); upstream gradient = (*grad_loss if non-NULL else 1) * grad_scale */
int findb200_chamfer_backward(const float* x, const float* y, const int64_t* x_len, const int64_t* y_len, int N,
                              int P1, int P2, const int64_t* idx_x, const int64_t* idx_y, const float* grad_loss,
                              float grad_scale, float* grad_x, float* grad_y, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Diagnostics: one-CTA tcgen05 self-test of the shared-memory image / descriptor / TMEM conventions used by
 * the field kernels (mode 0: K-major operands, forward and dX shape; mode 1: MN-major operands, dW shape).
 * a_img / b_img are fp16 swizzled tile images (see csrc/umma_selftest.cu); d_out is fp32 [128][256].
 * ------------------------------------------------------------------------------------------------------ */
int findb200_umma_selftest(int mode, const void* a_img, const void* b_img, float* d_out, void* stream);

/* FP32 CUDA-core peak probe (roofline denominator of the Chamfer kernel): launches blocks x 256 threads, each
 * issuing iters*64 dependent-chain FMAs with ILP 8 (mode 0: FFMA, mode 1: packed FFMA2 = 2 lanes/instr). */
int findb200_fp32_peak(int mode, int blocks, int iters, float* scratch, void* stream);

/* L2 -> shared-memory bulk-copy (cp.async.bulk) bandwidth probe: `blocks` CTAs x `issuers` warps each stream `iters`
 * chunks of `chunk_bytes` of the same `region_bytes` region through a `stages`-deep ring (how the field CTAs stream
 * packed weights). */
int findb200_l2_stream(const void* src, size_t region_bytes, int chunk_bytes, int blocks, int iters, int stages,
                       int issuers, float* scratch, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* FIND_B200_H */
