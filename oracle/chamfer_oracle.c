/*
 * oracle/chamfer_oracle.c — TEST INFRASTRUCTURE ONLY (never imported by find_b200/).
 *
 * CPU restatement of the third-party arithmetic FIND's 3-D loss reaches:
 *   pytorch3d @ 1706eb8216248e54f68cad86f7ea4125c79a3ca4  (requirements_mac_linux.txt:31)
 *     pytorch3d.ops.knn_points(K=1, norm=2)   <- reference call sites src/eval/eval_3d.py:176-177
 *     pytorch3d.loss.chamfer_distance(...)    <- reference call sites src/model/losses.py:77,85,88
 *                                                                 src/eval/eval_3d.py:151,159
 * PyTorch3D's source is NOT on this box, so this file restates its published algorithm
 * (SURVEY.md §8(a) rows 8-9).  PARITY UNPINNED: the reference ships no tests or golden vectors
 * for this path; the oracle is pinned only by the known-answer tests in tests/test_oracle_chamfer.py.
 *
 * knn (brute force, per query scan j = 0 .. len2-1 in ascending order):
 *     dist = sum_d (p1_d - p2_d)^2, accumulated in order d = 0,1,2
 *     the first candidate is always taken, later ones only on strict `<`  => lowest index wins ties
 * Two roundings of `dist` are provided:
 *     mode 0  FMA chain  fmaf(dz,dz, fmaf(dy,dy, dx*dx))   (what nvcc emits for `dist += diff*diff`
 *                                                           in pytorch3d/csrc/knn/knn.cu; THE CONTRACT)
 *     mode 1  separately rounded products                   (pytorch3d/csrc/knn/knn_cpu.cpp on x86)
 * Queries i >= len1[n] and batches with len2[n] == 0 keep dist = 0, idx = 0 (zero-initialised outputs).
 *
 * Build: see oracle/Makefile (gcc -O3 -ffp-contract=off -mavx2 -mfma -fopenmp).
 */
#include <math.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static inline float dist_fma(const float *a, const float *b) {
    float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    float d = dx * dx;
    d = fmaf(dy, dy, d);
    d = fmaf(dz, dz, d);
    return d;
}

static inline float dist_sep(const float *a, const float *b) {
    float dx = a[0] - b[0], dy = a[1] - b[1], dz = a[2] - b[2];
    volatile float xx = dx * dx, yy = dy * dy, zz = dz * dz; /* volatile: forbid re-association */
    float d = 0.0f + xx;
    d = d + yy;
    d = d + zz;
    return d;
}

/* knn_points(p1, p2, lengths1, lengths2, K=1).  len1/len2 may be NULL (= full).  threads<=1: serial. */
int oracle_knn1(const float *p1, const float *p2, const int64_t *len1, const int64_t *len2,
                int64_t N, int64_t P1, int64_t P2, int mode, int threads,
                float *dists /*[N,P1]*/, int64_t *idx /*[N,P1]*/) {
    if (N < 0 || P1 < 0 || P2 < 0) return -1;
    memset(dists, 0, sizeof(float) * (size_t)(N * P1));
    memset(idx, 0, sizeof(int64_t) * (size_t)(N * P1));
    for (int64_t n = 0; n < N; ++n) {
        int64_t l1 = len1 ? len1[n] : P1, l2 = len2 ? len2[n] : P2;
        if (l1 > P1 || l2 > P2 || l1 < 0 || l2 < 0) return -2;
        if (l2 == 0) continue;
        const float *q = p1 + n * P1 * 3, *t = p2 + n * P2 * 3;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(threads > 1 ? threads : 1)
#endif
        for (int64_t i = 0; i < l1; ++i) {
            float best = 0.f;
            int64_t bi = 0;
            for (int64_t j = 0; j < l2; ++j) {
                float d = mode == 0 ? dist_fma(q + 3 * i, t + 3 * j) : dist_sep(q + 3 * i, t + 3 * j);
                if (j == 0 || d < best) { best = d; bi = j; }
            }
            dists[n * P1 + i] = best;
            idx[n * P1 + i] = bi;
        }
    }
    return 0;
}

/* knn_points backward (norm=2, K=1): grad_p1[n,i] += 2 g (p1_i - p2_idx);  grad_p2[n,idx] -= same.
 * pytorch3d does the scatter with atomicAdd (order unspecified); here it is sequential in i. */
int oracle_knn1_backward(const float *p1, const float *p2, const int64_t *len1, const int64_t *len2,
                         const int64_t *idx, const float *grad_dists,
                         int64_t N, int64_t P1, int64_t P2,
                         float *grad_p1 /*[N,P1,3]*/, float *grad_p2 /*[N,P2,3]*/) {
    memset(grad_p1, 0, sizeof(float) * (size_t)(N * P1 * 3));
    memset(grad_p2, 0, sizeof(float) * (size_t)(N * P2 * 3));
    for (int64_t n = 0; n < N; ++n) {
        int64_t l1 = len1 ? len1[n] : P1, l2 = len2 ? len2[n] : P2;
        if (l2 == 0) continue;
        for (int64_t i = 0; i < l1; ++i) {
            int64_t j = idx[n * P1 + i];
            float g = grad_dists[n * P1 + i];
            for (int d = 0; d < 3; ++d) {
                float diff = 2.0f * g * (p1[(n * P1 + i) * 3 + d] - p2[(n * P2 + j) * 3 + d]);
                grad_p1[(n * P1 + i) * 3 + d] += diff;
                grad_p2[(n * P2 + j) * 3 + d] -= diff;
            }
        }
    }
    return 0;
}

/* chamfer_distance(x, y, x_lengths, y_lengths, batch_reduction="mean", point_reduction="mean"):
 *   cham_x[n] = sum_i dist_x[n,i] / clamp(len_x[n],1);  loss = (sum_n cham_x[n] + sum_n cham_y[n]) / max(N,1)
 * Sums are taken in double here (torch's fp32 reduction order is unspecified); compare with tolerance.
 * Optional outputs: NN indices/dists of both directions and d loss / d x, d loss / d y (upstream grad 1). */
int oracle_chamfer(const float *x, const float *y, const int64_t *xlen, const int64_t *ylen,
                   int64_t N, int64_t P1, int64_t P2, int mode, int threads,
                   double *loss, float *dist_x, int64_t *idx_x, float *dist_y, int64_t *idx_y,
                   float *grad_x, float *grad_y) {
    int rc = oracle_knn1(x, y, xlen, ylen, N, P1, P2, mode, threads, dist_x, idx_x);
    if (rc) return rc;
    rc = oracle_knn1(y, x, ylen, xlen, N, P2, P1, mode, threads, dist_y, idx_y);
    if (rc) return rc;
    double tot = 0.0;
    for (int64_t n = 0; n < N; ++n) {
        int64_t lx = xlen ? xlen[n] : P1, ly = ylen ? ylen[n] : P2;
        double sx = 0.0, sy = 0.0;
        for (int64_t i = 0; i < lx; ++i) sx += dist_x[n * P1 + i];
        for (int64_t j = 0; j < ly; ++j) sy += dist_y[n * P2 + j];
        tot += sx / (double)(lx < 1 ? 1 : lx) + sy / (double)(ly < 1 ? 1 : ly);
    }
    *loss = tot / (double)(N < 1 ? 1 : N);
    if (grad_x && grad_y) {
        memset(grad_x, 0, sizeof(float) * (size_t)(N * P1 * 3));
        memset(grad_y, 0, sizeof(float) * (size_t)(N * P2 * 3));
        for (int64_t n = 0; n < N; ++n) {
            int64_t lx = xlen ? xlen[n] : P1, ly = ylen ? ylen[n] : P2;
            float wx = 1.0f / ((float)(lx < 1 ? 1 : lx) * (float)(N < 1 ? 1 : N));
            float wy = 1.0f / ((float)(ly < 1 ? 1 : ly) * (float)(N < 1 ? 1 : N));
            if (ly > 0)
                for (int64_t i = 0; i < lx; ++i) {
                    int64_t j = idx_x[n * P1 + i];
                    for (int d = 0; d < 3; ++d) {
                        float g = 2.0f * wx * (x[(n * P1 + i) * 3 + d] - y[(n * P2 + j) * 3 + d]);
                        grad_x[(n * P1 + i) * 3 + d] += g;
                        grad_y[(n * P2 + j) * 3 + d] -= g;
                    }
                }
            if (lx > 0)
                for (int64_t j = 0; j < ly; ++j) {
                    int64_t i = idx_y[n * P2 + j];
                    for (int d = 0; d < 3; ++d) {
                        float g = 2.0f * wy * (y[(n * P2 + j) * 3 + d] - x[(n * P1 + i) * 3 + d]);
                        grad_y[(n * P2 + j) * 3 + d] += g;
                        grad_x[(n * P1 + i) * 3 + d] -= g;
                    }
                }
        }
    }
    return 0;
}

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
