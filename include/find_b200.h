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

/* findb200_knn1 for callers that cannot allocate or provide a workspace (CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.  Same inputs,
 * outputs and results as findb200_knn1; the scratch is a static device array of FINDB200_KNN1_STATIC_MAX_BYTES inside the
 * library (held by every GPU that loads it until the process exits), and the call performs no cudaMalloc / cudaFree and calls
 * no allocating routine.  Sizes whose findb200_knn_workspace_bytes(N, P1, P2) exceeds the limit are rejected with
 * FINDB200_ERR_WORKSPACE (8 MiB covers 8 clouds of 10 242 x 10 000 points or one of 50 000 x 100 000).  All calls share the
 * array: calls that overlap on the GPU (different streams or host threads) race — serialise them. */
#define FINDB200_KNN1_STATIC_MAX_BYTES (8u << 20)
int findb200_knn1_static(const float* p1, const float* p2, const int64_t* len1, const int64_t* len2, int N, int P1, int P2,
                         float* dists, int64_t* idx, void* stream);

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

/* The Chamfer + colour part of one training step in as few launches as possible (used by findb200_train_step): kNN both
 * ways, then ONE kernel for the indices, loss3 = [w_chamf * chamfer(x, y) + w_tex * mse(col, col_target), chamfer, mse]
 * (losses.py:87-88, model.py:1157-1158), d(w_chamf * chamfer)/dx and grad_col = d(w_tex * mse)/dcol (col == NULL: no colour
 * term).  The gradient with respect to x comes as 64-bit FIXED POINT, d/dx[n][i][k] = grad_x_fx[n][i][k] * grad_x_scale[n]
 * ([N,P1,3] int64 + [N] float): both Chamfer directions scatter into it with integer atomics, so two runs give the same bits
 * whatever the arrival order (PyTorch3D's knn_points_backward uses float atomicAdd).  mse_part: scratch of
 * >= 8 * ceil(max(P1,P2)/512) * N bytes.  zero_extra: an optional region (multiple of 4 bytes) cleared early in the call for
 * the kernels the caller enqueues next. */
int findb200_chamfer_train(const float* x, const float* y, const int64_t* x_len, const int64_t* y_len, int N, int P1, int P2,
                           float w_chamf, const float* col, const float* col_target, size_t n_col, float w_tex,
                           long long* grad_x_fx, float* grad_x_scale, float* grad_col, int64_t* idx_x, int64_t* idx_y,
                           float* loss3, double* mse_part, size_t mse_part_bytes, void* zero_extra, size_t zero_extra_bytes,
                           void* workspace, size_t workspace_bytes, void* stream);

/* d loss / d x (and d y if grad_y != NULL); upstream gradient = (*grad_loss if non-NULL else 1) * grad_scale */
int findb200_chamfer_backward(const float* x, const float* y, const int64_t* x_len, const int64_t* y_len, int N,
                              int P1, int P2, const int64_t* idx_x, const int64_t* idx_y, const float* grad_loss,
                              float grad_scale, float* grad_x, float* grad_y, void* stream);

/* Measurement hooks for the kNN kernel.  mode 0 (default): exact search with bounding-box culling over spatially
 * binned clouds; mode 1: the plain brute-force tile kernel (every pair evaluated), kept for A/B timing.  kq = queries
 * per lane of the culled kernel (0 = heuristic, 1, 2 or 4).  Results are bit-identical in every mode.
 * findb200_knn_stats copies {pair evaluations executed, queries that took the cross-chunk-tie path, most chunk
 * evaluations done for one query group, query groups} of the LAST culled call that used `workspace` into four host
 * uint64 (synchronises `stream`). */
int findb200_set_knn_mode(int mode, int kq);
/* Clouds of up to max_points points (and < 40000) are Hilbert-sorted by one 8-CTA cluster each in a single launch instead of
 * six kernels; 0 = never (the default: the single launch measured slower), negative = back to the default.  A/B measurement. */
int findb200_set_sort_cta_max(int max_points);
int findb200_knn_stats(const void* workspace, unsigned long long* out4_host, void* stream);
/* Side-stream schedule of the fused training step (A/B measurement and tests; both default on).  presort: 1 = the sort of the
 * target clouds and the curve order of the template run on a side stream beside the forward kernels and the deformed feet are laid
 * out in the template's order (bit-identical results), 0 = everything on the caller's stream, negative = environment default
 * (FINDB200_PRESORT).  dw_early_weight: share of an early CTA of the weight-gradient launch relative to a late one (the early CTAs
 * start on a side stream beside the trunk's backward chain), 0 = one launch, negative = default (FINDB200_DW_SPLIT, else 1.65); it
 * changes how the fp32 partial sums of dW are grouped, i.e. the last bits of the weight gradients. */
int findb200_set_step_overlap(int presort, float dw_early_weight);

/* ------------------------------------------------------------------------------------------------------
 * Neural displacement field   (replaces NeuralDisplacementField.forward + its autograd,
 *   src/model/model.py:393-453, the latent concat model.py:428-437, and the registration transform of
 *   get_meshes, model.py:481-491; called from ModelWithLoss.forward model.py:1009 and TextureLossGTSpace
 *   src/model/losses.py:50).  fp16 tensor-core GEMMs (tcgen05, fp32 accumulate), fp32 everywhere else.
 *
 * Parameter order in weight[] / bias[]: base.0,2,4,6,8 | mlp_disp.0,2,4,6 | mlp_col.0,2,4,6, each an
 * nn.Linear weight [out,in] / bias [out] exactly as in the reference state_dict (model.py:250-257,351-371).
 * mlp_disp.0 has in = 256 + shapevec_size + posevec_size, mlp_col.0 has in = 256 + texvec_size.
 * ------------------------------------------------------------------------------------------------------ */
typedef struct findb200_field_params {
    const float* weight[13];
    const float* bias[13];
    const float* fourier_B;   /* [3,256] Gaussian matrix of FourierFeatureTransform (fourier_feature_transform.py:23-26) */
    const float* avg_col;     /* [3] added to col, or NULL (model.py:449-451) */
    int shapevec_size, posevec_size, texvec_size;   /* widths of the latent blocks (0 = block absent) */
    int flags;                /* FINDB200_FIELD_* bits; 0 = recompute everything */
} findb200_field_params;
/* Caller's promise about the workspace it passes (same workspace, same `pos`, for_backward = 1 in an EARLIER forward):
 *   ENC_CACHED   the Fourier-encoded tiles of `pos` written by that forward are still there: the trunk loads them instead of
 *                recomputing 256 sin/cos pairs per vertex (the template is constant over training, model.py:285);
 *   TRUNK_CACHED also the trunk output is still valid (MLP weights unchanged since: latent-code fitting with a frozen MLP,
 *                train.py:218-224): the trunk forward is skipped altogether. */
#define FINDB200_FIELD_ENC_CACHED 1
#define FINDB200_FIELD_TRUNK_CACHED 2

typedef struct findb200_field_grads {   /* same order and shapes as findb200_field_params; every entry is overwritten */
    float* weight[13];
    float* bias[13];
} findb200_field_grads;

/* Scratch + saved-for-backward storage.  A forward with save_for_backward = 1 fills it; the matching backward
 * must be given the same, untouched buffer. */
size_t findb200_field_workspace_bytes(int nfeet, int V, int pos_shared, int for_backward);

/* pos: [V,3] if pos_shared (template shared by all feet, model.py:404-406,473-477) else [nfeet,V,3].
 * z_*: [nfeet, size] latent rows.  reg: [nfeet,9] = [t(3), euler XYZ(3), scale(3)] or NULL.
 * Outputs [nfeet,V,3]: disp = 0.1 tanh(.), col = 0.5 (1 + tanh(.)) (+avg_col) (NULL: colour head skipped),
 * verts = ((pos + disp) * scale) @ R + t (NULL: skipped; without reg: pos + disp). */
int findb200_field_forward(const findb200_field_params* p, const float* pos, int pos_shared, const float* z_shape,
                           const float* z_pose, const float* z_tex, const float* reg, int nfeet, int V,
                           int save_for_backward, float* disp, float* col, float* verts, void* workspace,
                           size_t workspace_bytes, void* stream);

/* Gradients of sum(grad_disp*disp) + sum(grad_col*col) + sum(grad_verts*verts) (each grad_* may be NULL).
 * disp / col: the forward outputs.  gw NULL = frozen MLP (latent-code fitting, src/train/train.py:218-224):
 * only gz_* and greg are produced.  gz_*: [nfeet,size]; greg: [nfeet,9]; NULL outputs are skipped. */
int findb200_field_backward(const findb200_field_params* p, const float* pos, int pos_shared, const float* z_shape,
                            const float* z_pose, const float* z_tex, const float* reg, int nfeet, int V,
                            const float* disp, const float* col, const float* grad_disp, const float* grad_col,
                            const float* grad_verts, const findb200_field_grads* gw, float* gz_shape, float* gz_pose,
                            float* gz_tex, float* greg, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Whole 3-D training step in one call (replaces, for the chamf(+tex) configuration, everything between
 * `loss, loss_dict = self.model(batch, ...)` and `loss.backward()` of Trainer.train_epoch,
 * src/train/trainer.py:106-121: ModelWithLoss.forward model.py:1009,1026-1030,1157-1158 -> get_meshes model.py:455-504
 * -> DisplacementLoss plain branch losses.py:87-88 -> autograd).  Chamfer is taken between the deformed template
 * vertices [nfeet,V,3] and target [nfeet,P2,3] (target_len int64 [nfeet] or NULL); col_target [nfeet,V,3] (or NULL)
 * adds w_tex * mean((col - col_target)^2).  loss3 = {total, chamfer, colour mse} (device).  disp/col/verts are
 * the forward outputs.  Gradients as in findb200_field_backward (gw NULL = frozen MLP).  Nothing synchronises; the
 * sequence can be captured in a CUDA graph.
 * ------------------------------------------------------------------------------------------------------ */
size_t findb200_train_step_workspace_bytes(int nfeet, int V, int P2);

int findb200_train_step(const findb200_field_params* p, const float* template_pos, const float* z_shape,
                        const float* z_pose, const float* z_tex, const float* reg, const float* target,
                        const int64_t* target_len, const float* col_target, int nfeet, int V, int P2, float w_chamf,
                        float w_tex, float* loss3, float* disp, float* col, float* verts,
                        const findb200_field_grads* gw, float* gz_shape, float* gz_pose, float* gz_tex, float* greg,
                        void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Mesh side of the 3-D loss.  faces / edges / incidence lists are int32 device arrays.
 * ------------------------------------------------------------------------------------------------------ */

/* pytorch3d.ops.sample_points_from_meshes(meshes, num_samples, return_textures)  — src/model/losses.py:39,63,67,
 * src/eval/eval_3d.py:149-150.  verts [N,V,3] (verts_shared: one [V,3] array for all meshes), faces [N,F,3] or shared
 * [F,3]; num_faces [N] int64 valid faces per mesh (NULL = F).  The draw: Philox4x32-10 keyed by `seed` over an integer
 * area CDF (see csrc/mesh.cu), counter words 2-3 = *draw_counter (device uint64, NULL = 0: lets a captured graph draw
 * fresh samples per replay); pass face_idx_in [N,S] (mesh-local) and uv_in [2,N,S] (PyTorch3D's
 * _rand_barycentric_coords layout) to inject a draw instead.  Outputs: samples [N,S,3]; face_idx [N,S] and bary [N,S,3]
 * (optional, needed by the backward); attr_out [N,S,C] = barycentric interpolation of vert_attr [N,V,C] (optional:
 * TexturesVertex colours for return_textures=True).  Workspace only needed when the draw is made here. */
size_t findb200_sample_workspace_bytes(int N, int F);
int findb200_sample_points_forward(const float* verts, int verts_shared, const int32_t* faces, int faces_shared,
                                   const int64_t* num_faces, int N, int V, int F, int S, unsigned long long seed,
                                   const unsigned long long* draw_counter, const int64_t* face_idx_in, const float* uv_in, float* samples, int64_t* face_idx,
                                   float* bary, const float* vert_attr, int attr_shared, int C, float* attr_out,
                                   void* workspace, size_t workspace_bytes, void* stream);
/* autograd of the interpolation: grad_verts[n, faces[f][k]] += bary[k] * grad_scale * grad_samples[n,s]
 * (grad_verts [N,V,3]; zeroed first unless accumulate != 0) */
int findb200_sample_points_backward(const float* grad_samples, const int32_t* faces, int faces_shared,
                                    const int64_t* face_idx, const float* bary, int N, int V, int F, int S,
                                    float grad_scale, int accumulate, float* grad_verts, void* stream);

/* MeshSmoothnessLoss (src/model/losses.py:93-99) = w_lap * mesh_laplacian_smoothing(meshes, "cot") +
 * w_edge * mesh_edge_loss(meshes, target_length) on N meshes sharing one topology (the deformed template):
 * faces [F,3], unique edges [E,2], vertex -> (face, corner) incidence in CSR form (vf_ptr [V+1], vf_face / vf_corner [nnz]).
 * loss2[0] = edge loss, loss2[1] = laplacian loss (unweighted, device).  grad_verts (optional, [N,V,3]) receives
 * w_edge * d edge + w_lap * d laplacian (added when accumulate != 0).  grad_loss must be NULL (fold it into the weights). */
size_t findb200_mesh_smooth_workspace_bytes(int N, int V, int nnz);
int findb200_mesh_smooth(const float* verts, int N, int V, const int32_t* faces, int F, const int32_t* edges, int E,
                         const int32_t* vf_ptr, const int32_t* vf_face, const int32_t* vf_corner, int nnz,
                         float target_length, float w_edge, float w_lap, const float* grad_loss, float* loss2,
                         float* grad_verts, int accumulate, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Soft-silhouette rasteriser (first slice): pytorch3d rasterize_meshes (faces_per_pixel = 100,
 * blur_radius = log(1/1e-4 - 1) * sigma, clipped perspective-correct barycentrics) + SoftSilhouetteShader as set up at
 * src/model/renderer.py:122-128 and called at renderer.py:208-245, 309-311 (FootRenderer.forward(return_mask=True)).
 * verts_ndc [N,V,3] fp32: x, y in PyTorch3D NDC (+x left, +y up), z = view-space depth — what MeshRasterizer.transform
 * produces; N counts meshes x views.  faces [F,3] int32 shared by all N.  alpha [N,H,W] = 1 - prod_k (1 - sigmoid(-signed
 * squared distance_k / sigma)) over the faces_per_pixel candidates of smallest depth (earlier face on ties).
 * prod [N,H,W] fp32 and sel [N,H,W] x 8 bytes are saved for the backward (may be NULL for inference).  pix_to_face / zbuf
 * [N,H,W] (optional) receive the nearest candidate face of each pixel (index within the mesh, -1 = none) and its depth: the
 * [..., 0] slice of the silhouette Fragments that FootRenderer's mask_out_faces reads (renderer.py:318-327).  The backward ADDS
 * d alpha / d (x, y) into grad_verts_ndc [N,V,3] (z receives no gradient) and must be given the workspace the forward
 * filled.  findb200_silhouette_status synchronises and returns 1 if a tile could not apply the K truncation exactly
 * (> 4096 faces on one 16 x 16 tile together with a pixel of more than faces_per_pixel candidates).
 * ------------------------------------------------------------------------------------------------------ */
size_t findb200_silhouette_workspace_bytes(int N, int F);
size_t findb200_silhouette_saved_bytes(int N, int F);   /* prefix of the workspace the backward needs; the rest is scratch */
int findb200_silhouette_forward(const float* verts_ndc, const int32_t* faces, int N, int V, int F, int H, int W,
                                float blur_radius, float sigma, int faces_per_pixel, int perspective_correct,
                                float* alpha, float* prod, void* sel, int32_t* pix_to_face, float* zbuf, void* workspace,
                                size_t workspace_bytes, void* stream);
int findb200_silhouette_backward(const float* grad_alpha, const int32_t* faces, int N, int V, int F, int H, int W,
                                 float blur_radius, float sigma, int faces_per_pixel, int perspective_correct,
                                 const float* prod, const void* sel, float* grad_verts_ndc, void* workspace,
                                 size_t workspace_bytes, void* stream);
int findb200_silhouette_status(const void* workspace, int N, int F, void* stream);
/* Hard rasterisation, faces_per_pixel = 1 (the image rasteriser of src/model/renderer.py:116-120, called at :208-245): for every
 * pixel the nearest candidate face (index within the mesh, -1 = none), its depth, barycentrics [N,H,W,3] (perspective-correct
 * when asked, clipped + renormalised when asked) and signed squared distance (negative inside).  Same workspace size as the
 * silhouette.  Forward only. */
int findb200_rasterize_nearest(const float* verts_ndc, const int32_t* faces, int N, int V, int F, int H, int W,
                               float blur_radius, int perspective_correct, int clip_barycentric_coords,
                               int32_t* pix_to_face, float* zbuf, float* bary, float* dists, void* workspace,
                               size_t workspace_bytes, void* stream);

/* Phong image of the hard fragments, forward only: SoftPhongShader (TexturesVertex + phong_shading with PointLights and default
 * Materials + softmax_rgb_blend) as FootRenderer builds it (src/model/renderer.py:110-115, 133) and calls it (:273-275).
 * findb200_vertex_normals = Meshes.verts_normals_padded() (area-weighted, normalised) over the vertex -> (face, corner) CSR list.
 * Images are mesh-major (image = mesh * num_views + view); verts / normals / colours are per mesh [M,V,3]; cam_centres
 * [num_views,3]; shading = 20 HOST floats: light xyz, ambient rgb, diffuse rgb, specular rgb, background rgb, shininess, sigma,
 * gamma, znear, zfar. */
int findb200_vertex_normals(const float* verts, int N, int V, const int32_t* faces, const int32_t* vf_ptr, const int32_t* vf_face,
                            const int32_t* vf_corner, float* normals, void* stream);
int findb200_phong_shade(const int32_t* pix_to_face, const float* bary, const float* zbuf, const float* dists, int n_images, int H,
                         int W, int num_views, const float* verts_world, const float* vert_normals, const float* vert_colours,
                         const int32_t* faces, int V, const float* cam_centres, const float* shading, float* image, void* stream);

/* Backward of the Phong image (FootRenderer.forward(return_images=True) with gradients: the predicted renders of the pixel /
 * perceptual losses, model.py:1089-1107): d/d(world vertices), d/d(vertex normals), d/d(vertex colours) and d/d(NDC vertices of
 * every image) from d/d(image), ACCUMULATED into the caller's pre-zeroed buffers.  findb200_vertex_normals_backward then carries
 * d/d(normals) to the vertices (scratch: N * V * 3 floats), and the camera projection's backward is the caller's (torch autograd
 * of FootRenderer.project).  Same arguments as the forward plus verts_ndc [n_images, V, 3]. */
int findb200_phong_shade_backward(const float* grad_image, const int32_t* pix_to_face, const float* bary, const float* zbuf,
                                  const float* dists, int n_images, int H, int W, int num_views, const float* verts_ndc,
                                  const float* verts_world, const float* vert_normals, const float* vert_colours,
                                  const int32_t* faces, int V, const float* cam_centres, const float* shading,
                                  float* grad_verts_world, float* grad_normals, float* grad_colours, float* grad_verts_ndc,
                                  void* stream);
int findb200_vertex_normals_backward(const float* verts, int N, int V, const int32_t* faces, const int32_t* vf_ptr,
                                     const int32_t* vf_face, const int32_t* vf_corner, const float* grad_normals, float* scratch,
                                     float* grad_verts, void* stream);

/* Keypoint overlay (FootRenderer.forward(keypoints=...), /root/reference/src/model/renderer.py:139-142, 365-373; call site
 * src/eval/eval_3d.py:97): PyTorch3D's PointsRenderer — PointsRasterizer(radius, points_per_pixel) + AlphaCompositor without a
 * background.  points_ndc [N, P, 3] = NDC x, y and view-space depth (FootRenderer.project), features [N, P, C], C <= 8,
 * points_per_pixel <= 16 -> image [N, H, W, C].  A point covers the pixels whose centre is closer than `radius` (NDC units) and
 * has depth >= 0; the nearest `points_per_pixel` are composited front to back with weight 1 - d^2 / radius^2.  Forward only. */
int findb200_points_render(const float* points_ndc, const float* features, int N, int P, int C, int H, int W, float radius,
                           int points_per_pixel, float* image, void* stream);

/* FeatureShader (renderer.py:74-105) = the reference's own softmax_blend (renderer.py:23-72, default znear = 1, zfar = 100) over the
 * fragments of a findb200_silhouette_forward call: per-vertex features [M, V, C] (M = N / num_views) interpolated with the clipped
 * barycentrics of each kept face and blended with weights sigmoid(-dist / sigma) * exp((zinv - max zinv) / gamma).  `ws`, `sel`,
 * `zbuf`, `pix_to_face` are that call's workspace and outputs.  Forward: out [N, H, W, C], den [N, H, W] (kept for the backward).
 * Backward: d/d(features) and d/d(NDC vertices), ACCUMULATED into pre-zeroed buffers.  background: [C] on the device or NULL. */
int findb200_feature_blend_forward(const float* features, int C, int num_views, const int32_t* faces, int N, int V, int F, int H, int W,
                                   float blur_radius, float sigma, int faces_per_pixel, int perspective_correct, float gamma,
                                   float znear, float zfar, const float* background, const void* sel, const float* zbuf, float* out,
                                   float* den, void* workspace, size_t workspace_bytes, void* stream);
int findb200_feature_blend_backward(const float* grad_out, const float* features, int C, int num_views, const int32_t* faces, int N,
                                    int V, int F, int H, int W, float blur_radius, float sigma, int faces_per_pixel,
                                    int perspective_correct, float gamma, float znear, float zfar, const float* background,
                                    const void* sel, const float* zbuf, const int32_t* pix_to_face, const float* out, const float* den,
                                    float* grad_features, float* grad_verts_ndc, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * Optimiser updates: torch.optim.Adam(main_params / latent_params) and
 * torch.optim.SGD(reg_params, momentum=0.9) as stepped at src/train/train.py:161-168, src/train/trainer.py:122-123.
 * One launch updates every tensor of the table (dense updates, like torch on dense .grad).  `tensors` is a HOST array of
 * device pointers; step_dev is a device int64 holding the number of steps taken so far (read, then incremented on the
 * stream), which keeps the call legal inside a captured CUDA graph.
 * ------------------------------------------------------------------------------------------------------ */
typedef struct findb200_optim_tensor {
    float* param;         /* updated in place */
    const float* grad;
    float* m;             /* Adam exp_avg / SGD momentum_buffer */
    float* v;             /* Adam exp_avg_sq (unused by SGD, may be NULL) */
    long long numel;
} findb200_optim_tensor;
int findb200_adam_step(const findb200_optim_tensor* tensors, int n_tensors, double lr, double beta1, double beta2, double eps,
                       double weight_decay, long long* step_dev, void* stream);
int findb200_sgd_momentum_step(const findb200_optim_tensor* tensors, int n_tensors, double lr, double momentum, double weight_decay,
                               long long* step_dev, void* stream);

/* The shipped 3-D configuration (cfgs/train_3d.yaml: chamf + smooth + tex) in one call: findb200_train_step plus
 *   - Chamfer on num_samples area-weighted surface samples of the deformed template instead of its vertices
 *     (DisplacementLoss, src/model/losses.py:67,88; 0 = on the vertices), gradient gathered back to the vertices;
 *   - w_smooth * MeshSmoothnessLoss(deformed template) (src/model/model.py:1032-1033, losses.py:93-99; opts.py:97 default
 *     weight 1000; 0 = off).
 * The target cloud is given as points (the GT-mesh samples of losses.py:63 are drawn by findb200_sample_points_forward on
 * the data side).  loss3[0] includes the smoothness term; loss_extra (optional) = {edge loss, laplacian loss}. */
typedef struct findb200_mesh_terms {
    const int32_t* faces; int F;               /* template topology (device, int32): faces [F,3]            */
    const int32_t* edges; int E;               /* unique edges [E,2]                       (smoothness only) */
    const int32_t* vf_ptr;                     /* [V+1] CSR of vertex -> (face, corner)    (smoothness only) */
    const int32_t* vf_face;
    const int32_t* vf_corner;
    int nnz;
    int num_samples;
    unsigned long long seed;
    unsigned long long* draw_counter;          /* device uint64, bumped after every draw (NULL: same draw every call) */
    float w_smooth;
    float* loss_extra;                         /* device [2] or NULL */
} findb200_mesh_terms;
size_t findb200_train_step_mesh_workspace_bytes(int nfeet, int V, int P2, int num_samples, int nnz, int F);
int findb200_train_step_mesh(const findb200_field_params* p, const float* template_pos, const float* z_shape,
                             const float* z_pose, const float* z_tex, const float* reg, const float* target,
                             const int64_t* target_len, const float* col_target, int nfeet, int V, int P2, float w_chamf,
                             float w_tex, float* loss3, float* disp, float* col, float* verts,
                             const findb200_field_grads* gw, float* gz_shape, float* gz_pose, float* gz_tex, float* greg,
                             const findb200_mesh_terms* mesh_terms, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------------------
 * The exchange step of the sharded training step: in-place SUM (times `scale`) of the flat MLP-gradient
 * bucket over `world` ranks, one kernel over NVLink / NVSwitch peer memory (two-shot, deterministic, bit-identical on
 * every rank).  peer_bufs / peer_flags: HOST arrays of `world` device pointers — every rank's bucket (n_floats fp32,
 * multiple of 4, 16-byte aligned) and flag page (findb200_allreduce_p2p_flag_bytes() bytes, zeroed once) as mapped into THIS
 * process (e.g. torch symmetric memory); epoch_dev: device u32[2], zero-initialised, owned by the call sequence.  Graph-capturable.
 * Every rank must enqueue the call the same number of times.
 * ------------------------------------------------------------------------------------------------------ */
size_t findb200_allreduce_p2p_flag_bytes(void);
int findb200_allreduce_p2p(void* const* peer_bufs, void* const* peer_flags, int rank, int world, long long n_floats,
                           float scale, unsigned int* epoch_dev, int n_ctas, void* stream);
/* The same exchange with the sum taken inside the NVSwitch (NVLS): mc_buf is the MULTICAST mapping of the bucket
 * (multimem.ld_reduce adds the addressed bytes over every rank's copy, multimem.st writes the result into all of them); the
 * barriers use the same peer flag pages.  Needs a multicast object over the ranks (torch symmetric memory reports it as
 * handle.multicast_ptr != 0); results are bit-identical on every rank, the order of the additions is the switch's. */
int findb200_allreduce_mc(void* mc_buf, void* const* peer_flags, int rank, int world, long long n_floats, float scale,
                          unsigned int* epoch_dev, int n_ctas, void* stream);

/* Kernel counters of the last findb200_train_step that used this workspace: same four numbers as findb200_knn_stats
 * ({pair evaluations executed, cross-chunk-tie queries, most chunk evaluations of one group, groups}); synchronises. */
int findb200_train_step_knn_stats(const void* workspace, int nfeet, int V, int P2, unsigned long long* out4_host,
                                  void* stream);

/* Latent-table access (replaces LatentVector.__getitem__ with an index tensor, src/model/model.py:130-152 as called
 * from sample_latent_vectors src/train/trainer.py:29-46, and the index scatter-add of its autograd).
 * table / grad_table: [n_rows,width]; idx: int64 [B] (device); out / g: [B,width].  Duplicate rows accumulate.
 * An index outside [0, n_rows) traps the kernel with a message, like torch's device-side assert for `self.data[idx]`. */
int findb200_gather_rows(const float* table, int n_rows, int width, const int64_t* idx, int B, float* out, void* stream);
int findb200_scatter_add_rows(float* grad_table, int n_rows, int width, const int64_t* idx, int B, const float* g,
                              void* stream);
/* the same gather for up to 8 tables in one launch (shapevec / posevec / texvec / reg rows of a batch).  A job's own `idx`
 * (non-NULL) overrides the common one: labelled tables are indexed by batch[vec.key], the others by batch['idx']
 * (sample_latent_vectors, src/train/trainer.py:29-46). */
typedef struct findb200_gather_job { const float* table; int n_rows; int width; float* out; const int64_t* idx; } findb200_gather_job;
int findb200_gather_rows_multi(const findb200_gather_job* jobs, int n_jobs, const int64_t* idx, int B, void* stream);
/* Dense table gradients of one step for up to 8 tables in one launch, WITHOUT atomics (bit-reproducible): row r of
 * grad_table = (accumulate ? itself : 0) + the sum, in batch order, of g[b] over the batch entries with idx[b] == r — what
 * loss.backward() leaves in LatentVector.data.grad after optimiser.zero_grad() (src/train/trainer.py:99-100,121). */
typedef struct findb200_scatter_job { const float* g; int n_rows; int width; float* grad_table; const int64_t* idx; } findb200_scatter_job;
int findb200_scatter_rows_multi(const findb200_scatter_job* jobs, int n_jobs, const int64_t* idx, int B, int accumulate,
                                void* stream);


/* ------------------------------------------------------------------------------------------------------
 * Diagnostics: one-CTA tcgen05 self-test of the shared-memory image / descriptor / TMEM conventions used by
 * the field kernels (mode 0: K-major operands, forward and dX shape; mode 1: MN-major operands, dW shape;
 * modes 2 / 3: the CTA-pair form, one tcgen05.mma.cta_group::2 stream with M = 256 over a cluster of two CTAs,
 * each staging half of the weight operand — 2: peer -> leader relay by remote mbarrier arrive, 3: the peer's bulk
 * copies signal the leader's mbarrier directly).
 * a_img / b_img are fp16 swizzled tile images (see csrc/umma_selftest.cu); d_out is fp32 [128][256]
 * ([256][256] for modes 2 / 3).
 * ------------------------------------------------------------------------------------------------------ */
int findb200_umma_selftest(int mode, const void* a_img, const void* b_img, float* d_out, void* stream);
/* Tensor-pipe rate probe (csrc/umma_selftest.cu): one CTA (pair) issues iters x 16 K = 16 MMAs of the chain kernels' shapes
 * back to back; out2_dev[0] = SM cycles from first issue to completion, [1] = MMAs issued.  mode 0: cta_group::1 M128 N256
 * operands from shared memory, 1: N128, 2: cta_group::2 M256 N256, 3 / 4: as 0 / 2 with A from TMEM.  flags add what the chain
 * kernels do around the MMAs (1 commit per 4 MMAs, 2 tcgen05 fence per 4, 4 mbarrier wait per 4, 8 background bulk-copy stream from
 * gsrc_1mb (>= 1 MiB, pair modes), 16 background shared-memory writers). */
int findb200_umma_rate(int mode, int iters, int flags, const void* gsrc_1mb, unsigned long long* out2_dev, void* stream);
/* Cost of one synchronisation primitive of the chain kernels' issue path for one elected thread (csrc/umma_selftest.cu lists
 * the ops); out2_dev[0] = cycles for `iters` repetitions, out2_dev[1] = iters.  Diagnostics only. */
int findb200_sync_probe(int op, int iters, unsigned long long* out2_dev, void* stream);

/* Launch profiler used by bench.py for the live roofline: after findb200_profile_enable(1) every kernel launch of the
 * library is bracketed by CUDA events on its launching stream; findb200_profile_read synchronises and writes
 * "kernel,launches,total_ms" lines into the HOST buffer buf (returns bytes written) and clears the record.
 * In-graph timing: enable(1) just before capturing a step into a CUDA graph makes the marks event-record NODES of the graph;
 * enable(2) after the capture stops recording but keeps the marks, and findb200_profile_read after every replay (it
 * synchronises) returns that replay's kernel durations without the host's launch gaps.  enable(0) ends either mode. */
int findb200_chain_trace(void* dev_buf_256_u64);   /* diagnostics: timeline of one head_fwd item (csrc/diag.cu) */
int findb200_profile_enable(int on);
int findb200_profile_read(char* buf, size_t cap);

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
