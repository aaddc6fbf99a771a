/*
 * gsplat_b200.h -- C ABI of the B200-native 3DGS-Avatar hot path (libgsplat_b200.so).
 *
 * Plain pointers and sizes only: no torch types, no C++ types.  All pointers are DEVICE pointers unless a
 * name ends in _host.  The caller owns every buffer (outputs and workspaces); the library never
 * allocates device memory and keeps no global state, so it is safe for one stream per process-per-GPU.
 * Every entry point takes the CUDA stream to launch on (a cudaStream_t passed as void*).
 *
 * Which reference interface each entry point replaces (paths under
 * /root/reference/submodules/diff-gaussian-rasterization/):
 *   gs_forward        <- CudaRasterizer::Rasterizer::forward   cuda_rasterizer/rasterizer.h:31-58,
 *                        cuda_rasterizer/rasterizer_impl.cu:198-336 (called by RasterizeGaussiansCUDA,
 *                        rasterize_points.cu:35-115)
 *   gs_backward       <- CudaRasterizer::Rasterizer::backward  cuda_rasterizer/rasterizer.h:60-84,
 *                        rasterizer_impl.cu:340-434 (RasterizeGaussiansBackwardCUDA, rasterize_points.cu:117-196)
 *   gs_mark_visible   <- CudaRasterizer::Rasterizer::markVisible rasterizer.h:24-29, rasterizer_impl.cu:141-153
 *   gs_*_bytes        <- required<GeometryState/ImageState/BinningState>() rasterizer_impl.h:66-72 and the three
 *                        std::function<char*(size_t)> resize callbacks of forward()
 *   gs_debug_*        <- no counterpart; exposes the private workspace (keys, point_list, ranges, n_contrib,
 *                        final_T) that the bit-exact parity tests compare
 *   pose_forward / pose_backward <- the PyTorch posing chain models/deformer/rigid.py:49-71 / 215-236,
 *                        scene/gaussian_model.py:28-32,154-157, models/texture/texture.py:21-38 (no native
 *                        counterpart in the reference)
 *
 * Null pointers select code paths exactly as the reference's nullptr tests do: colors_precomp == NULL -> SH
 * colours are evaluated (forward.cu:241); cov3D_precomp == NULL -> covariance from scales/rotations
 * (forward.cu:205).
 */
#ifndef GSPLAT_B200_H
#define GSPLAT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* status codes (0 = ok; negative = error) */
#define GS_OK 0
#define GS_ERR_BAD_ARG (-1)        /* null/shape problem; mirrors AT_ERROR in rasterize_points.cu:57-59 */
#define GS_ERR_WORKSPACE_SMALL (-2) /* a caller buffer is smaller than gs_*_bytes() says */
#define GS_ERR_CUDA (-3)           /* a CUDA call failed; gs_last_cuda_error() has the code */
#define GS_ERR_NON_RGB (-4)        /* rasterizer_impl.cu:242-245 */
#define GS_ERR_STATIC_LIMIT (-5)   /* l1_loss_forward_static: n exceeds l1_loss_static_max_n() (the built-in scratch limit) */

typedef struct GsForwardArgs {
  int32_t P;             /* number of Gaussians */
  int32_t D;             /* active SH degree */
  int32_t M;             /* SH coefficients per channel held in `shs` ((max_degree+1)^2), 0 if none */
  int32_t width, height; /* image size */
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int32_t prefiltered;
  int32_t debug;           /* !=0: synchronise + check after every stage (CHECK_CUDA, auxiliary.h:166-173) */
  const float* background; /* [3] */
  const float* means3D;    /* [P,3] */
  const float* shs;        /* [P,M,3] or NULL */
  const float* colors_precomp; /* [P,3] or NULL */
  const float* opacities;  /* [P] */
  const float* scales;     /* [P,3] or NULL */
  const float* rotations;  /* [P,4] (w,x,y,z) or NULL */
  const float* cov3D_precomp; /* [P,6] or NULL */
  const float* viewmatrix; /* [16] */
  const float* projmatrix; /* [16] */
  const float* campos;     /* [3] */
  float* out_color;        /* [3,H,W] */
  int32_t* radii;          /* [P] */
  /* Fused opacity pass.  gaussian_renderer/__init__.py:131-142 rasterises the same geometry a
   * second time with colors_precomp = ones and keeps channel 0 (`opacity_image[:1]`).  Non-NULL: the one forward blend
   * carries that channel too and writes it here, bit-identical to the second call.  NULL: plain forward. */
  float* out_opacity;      /* [H,W] or NULL */
} GsForwardArgs;

typedef struct GsBackwardArgs {
  int32_t P, D, M;
  int32_t width, height;
  float tan_fovx, tan_fovy;
  float scale_modifier;
  int32_t debug;
  const float* background;
  const float* means3D;
  const float* shs;
  const float* colors_precomp;
  const float* scales;
  const float* rotations;
  const float* cov3D_precomp;
  const float* viewmatrix;
  const float* projmatrix;
  const float* campos;
  const int32_t* radii;
  const float* dL_dpix;  /* [3,H,W] */
  /* outputs: gs_backward writes EVERY element (culled Gaussians get explicit zeros), so the caller need not zero-fill them
   * as rasterize_points.cu:151-159 does.  The per-pixel sums are accumulated in a 48-byte record per Gaussian inside the geom
   * workspace, which the forward zeroes and the backward zeroes again after use. */
  float* dL_dmean2D;     /* [P,3]  (.z = 0) */
  float* dL_dconic;      /* [P,4]  (2x2; .z = 0) optional: NULL when the caller does not need it (the reference keeps it internal) */
  float* dL_dopacity;    /* [P] */
  float* dL_dcolor;      /* [P,3] */
  float* dL_dmean3D;     /* [P,3] */
  float* dL_dcov3D;      /* [P,6] */
  float* dL_dsh;         /* [P,M,3] or NULL when M == 0 */
  float* dL_dscale;      /* [P,3] */
  float* dL_drot;        /* [P,4] */
  /* upstream gradient of the fused opacity image (forward must have been run with out_opacity); the geometry
   * gradients come out as the SUM of the two backward calls of the reference, in one backward blend */
  const float* dL_dopacity_pix; /* [H,W] or NULL */
} GsBackwardArgs;

/* Workspace sizes.  geom and image are kept for backward (like geomBuffer / imgBuffer); `list` holds the
 * sorted Gaussian index per (Gaussian, tile) pair (like binningBuffer) plus the backward work-item table and the
 * forward's per-pixel checkpoints, and is kept too (gs_backward must be given the same pair_capacity); `scratch`
 * is only used inside gs_forward and may be reused by the next call. */
size_t gs_geom_bytes(int32_t P);
size_t gs_image_bytes(int32_t width, int32_t height);
size_t gs_list_bytes(int64_t pair_capacity, int32_t width, int32_t height);
/* inverse of gs_list_bytes for capacities that are multiples of 4096 (the largest one that fits list_bytes) */
int64_t gs_list_capacity(size_t list_bytes, int32_t width, int32_t height);
size_t gs_scratch_bytes(int32_t P, int32_t width, int32_t height, int64_t pair_capacity);

/* Forward.  Launches everything on `stream` without synchronising (unless args->debug).  The number of
 * (Gaussian, tile) pairs N is produced on the device; when it exceeds pair_capacity the image is NOT valid
 * (the kernels stop at capacity): the caller reads N back with gs_read_num_rendered() (or from
 * num_rendered_pinned_host, filled by an async copy on `stream`) and, if N > pair_capacity, calls gs_forward
 * again with larger `list` / `scratch` buffers.  Returns GS_OK or a negative status. */
int gs_forward(const GsForwardArgs* args, void* geom, size_t geom_bytes, void* image, size_t image_bytes,
               void* list, size_t list_bytes, void* scratch, size_t scratch_bytes, int64_t pair_capacity,
               int64_t* num_rendered_pinned_host, void* stream);

/* Blocking read of N from a geom workspace filled by gs_forward on `stream`. */
int gs_read_num_rendered(const void* geom, int64_t* num_rendered_host, void* stream);

int gs_backward(const GsBackwardArgs* args, const void* geom, size_t geom_bytes, const void* image,
                size_t image_bytes, const void* list, size_t list_bytes, int64_t pair_capacity, void* stream);

/* present[P] (bytes 0/1) = p_view.z > 0.2 */
int gs_mark_visible(int32_t P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                    uint8_t* present, void* stream);

/* Debug accessors (device -> device copies into caller buffers; used by the parity tests).
 * which: 0 ranges uint32[tiles*2] | 1 n_contrib uint32[H*W] | 2 final_T float[H*W] | 3 point_list uint32[N]
 *        4 sorted keys uint64[N] (tile<<32 | depth bits) | 5 means2D float[P*2] | 6 depths float[P]
 *        7 conic_opacity float[P*4] | 8 tiles_touched uint32[P] | 9 rgb float[P*3] | 10 cov3D float[P*6]
 *        11 clamped uint8[P*3] */
int gs_debug_read(int32_t which, const GsForwardArgs* args, const void* geom, const void* image, const void* list,
                  int64_t num_rendered, void* dst, size_t dst_bytes, void* stream);

/* ---- posing (fused rigid LBS + covariance + SH colour) ---- */
typedef struct PoseArgs {
  int32_t P;
  int32_t sh_degree;      /* active degree 0..3 */
  int32_t M;              /* coefficients per channel stored (16 for max degree 3) */
  int32_t cano_view_dir;  /* texture.py:24 */
  float scaling_modifier;
  const float* xyz;       /* [P,3] canonical means */
  const float* rotation;  /* [P,4] raw quaternions (normalised inside, general_utils.py:88-90) */
  const float* scaling;   /* [P,3] log-scales */
  const float* opacity;   /* [P] logits */
  const float* features_dc;   /* [P,1,3] */
  const float* features_rest; /* [P,M-1,3] (may be NULL when M == 1) */
  const float* weights;   /* [P,24] skinning weights */
  const float* bone_transforms; /* [24,4,4] row-major */
  const float* camera_center;   /* [3] */
  /* outputs */
  float* means3D;   /* [P,3] */
  float* cov3D;     /* [P,6] */
  float* colors;    /* [P,3] */
  float* opacity_out; /* [P] sigmoid */
  float* T_fwd;     /* [P,16] or NULL */
  float* rotation_bar; /* [P,9] or NULL */
  const float* view_noise; /* [9] row-major 3x3 or NULL: training-time view augmentation of SH2RGB (texture.py:28-33), applied to the
                            * canonical view direction as dir @ view_noise; only with cano_view_dir */
} PoseArgs;

typedef struct PoseBackwardArgs {
  PoseArgs fwd;             /* same inputs as the forward call (outputs ignored) */
  const float* dL_dmeans3D; /* [P,3] */
  const float* dL_dcov3D;   /* [P,6] */
  const float* dL_dcolors;  /* [P,3] */
  const float* dL_dopacity; /* [P] */
  /* outputs (assigned, not accumulated) */
  float* dL_dxyz;      /* [P,3] */
  float* dL_drotation; /* [P,4] */
  float* dL_dscaling;  /* [P,3] */
  float* dL_dopacity_logit; /* [P] */
  float* dL_dfeatures_dc;   /* [P,1,3] */
  float* dL_dfeatures_rest; /* [P,M-1,3] */
  float* dL_dweights;  /* [P,24] or NULL */
  float* dL_dbone_transforms; /* [24,4,4] or NULL: sum_p w[p,j] dL/dT_fwd[p] (rigid.py:215-236), the gradient that reaches the
                               * per-frame pose correction (models/pose_correction); zeroed by the call, bottom rows stay 0 */
} PoseBackwardArgs;

int pose_forward(const PoseArgs* args, void* stream);
int pose_backward(const PoseBackwardArgs* args, void* stream);

/* smpl_nn skinning weights: index of the nearest of the V (6890) SMPL vertices per Gaussian and the gather of that vertex's
 * 24 skinning weights.  Replaces SMPLNN.query_weights, models/deformer/rigid.py:41-47 (pytorch3d knn_points K = 1, P x V).
 * Exact brute force in FP32 (dx*dx + dy*dy + dz*dz), the first minimum wins.  `idx_out` [P] and `weights_out` [P,24] may each
 * be NULL; `vert_weights` [V,24] is needed for weights_out. */
int smpl_nn_forward(int32_t P, const float* xyz, int32_t V, const float* verts, const float* vert_weights, int32_t* idx_out,
                    float* weights_out, void* stream);

/* ---- non-rigid deformation: multiresolution hash grid + conditional MLP (forward) ----
 * Replaces HashGridwithMLP.forward, models/deformer/non_rigid.py:226-286: AABB.normalize (utils/dataset_utils.py:66-76),
 * HashGrid (models/network_utils.py:329-344; tiny-cuda-nn's multiresolution hash encoding, Mueller et al. 2022 -- the
 * dependency is not vendored, so the published algorithm is restated: parity unpinned), VanillaCondMLP
 * (models/network_utils.py:182-249: lin0 (F+C)->H, lin1, lin2 H->H, lin3 H->n_out, LeakyReLU(0.01) between) and the
 * delta application (xyz += d[0:3]; scaling += d[3:6]; rotation = (1, d[7:10]) * rotation, Hamilton product).
 * The four linear layers run on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM) with a 2-way BF16
 * split of activations and weights (hi*hi + lo*hi + hi*lo, FP32 accumulate): ~2^-16 relative error per product.
 * The per-frame condition (pose feature, C floats) is constant over the Gaussians and is folded into the first bias. */
#define NR_MAX_LEVELS 32
typedef struct NonRigidArgs {
  int32_t P;
  int32_t n_levels;        /* L (16) */
  int32_t n_feat;          /* features per level: must be 2 */
  int32_t hidden;          /* H: must be 128 */
  int32_t n_cond;          /* C (144) */
  int32_t n_out;           /* 10 + feature_dim, <= 32 */
  uint32_t level_offset[NR_MAX_LEVELS + 1]; /* entry offsets of the levels in `table` (entries of n_feat floats) */
  uint32_t level_res[NR_MAX_LEVELS];        /* grid resolution per level */
  float level_scale[NR_MAX_LEVELS];         /* base_resolution * per_level_scale^l - 1 */
  const float* aabb_min;   /* [3] */
  const float* aabb_max;   /* [3] */
  const float* xyz;        /* [P,3] canonical means */
  const float* scaling;    /* [P,3] */
  const float* rotation;   /* [P,4] */
  const float* table;      /* [level_offset[L], n_feat] hash-grid parameters (FP32) */
  const float* cond;       /* [C] per-frame pose feature */
  const float* W0; const float* b0;  /* [H, L*n_feat + C], [H] */
  const float* W1; const float* b1;  /* [H, H], [H] */
  const float* W2; const float* b2;
  const float* W3; const float* b3;  /* [n_out, H], [n_out] */
  /* outputs */
  float* xyz_out;          /* [P,3] */
  float* scaling_out;      /* [P,3] */
  float* rotation_out;     /* [P,4] */
  float* deltas;           /* [P,n_out] raw MLP output (delta_rot[:,0] left as produced), or NULL */
  float* features;         /* [P, L*n_feat] hash-grid encoding, or NULL (debug / tests) */
} NonRigidArgs;

size_t nonrigid_workspace_bytes(const NonRigidArgs* args);
int nonrigid_forward(const NonRigidArgs* args, void* workspace, size_t workspace_bytes, void* stream);

/* ---- non-rigid deformation, training: forward that keeps its activations + hand-written backward ----
 * Replaces the autograd backward of HashGridwithMLP.forward (models/deformer/non_rigid.py:226-300): VanillaCondMLP
 * (models/network_utils.py:182-249: dX = G.W, dW = G^T.H, db = column sums, LeakyReLU' from the saved activations),
 * HashGrid (models/network_utils.py:329-344 -> tiny-cuda-nn: dL/dtable scattered with red.global.add over the 8 corners of
 * every level, dL/dx through the trilinear weights) and the delta application (identity for xyz / scaling, Hamilton product
 * for the rotation).  nonrigid_forward_train = nonrigid_forward + `saved` (nonrigid_saved_bytes(P) bytes: per tile of 128
 * Gaussians the input features and the three hidden activations in the BF16 hi/lo operand layout of the tensor cores).
 * nonrigid_backward: `fwd` as in the forward call with fwd.deltas = the forward's raw MLP output (read here);  upstream
 * gradients may be NULL (= zero).  Per-Gaussian outputs are assigned; parameter gradients (table, W*, b*, cond) are zeroed
 * by the call and then accumulated.  All GEMMs run on tcgen05 (dW contracts over the tile's 128 Gaussians with MN-major
 * descriptors on the very buffers dX reads K-major), 2-way BF16 split, FP32 accumulate in TMEM. */
typedef struct NonRigidBackwardArgs {
  NonRigidArgs fwd;
  const float* dL_dxyz_out;       /* [P,3] or NULL */
  const float* dL_dscaling_out;   /* [P,3] or NULL */
  const float* dL_drotation_out;  /* [P,4] or NULL */
  const float* dL_ddeltas;        /* [P,n_out] or NULL (regularisers on the raw deltas, non_rigid.py:283-298) */
  float* dL_dxyz;                 /* [P,3] */
  float* dL_dscaling;             /* [P,3] */
  float* dL_drotation;            /* [P,4] */
  float* dL_dcond;                /* [C] or NULL */
  float* dL_dtable;               /* [level_offset[L], 2] */
  float* dL_dW0; float* dL_db0;   /* [H, L*n_feat + C], [H] */
  float* dL_dW1; float* dL_db1;
  float* dL_dW2; float* dL_db2;
  float* dL_dW3; float* dL_db3;   /* [n_out, H], [n_out] */
} NonRigidBackwardArgs;
size_t nonrigid_saved_bytes(int32_t P);
int nonrigid_forward_train(const NonRigidArgs* args, void* workspace, size_t workspace_bytes, void* saved, size_t saved_bytes,
                           void* stream);
int nonrigid_backward(const NonRigidBackwardArgs* args, void* workspace, size_t workspace_bytes, const void* saved,
                      size_t saved_bytes, void* stream);

/* ---- skinning field: per-Gaussian LBS weights from the skinning MLP (forward) ----
 * Replaces SkinningField.get_forward_transform's weight part, models/deformer/rigid.py:160-171,189-196 (non-distilled
 * path): AABB.normalize(sym=True) -> VanillaCondMLP(3 -> H x4 -> n_out) (models/network_utils.py:182-249,251-257;
 * configs/rigid/skinning_field.yaml) -> hierarchical_softmax over the kinematic tree when n_out == 25
 * (rigid.py:85-129; logit 24 gates joint 9 -> {12,13,14}) or a plain softmax when n_out == 24.  The [P,24] result is the
 * `weights` input of pose_forward.  Same tensor-core building block as nonrigid_forward (tcgen05, 2-way BF16 split). */
typedef struct SkinArgs {
  int32_t P;
  int32_t hidden;          /* must be 128 */
  int32_t n_out;           /* 25 (hierarchical softmax) or 24 (softmax) */
  int32_t pad_;
  const float* aabb_min;   /* [3] */
  const float* aabb_max;   /* [3] */
  const float* xyz;        /* [P,3] canonical (non-rigidly displaced) means */
  const float* W[5];       /* lin0 [H,3], lin1..lin3 [H,H], lin4 [n_out,H] */
  const float* b[5];
  float* weights;          /* [P,24] */
  float* logits;           /* [P,n_out] or NULL */
} SkinArgs;
size_t skinning_workspace_bytes(const SkinArgs* args);
int skinning_forward(const SkinArgs* args, void* workspace, size_t workspace_bytes, void* stream);

/* ---- skinning field, training: forward that keeps its activations + hand-written backward ----
 * Replaces the autograd backward of the weight part of SkinningField (rigid.py:85-129 hierarchical_softmax, :160-171 softmax,
 * :189-196 lbs_network): adjoint of the (hierarchical) softmax per Gaussian, then the MLP backward on tcgen05 as in
 * nonrigid_backward; dL/dxyz through AABB.normalize.  `fwd.logits` must hold the forward's logits [P,n_out] (read here).
 * Parameter gradients are zeroed by the call and accumulated; dL_dxyz is assigned. */
typedef struct SkinBackwardArgs {
  SkinArgs fwd;
  const float* dL_dweights;  /* [P,24] */
  float* dL_dxyz;            /* [P,3] */
  float* dL_dW[5];
  float* dL_db[5];
} SkinBackwardArgs;
size_t skinning_saved_bytes(int32_t P);
int skinning_forward_train(const SkinArgs* args, void* workspace, size_t workspace_bytes, void* saved, size_t saved_bytes,
                           void* stream);
int skinning_backward(const SkinBackwardArgs* args, void* workspace, size_t workspace_bytes, const void* saved,
                      size_t saved_bytes, void* stream);

/* ---- colour MLP texture (forward) ----
 * Replaces ColorMLP.forward, models/texture/texture.py:40-125 (registry key "mlp", the repository's default texture,
 * configs/texture/shallow_mlp.yaml): compose_input = [ per-Gaussian feature (feature_dim) | eval_sh_bases(sh_degree,
 * view dir)[1:] (utils/sh_utils.py:115-161) | non_rigid_feature (non_rigid.py:87,119-120) | per-frame latent code ],
 * view dir = (xyz - camera_center), rotated into the canonical frame by T_fwd[:, :3, :3]^T when cano_view_dir
 * (texture.py:90-93), times the optional training-time view-noise rotation (:94-99), divided by (norm + 1e-12) (:100);
 * VanillaCondMLP(d_in -> H -> H -> 3) with LeakyReLU(0.01) (models/network_utils.py:182-249) and a sigmoid (:121-123).
 * The latent code is the same for every Gaussian of a frame and is folded into the first bias; the three linear layers
 * run on tcgen05 tensor cores (64-wide layers, accumulators in TMEM, 2-way BF16 split as in nonrigid_forward) in a
 * persistent kernel whose weights stay resident in shared memory. */
typedef struct ColorArgs {
  int32_t P;
  int32_t feature_dim;     /* F (32) */
  int32_t sh_degree;       /* 0..3; contributes (sh_degree+1)^2 - 1 inputs */
  int32_t non_rigid_dim;   /* R (16), 0 = none */
  int32_t latent_dim;      /* Z (16), 0 = none */
  int32_t hidden;          /* H: must be 64 */
  int32_t cano_view_dir;   /* texture.py:90 */
  int32_t pad_;
  const float* features;   /* [P,F] gaussians.get_features.squeeze(-1) */
  const float* xyz;        /* [P,3] deformed means (gaussians.get_xyz of the posed set) */
  const float* T_fwd;      /* [P,16] row-major forward transforms (needed when cano_view_dir), else NULL */
  const float* camera_center; /* [3] */
  const float* view_noise; /* [9] row-major 3x3, dir <- dir @ view_noise (training augmentation), or NULL */
  const float* non_rigid_feature; /* [P,R] or NULL when R == 0 */
  const float* latent;     /* [Z] latent code of the frame, or NULL when Z == 0 */
  const float* W0; const float* b0;  /* [H, F + S + R + Z], [H]   (S = (sh_degree+1)^2 - 1; F + S + R <= 64) */
  const float* W1; const float* b1;  /* [H, H], [H] */
  const float* W2; const float* b2;  /* [3, H], [3] */
  float* colors;           /* [P,3] sigmoid(MLP) = colors_precomp of the rasterizer */
  float* inputs;           /* [P, F + S + R] composed per-Gaussian input (debug / tests), or NULL */
} ColorArgs;
size_t color_workspace_bytes(const ColorArgs* args);
int color_forward(const ColorArgs* args, void* workspace, size_t workspace_bytes, void* stream);

/* ---- colour MLP texture, training: forward that keeps its activations + hand-written backward ----
 * Replaces the autograd backward of ColorMLP.forward (models/texture/texture.py:74-125): sigmoid', the three linear layers
 * on tcgen05 (dX = G.W, dW = G^T.A over the tile's 128 Gaussians, MN-major descriptors; see nonrigid_backward), LeakyReLU'
 * from the saved activations, then the adjoint of compose_input: per-Gaussian feature, SH basis -> view direction ->
 * (view noise) -> canonical rotation T_fwd[:3,:3]^T -> xyz, non-rigid feature; the latent code through the folded bias.
 * `fwd.colors` must hold the forward's colours (read here).  Per-Gaussian outputs (each may be NULL) are assigned; parameter
 * gradients are zeroed by the call and accumulated. */
typedef struct ColorBackwardArgs {
  ColorArgs fwd;
  const float* dL_dcolors;        /* [P,3] */
  float* dL_dfeatures;            /* [P,F] or NULL */
  float* dL_dxyz;                 /* [P,3] or NULL */
  float* dL_dT_fwd;               /* [P,16] or NULL */
  float* dL_dnon_rigid_feature;   /* [P,R] or NULL */
  float* dL_dlatent;              /* [Z] or NULL */
  float* dL_dW0; float* dL_db0;
  float* dL_dW1; float* dL_db1;
  float* dL_dW2; float* dL_db2;
} ColorBackwardArgs;
size_t color_saved_bytes(int32_t P);
int color_forward_train(const ColorArgs* args, void* workspace, size_t workspace_bytes, void* saved, size_t saved_bytes,
                        void* stream);
int color_backward(const ColorBackwardArgs* args, void* workspace, size_t workspace_bytes, const void* saved, size_t saved_bytes,
                   void* stream);

/* ---- optimiser-side per-Gaussian work of a training iteration ----
 * adam_step: torch.optim.Adam as configured by GaussianModel.training_setup (scene/gaussian_model.py:201-221: six
 * parameter groups with their own learning rates, betas (0.9, 0.999), eps 1e-15, no weight decay / amsgrad), applied by
 * scene.optimize at train.py:231-232, as ONE launch over a flat parameter slab.  The slab layout is the flat gradient
 * slab of pose_backward (xyz 3P | rotation 4P | scaling 3P | opacity P | f_dc 3P | f_rest 3(M-1)P), which is also the
 * buffer the data-parallel all-reduce runs over; `group_end` are the exclusive end offsets of the groups, `lrs` a
 * DEVICE array (the xyz learning rate is rescheduled every iteration, gaussian_model.py:223-229; the launch can sit in a
 * CUDA graph).  `state` = adam_state_bytes() zero-initialised device bytes holding the step counter.  Operation order is
 * torch's _single_tensor_adam. */
#define ADAM_MAX_GROUPS 8
typedef struct AdamArgs {
  int64_t n;                 /* scalars in the slab */
  int32_t n_groups;
  int32_t pad_;
  int64_t group_end[ADAM_MAX_GROUPS];
  const float* lrs;          /* device [n_groups] */
  float beta1, beta2, eps;
  float pad2_;
  float* params;             /* [n] updated in place */
  const float* grads;        /* [n] */
  float* exp_avg;            /* [n] */
  float* exp_avg_sq;         /* [n] */
  void* state;               /* device, adam_state_bytes() */
} AdamArgs;
size_t adam_state_bytes(void);
int adam_step(const AdamArgs* args, void* stream);

/* densify_stats: for every visible Gaussian (radii > 0): max_radii2D = max(max_radii2D, radii) (train.py:219),
 * xyz_gradient_accum += ||viewspace_grad[:, :2]||, denom += 1 (scene/gaussian_model.py:464-466), one pass. */
int densify_stats(int32_t P, const int32_t* radii, const float* viewspace_grad /*[P,3]*/, float* max_radii2D /*[P]*/,
                  float* xyz_gradient_accum /*[P]*/, float* denom /*[P]*/, void* stream);

/* ---- hierarchical pose encoder (forward): the per-frame condition of the non-rigid MLP ----
 * Replaces HierarchicalPoseEncoder.forward, models/network_utils.py:152-178 (constructed from cfg.pose_encoder at
 * models/deformer/non_rigid.py:33,137,206; num_joints 24, dim_per_joint 6, out_dim -1 -> identity output layer):
 * global feature = layer_0([rots | Jtrs]); per joint, in kinematic-tree order, out[j] = Linear(ReLU(Linear([rot_j | Jtr_j |
 * bone length | out[parent] (global feature for the root)]))); result = concatenation [24*6].  One launch. */
typedef struct PoseEncoderArgs {
  int32_t num_joints;      /* must be 24 */
  int32_t dim_per_joint;   /* must be 6 */
  int32_t rel_joints;      /* network_utils.py:155-159 */
  int32_t pad_;
  int32_t parents[24];     /* ktree_parents (:128-129), parents[j] < j, -1 for the root */
  const float* rots;       /* [24,9] */
  const float* Jtrs;       /* [24,3] */
  const float* W0; const float* b0;  /* layer_0: [6, 288], [6] */
  const float* Wa; const float* ba;  /* layers[j][0]: [24,19,19], [24,19] */
  const float* Wb; const float* bb;  /* layers[j][2]: [24,6,19], [24,6] */
  float* out;              /* [144] */
} PoseEncoderArgs;
int pose_encoder_forward(const PoseEncoderArgs* args, void* stream);
/* Backward of pose_encoder_forward (the reference: autograd over the ~150 nn.Linear nodes of network_utils.py:152-178) in one
 * launch: re-evaluates the forward from `fwd`'s inputs (fwd.out is not read) and writes every element of every gradient (no
 * clearing needed; deterministic).  dL_drots / dL_dJtrs may be NULL; with rel_joints the reference forms the relative joints
 * under no_grad (:155-159), so dL_dJtrs is all zeros. */
typedef struct PoseEncoderBackwardArgs {
  PoseEncoderArgs fwd;
  const float* dL_dout;    /* [144] */
  float* dL_drots;         /* [24,9] or NULL */
  float* dL_dJtrs;         /* [24,3] or NULL */
  float* dL_dW0; float* dL_db0;  /* [6,288], [6] */
  float* dL_dWa; float* dL_dba;  /* [24,19,19], [24,19] */
  float* dL_dWb; float* dL_dbb;  /* [24,6,19], [24,6] */
} PoseEncoderBackwardArgs;
int pose_encoder_backward(const PoseEncoderBackwardArgs* args, void* stream);

/* ---- data-parallel training: in-place sum of the per-rank gradient slabs over NVLink / NVSwitch peer memory ----
 * The reference has no multi-GPU path; this is the one collective the data-parallel step needs, as one kernel
 * behind pose_backward instead of an NCCL call.  The slab must live in symmetric memory (same allocation on every rank,
 * peers' addresses exchanged once): `peer_ptrs[world]` / `signal_pads[world]` are DEVICE arrays of the peers' buffer and
 * signal-pad addresses (zero-initialised pads of >= (256 + world) * 4 bytes; words [256, 256 + world) belong to this kernel and
 * hold barrier epochs that only grow -- the first 1 KB is left to torch's own symmetric-memory barriers), `multicast_ptr` the NVLS multicast address of
 * the buffer (required: without NVLS multicast support the caller uses NCCL).  Every rank must call it with the same
 * n / blocks; ranks meet at in-kernel barriers.  `error_flag`: 16 device u32; [0] is set if a barrier gives up after tens of
 * seconds (a peer never arrived): the slab is then left UNTOUCHED, this rank's flags are cleared, and every later call is a
 * no-op until the caller resets [0] (after a timeout the ranks' epoch counters may disagree: re-create the slab); [2:10] receive
 * four 64-bit %globaltimer stamps of block 0 (start, after the first barrier, after the data phase, end); [12:15] are the kernel's
 * own control words (gate, block counter, barrier epoch) and must start at zero. */
typedef struct SlabAllReduceArgs {
  int64_t n;               /* floats, multiple of 4 */
  int32_t rank, world;
  void* const* peer_ptrs;
  uint32_t* const* signal_pads;
  void* multicast_ptr;
  int64_t byte_offset;     /* of the slab inside the symmetric buffer (multiple of 16) */
  int32_t blocks;
  int32_t pad_;
  uint32_t* error_flag;
} SlabAllReduceArgs;
int slab_allreduce(const SlabAllReduceArgs* args, void* stream);

/* ---- photometric L1 loss of the training step, forward + per-pixel gradient in one pass ----
 * Replaces l1_loss (utils/loss_utils.py:19-20: torch.abs(network_output - gt).mean(), called at train.py:121) and its
 * autograd backward.  `loss` [1] = mean |image - target|; `dL_dimage` [n] = sign(image - target) / n (sign(0) = 0), i.e.
 * the gradient for dL/dloss = 1; l1_loss_backward scales it by the device scalar dL_dloss.  Deterministic summation.
 * `workspace`: l1_loss_workspace_bytes() bytes, zero-filled once by the caller before the first use.  Pointers 16-byte
 * aligned. */
size_t l1_loss_workspace_bytes(void);
int l1_loss_forward(const float* image, const float* target, int64_t n, float* loss, float* dL_dimage, void* workspace,
                    size_t workspace_bytes, void* stream);
int l1_loss_backward(const float* dL_dimage_unit, const float* dL_dloss, int64_t n, float* dL_dimage, void* stream);
/* l1_loss_forward without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same inputs, outputs and results as l1_loss_forward; the kernel's scratch (per-block partial sums + the
 * completion ticket, 2372 bytes) is a file-scope `static __device__` array inside the library, which every GPU that loads the
 * library holds until exit.  Limit: n <= l1_loss_static_max_n() = 2^26 elements, beyond that GS_ERR_STATIC_LIMIT and nothing is
 * launched.  All calls on a device share that one array (it keeps the last call's contents; the entry point clears the ticket
 * word on `stream` before each launch): calls that OVERLAP on the GPU -- different streams or host threads on one device -- race
 * and must be serialised by the caller.  Calls no cudaMalloc / cudaFree / allocating routine (cudaGetSymbolAddress once per
 * device, cudaMemsetAsync of 4 bytes, one kernel launch). */
int64_t l1_loss_static_max_n(void);
int l1_loss_forward_static(const float* image, const float* target, int64_t n, float* loss, float* dL_dimage, void* stream);

/* Bench support (no reference counterpart): per-stage CUDA-event timing of the last forward/backward call, and an
 * FP32 FMA throughput probe (8 independent FMA chains per thread, `iters` rounds: 16*iters flop per thread). */
void gs_profile_enable(int on);
int gs_profile_read(float* ms_host /* [7]: preprocess, binning, blend fwd, blend bwd, preprocess bwd, pose fwd, pose bwd */);
int gs_fma_probe(int blocks, int threads, int iters, float* out_dev, void* stream);

/* exp() of the blend kernels (cuda_rasterizer/forward.cu:344 `exp(power)`, backward.cu:483).  mode 0 (default): CUDA's expf,
 * the instruction sequence nvcc emits for the reference, so that the alpha >= 1/255 and T < 1e-4 decisions -- n_contrib -- are
 * the reference build's bit for bit.  mode 1 ("oracle mode"): a fixed FMA polynomial (<= 1 ulp) that oracle/gs_oracle.c
 * reproduces bit for bit on the CPU, used by the parity tests that compare integer artefacts AND image bits with the
 * oracle.  Process-wide; returns the previous mode.  GS_EXP_MODE=1 in the environment selects mode 1 initially. */
int gs_set_exp_mode(int mode);
/* counters_host[0] = number of (pixel, entry) pairs blended in the frame of `geom` (E_bwd of SURVEY.md 8(d)), counted by
 * gs_backward -- the one word of the forward's workspaces the backward writes: it adds to this profiling counter in the
 * header, so the value is E_bwd times the number of backward calls since the forward;
 * counters_host[1] = forward work items (tile, 512-entry chunk) of that frame.  Synchronises `stream`. */
int gs_read_counters(const void* geom, uint64_t* counters_host, void* stream);

/* sizeof of the ABI structs: 0 GsForwardArgs, 1 GsBackwardArgs, 2 PoseArgs, 3 PoseBackwardArgs, 4 NonRigidArgs, 5 SkinArgs,
 * 6 ColorArgs, 7 AdamArgs, 8 PoseEncoderArgs, 9 SlabAllReduceArgs, 10 NonRigidBackwardArgs, 11 SkinBackwardArgs,
 * 12 ColorBackwardArgs, 13 PoseEncoderBackwardArgs */
size_t gs_sizeof(int32_t which);
int gs_last_cuda_error(void);
const char* gs_version(void);

#ifdef __cplusplus
}
#endif
#endif
