/*
 * moss_splat.h -- C ABI of the B200-native differentiable Gaussian-splat rasterizer.
 *
 * Drop-in boundary for MOSS's vendored `diff_gaussian_rasterization` (DGR =
 * /root/reference/submodules/diff-gaussian-rasterization).  Each entry point names the
 * reference interface it replaces.  Plain pointers and sizes only: every pointer is a
 * DEVICE pointer on the current CUDA device unless stated otherwise; the caller owns
 * every buffer (outputs, gradients and the three workspaces).  No C++ exceptions cross
 * this boundary: every call returns a MOSS_* status code.
 *
 * All launches go to the `stream` argument (a cudaStream_t passed as void*), never to
 * the legacy default stream, so calls are CUDA-graph capturable when MOSS_FLAG_ASYNC
 * is set (no host synchronisation at all in that mode).
 */
#ifndef MOSS_SPLAT_H_INCLUDED
#define MOSS_SPLAT_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MOSS_ABI_VERSION 10

/* status codes */
#define MOSS_OK                  0
#define MOSS_ERR_BINNING_OVERFLOW 1  /* num_rendered > binning capacity: outputs invalid; grow the
                                        binning workspace to >= moss_binning_workspace_bytes(*num_rendered)
                                        and call again with MOSS_FLAG_REBIN */
#define MOSS_ERR_INVALID_ARGUMENT 2
#define MOSS_ERR_WORKSPACE_TOO_SMALL 3
#define MOSS_ERR_CUDA            4   /* a CUDA runtime error; see moss_last_cuda_error() */
#define MOSS_ERR_STATIC_SCRATCH_LIMIT 5  /* moss_image_losses_static(): image larger than the library's static scratch */

/* flags */
#define MOSS_FLAG_ASYNC       1u  /* never block the host: *num_rendered is set to -1; the instance count and an
                                     overflow flag stay on the device (moss_export_state which=MOSS_STATE_COUNTERS).
                                     Required under CUDA-graph capture. */
#define MOSS_FLAG_REBIN       2u  /* geometry + tile counts in the workspaces are valid (previous call returned
                                     MOSS_ERR_BINNING_OVERFLOW): redo only scatter + sort + blend */
#define MOSS_FLAG_EXPORT_KEYS 4u  /* also materialise the sorted 64-bit (tile<<32 | depth_bits) key list (parity
                                     checks against DGR/cuda_rasterizer/rasterizer_impl.cu:305-310); needs the
                                     larger binning workspace moss_binning_workspace_bytes_ex(R, 1) */
#define MOSS_FLAG_DEBUG       8u  /* synchronise + check after every launch (DGR/cuda_rasterizer/auxiliary.h:166-173) */
#define MOSS_FLAG_DETERMINISTIC 16u /* backward: fixed-order gradient reduction (bit-reproducible run to run) */
#define MOSS_FLAG_SPLIT_SHIFT  8   /* bits 8-9: blend block shape; 0 = automatic, 1 = whole tile (8 warps), 2 = half
                                     tile (4 warps), 3 = quarter tile (2 warps).  Results do not depend on it. */
#define MOSS_FLAG_SPLIT_MASK   (3u << MOSS_FLAG_SPLIT_SHIFT)
#define MOSS_FLAG_NO_CULL     32u /* forward: disable the per-patch culling of instances (every instance is offered to
                                       every pixel of its tile, as DGR forward.cu:330-365 does); results are identical
                                       by construction -- the flag exists so tests can prove it at full sizes */

/* Workspace sizes.  Replace the reference's `required<GeometryState|ImageState|BinningState>()`
 * (DGR/cuda_rasterizer/rasterizer_impl.h:64-72) and its three grow-callbacks
 * (DGR/cuda_rasterizer/rasterizer.h:33-35, DGR/rasterize_points.cu:27-33): the caller allocates. */
size_t moss_geom_workspace_bytes(int P);
size_t moss_image_workspace_bytes(int W, int H);
size_t moss_binning_workspace_bytes(long long capacity);
size_t moss_binning_workspace_bytes_ex(long long capacity, int export_keys);
size_t moss_ckpt_workspace_bytes(long long capacity, int W, int H);
/* largest instance count a binning workspace of `bytes` can hold */
long long moss_binning_capacity(size_t bytes, int export_keys);

/* Per-view rasterization inputs (GaussianRasterizationSettings,
 * DGR/diff_gaussian_rasterization/__init__.py:160-172, plus the tensors of
 * GaussianRasterizer.forward :190-223).  Exactly one of shs / colors_precomp and exactly one of
 * (scales, rotations) / cov3D_precomp must be non-NULL. */
typedef struct MossScene {
    int P;                       /* number of Gaussians */
    int D;                       /* active SH degree 0..3 */
    int M;                       /* SH coefficient stride (16 for degree-3 storage); 0 if shs == NULL */
    int W, H;                    /* image size */
    const float* bg;             /* [3] */
    const float* means3D;        /* [P,3] */
    const float* shs;            /* [P,M,3] or NULL */
    const float* colors_precomp; /* [P,3] or NULL */
    const float* opacities;      /* [P] */
    const float* scales;         /* [P,3] or NULL */
    const float* rotations;      /* [P,4] (r,x,y,z), used un-normalised as in DGR forward.cu:127, or NULL */
    const float* cov3D_precomp;  /* [P,6] xx,xy,xz,yy,yz,zz or NULL */
    const float* viewmatrix;     /* [16], W2C transposed (row-major tensor), DGR auxiliary.h:58-77 */
    const float* projmatrix;     /* [16] */
    const float* campos;         /* [3] */
    float scale_modifier;
    float tanfovx, tanfovy;
    int prefiltered;
    void* geom_ws;   size_t geom_bytes;    /* >= moss_geom_workspace_bytes(P) */
    void* image_ws;  size_t image_bytes;   /* >= moss_image_workspace_bytes(W,H) */
    void* binning_ws; size_t binning_bytes; /* capacity derived with moss_binning_capacity() */
    void* ckpt_ws;   size_t ckpt_bytes;    /* optional (NULL/0 = off), >= moss_ckpt_workspace_bytes(capacity, W, H): the
                                              forward stores per-pixel blend state every 128 instances of a tile so that
                                              the backward walks a deep tile as independent 128-instance segments
                                              (shorter critical path at ZJU/MonoCap shapes).  Forward outputs are
                                              unaffected; gradients agree with the unsegmented walk to fp32 rounding.
                                              Must be the same buffer in the forward and the backward call. */
} MossScene;

/* One camera of a batched call: the per-view fields of GaussianRasterizationSettings
 * (DGR/diff_gaussian_rasterization/__init__.py:160-172).  Image size, SH degree, scale modifier and `prefiltered` are
 * shared by the batch and come from the MossScene. */
#define MOSS_MAX_VIEWS 32
typedef struct MossView {
    const float* viewmatrix;     /* [16] device */
    const float* projmatrix;     /* [16] device */
    const float* campos;         /* [3]  device (may be NULL with colors_precomp) */
    const float* bg;             /* [3]  device */
    float tanfovx, tanfovy;
} MossView;

/* Forward.  Replaces CudaRasterizer::Rasterizer::forward (DGR/cuda_rasterizer/rasterizer.h:33-60,
 * rasterizer_impl.cu:198-341) and the torch glue RasterizeGaussiansCUDA (DGR/rasterize_points.cu:35-119).
 * out_color [3,H,W], out_depth [H,W], out_alpha [H,W], radii int32[P] are fully written.
 * *num_rendered (HOST pointer, may be NULL) receives the instance count R. */
int moss_rasterize_forward(void* stream, const MossScene* scene, unsigned flags,
                           float* out_color, float* out_depth, float* out_alpha, int* radii,
                           int* num_rendered);

/* Backward.  Replaces CudaRasterizer::Rasterizer::backward (DGR/cuda_rasterizer/rasterizer.h:62-98,
 * rasterizer_impl.cu:345-447) and RasterizeGaussiansBackwardCUDA (DGR/rasterize_points.cu:121-206).
 * The workspaces must be the ones the forward call filled.  dL_ddepth and dL_dalpha may be NULL (no gradient
 * reaches that output: the same arithmetic as an all-zero array, without the array).
 * EVERY output is fully written (rows of culled Gaussians with zeros): dL_dmean2D [P,3], dL_dconic [P,4],
 * dL_dopacity [P], dL_dcolor [P,3], dL_dmean3D [P,3], dL_dcov3D [P,6], dL_dsh [P,M,3] (if shs), dL_dscale [P,3]
 * and dL_drot [P,4] (if scales).  Nothing needs zero-filling beforehand (the reference zero-fills nine tensors,
 * DGR/rasterize_points.cu:158-166): the atomically accumulated blend gradients live in a scratch area of the
 * geometry workspace that the forward clears and the backward leaves cleared, so the backward may be repeated
 * on the same forward state. */
int moss_rasterize_backward(void* stream, const MossScene* scene, unsigned flags, const int* radii,
                            const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
                            float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                            float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot);

/* Batched multi-view calls (no reference counterpart: MOSS renders one view per iteration, train_ZJU.py:93-100; SURVEY.md
 * section 7 step 7 / BASELINE.json config 4).  ONE launch sequence renders `V` (1..MOSS_MAX_VIEWS) cameras of the SAME
 * Gaussians: the view index is folded into the tile index (tile' = v*T + tile) and into the Gaussian index of the
 * per-view state (id' = v*P + i), so every view's outputs -- radii, sorted lists, ranges, image -- are bit-identical to V
 * separate single-view calls, while the blend kernels see V times the busy tiles (a single 512x512 body view fills
 * ~1/4 of the GPU).  `scene` supplies the Gaussians, image size and workspaces (its viewmatrix / projmatrix / campos /
 * bg / tanfov fields are ignored); workspaces are sized with the *_batch queries; the binning capacity counts the
 * instances of all views together.  `views` is a HOST array.
 *   forward : out_color [V,3,H,W], out_depth [V,H,W], out_alpha [V,H,W], radii int32 [V,P]; *num_rendered = total.
 *   backward: dL_dpix [V,3,H,W], dL_ddepth / dL_dalpha [V,H,W] or NULL.  Per view: dL_dmean2D [V,P,3] and (optional,
 *             NULL to skip) dL_dconic [V,P,4].  SUMMED over the views, in view order (reproducible): dL_dopacity [P],
 *             dL_dcolor [P,3], dL_dmean3D [P,3], dL_dcov3D [P,6], dL_dsh [P,M,3], dL_dscale [P,3], dL_drot [P,4].
 *             `tmp_ws` (>= moss_batch_grad_workspace_bytes, only needed for V > 1) holds the per-view rows. */
size_t moss_geom_workspace_bytes_batch(int P, int V);
size_t moss_image_workspace_bytes_batch(int W, int H, int V);
size_t moss_ckpt_workspace_bytes_batch(long long capacity, int W, int H, int V);
size_t moss_batch_grad_workspace_bytes(int P, int M, int V);
int moss_rasterize_forward_batch(void* stream, const MossScene* scene, int V, const MossView* views, unsigned flags,
                                 float* out_color, float* out_depth, float* out_alpha, int* radii, int* num_rendered);
int moss_rasterize_backward_batch(void* stream, const MossScene* scene, int V, const MossView* views, unsigned flags,
                                  const int* radii, const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
                                  float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                                  float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                  void* tmp_ws, size_t tmp_bytes);
size_t moss_sizeof_view(void);

/* markVisible.  Replaces CudaRasterizer::Rasterizer::markVisible (DGR/cuda_rasterizer/rasterizer.h:26-31,
 * rasterizer_impl.cu:54-66,141-153).  present: uint8[P]. */
int moss_mark_visible(void* stream, int P, const float* means3D, const float* viewmatrix,
                      const float* projmatrix, uint8_t* present);

/* Fused skinning + covariance front end (optional entry points).  Replaces, per render()
 * call, the per-Gaussian half of GaussianModel.coarse_deform_c2source (scene/gaussian_model.py:829-846, 857-868,
 * 889-921) and GaussianModel.get_covariance(scaling_modifier, transforms) (scene/gaussian_model.py:37-44,168-169;
 * utils/general_utils.py:62-118) as gaussian_renderer/__init__.py:57-85 calls them: ONE kernel instead of ~60 torch
 * launches.  The per-frame, Gaussian-independent tables (joint transforms A, blend-shape offsets) are inputs. */
typedef struct MossSkin {
    int P;                      /* Gaussians */
    int V;                      /* rows of the per-vertex tables (6890 for SMPL) */
    const float* xyz;           /* [P,3] canonical centres (pc.get_xyz) */
    const int*   vert_ids;      /* [P] nearest SMPL vertex, 0 <= id < V (the KNN result, scene/gaussian_model.py:827) */
    const float* weights;       /* [V,24] SMPL_NEUTRAL['weights'], 16-byte aligned */
    const float* lbs_delta;     /* [P,24] or NULL: bw = softmax(log(w + 1e-9) + delta) (:831-833), 16-byte aligned */
    const float* A_big;         /* [24,16] joint transforms of the big pose, row-major 4x4 (:838) */
    const float* A_tgt;         /* [24,16] joint transforms of the target pose (:895) */
    const float* poseoff_big;   /* [V,3] or NULL (:849-857) */
    const float* shapeoff;      /* [V,3] or NULL (:862-866) */
    const float* poseoff_tgt;   /* [V,3] or NULL (:872-889) */
    const float* Rg;            /* [9] params['R'] */
    const float* Th;            /* [3] params['Th'] */
    const float* scaling;       /* [P,3] pc.get_scaling, or NULL: no covariance */
    const float* rotation;      /* [P,4] pc._rotation (raw; normalised inside like build_rotation) or NULL */
    float scale_modifier;
} MossSkin;
/* means3D [P,3] is always written; cov3D [P,6] (needs scaling+rotation), transforms [P,9], bweights [P,24],
 * translation [P,3] (return_transl) are optional (NULL to skip). */
int moss_skin_forward(void* stream, const MossSkin* skin, float* means3D, float* cov3D, float* transforms,
                      float* bweights, float* translation);
/* Backward (the reference's is torch autograd through the python ops).  g_means3D [P,3], g_cov3D [P,6] or NULL.
 * Fully written: d_xyz [P,3], d_scaling [P,3], d_rotation [P,4] (NULL allowed without scaling), d_lbs_delta [P,24]
 * (NULL allowed).  Optional accumulators, cleared by this call: d_A_tgt [24,16], d_poseoff_tgt [V,3]. */
int moss_skin_backward(void* stream, const MossSkin* skin, const float* g_means3D, const float* g_cov3D,
                       float* d_xyz, float* d_scaling, float* d_rotation, float* d_lbs_delta,
                       float* d_A_tgt, float* d_poseoff_tgt);
size_t moss_sizeof_skin(void);

/* Fused skin + rasterize (north_star subsystem 1): MOSS's one-view-per-iteration loop,
 * gaussian_renderer/__init__.py:57-93 (coarse_deform_c2source + get_covariance) feeding the rasterizer's preprocess
 * (DGR/cuda_rasterizer/forward.cu:155-256), with the skinned centre and the transformed covariance held in registers of
 * ONE per-Gaussian kernel each way: means3D and cov3D_precomp (36 B per Gaussian, written and read back in the two-call
 * form, both directions) never touch HBM, and one launch disappears from the forward and from the backward.
 * `scene` carries the camera, colours (shs / colors_precomp), opacities and the workspaces; its means3D, scales, rotations
 * and cov3D_precomp must be NULL.  `skin` as for moss_skin_forward() with scaling and rotation present, skin->P == scene->P.
 * The remaining arguments are those of moss_rasterize_forward() plus two optional per-Gaussian outputs: means3D [P,3]
 * (the skinned centres) and transforms [P,9] (render()'s "transforms"); NULL skips the store. */
int moss_skin_rasterize_forward(void* stream, const MossScene* scene, const MossSkin* skin, unsigned flags,
                                float* out_color, float* out_depth, float* out_alpha, int* radii, int* num_rendered,
                                float* means3D, float* transforms);
/* Backward of the fused call: the rasterizer's dL/dmeans3D and dL/dcov3D stay in registers and flow straight into the
 * skinning backward.  Image-side arguments and dL_dmean2D / dL_dconic / dL_dopacity / dL_dcolor / dL_dsh as for
 * moss_rasterize_backward(); front-end gradients as for moss_skin_backward() (d_xyz, d_scaling, d_rotation fully written;
 * d_lbs_delta, d_A_tgt, d_poseoff_tgt optional, the last two cleared by this call). */
int moss_skin_rasterize_backward(void* stream, const MossScene* scene, const MossSkin* skin, unsigned flags, const int* radii,
                                 const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
                                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor, float* dL_dsh,
                                 float* d_xyz, float* d_scaling, float* d_rotation, float* d_lbs_delta,
                                 float* d_A_tgt, float* d_poseoff_tgt);

/* Per-frame SMPL tables: what render() recomputes for every frame before it touches a
 * Gaussian, as ONE launch forward and ONE backward instead of ~60 torch launches and two Python loops over the joints.
 * Replaces batch_rodrigues (scene/gaussian_model.py:1033-1064), the optional pose correction rot[1:] @ correct_Rs
 * (:885-888, :1015-1018), get_transform_params_torch (:998-1031: v_shaped, joints = J_regressor . v_shaped),
 * get_rigid_transformation_torch (:965-995: the 24-joint kinematic chain), the pose blend shapes
 * pose_offsets = (rot[1:] - I).flatten() @ posedirs^T (:855-861, :878-893) and the shape blend shapes (:870-871).
 * All arrays float32 on the device unless noted. */
typedef struct MossSmplFrame {
    int V;                      /* SMPL vertices (6890) */
    int n_shapes;               /* betas used, <= shape_stride (params['shapes'].shape[-1]) */
    int shape_stride;           /* last dimension of shapedirs (10) */
    const float* shapedirs;     /* [V,3,shape_stride] */
    const float* posedirs;      /* [V,3,207] (the reference's layout: one row of 207 per vertex coordinate) */
    const float* J_template;    /* [24,3]              J_regressor . v_template   } from moss_smpl_joint_tables(), */
    const float* J_shapedirs;   /* [24,3,shape_stride] J_regressor . shapedirs    } once per SMPL model            */
    const int*   parents;       /* int32 [24], kintree_table[0]; parents[j] < j, parents[0] ignored */
    const float* poses;         /* [72] axis-angle */
    const float* shapes;        /* [n_shapes] */
    const float* correct_Rs;    /* [23,9] or NULL */
    void* scratch;              /* >= moss_smpl_frame_scratch_bytes(V) bytes, zeroed ONCE by the caller (self-cleaning);
                                   used by the backward only */
    size_t scratch_bytes;
} MossSmplFrame;
size_t moss_smpl_frame_scratch_bytes(int V);
size_t moss_sizeof_smpl_frame(void);
/* The two joint-regression tables of an SMPL model (constants of a run): joints = J_regressor . (v_template +
 * shapedirs . betas) (scene/gaussian_model.py:1007, :1021) = J_template + J_shapedirs . betas, so the per-frame kernel
 * needs no pass over the vertices.  v_template [V,3], shapedirs [V,3,S], J_regressor [24,V] -> J_template [24,3],
 * J_shapedirs [24,3,S]. */
int moss_smpl_joint_tables(void* stream, int V, int S, const float* v_template, const float* shapedirs,
                           const float* J_regressor, float* J_template, float* J_shapedirs);
/* A [24,16] (row-major 4x4 per joint) is always written; rot_mats [24,9] (after the correction), joints [24,3],
 * poseoff [V,3], shapeoff [V,3] are optional (NULL to skip; a NULL poseoff skips reading posedirs altogether). */
int moss_smpl_frame_forward(void* stream, const MossSmplFrame* frame, float* A, float* rot_mats, float* joints,
                            float* poseoff, float* shapeoff);
/* Gradients w.r.t. poses [72] (always written) and correct_Rs [23,9] (NULL to skip; zeros when frame->correct_Rs is
 * NULL) from d_A [24,16] and d_poseoff [V,3] (either may be NULL = zero).  A and joints are the forward's outputs.
 * shapes and the SMPL tables are constants of a run (no gradient), as in the reference's training loop. */
int moss_smpl_frame_backward(void* stream, const MossSmplFrame* frame, const float* A, const float* joints,
                             const float* d_A, const float* d_poseoff, float* d_poses, float* d_correct_Rs);

/* Image-space losses fused behind the rasterizer.  Replaces, per iteration, the masked
 * L1 / L2 terms, the SSIM and the S3IM loss of train_ZJU.py:107-128 (utils/loss_utils.py:16-87: l1_loss, l2_loss, ssim
 * with an 11x11 sigma-1.5 window, s3im_fun) AND their autograd backward: ONE launch writes the loss values and
 * dL/dimage, dL/dalpha -- the upstream gradients moss_rasterize_backward() takes -- for
 *     loss = w_l1 * Ll1 + w_mask * mask_loss + w_ssim * (1 - ssim) + w_s3im * s3im_loss   (1, 0.5, 0.2, 0.3 at :131).
 * LPIPS (a VGG network) and the pose NLL stay on the caller's side and add their own gradient. */
typedef struct MossImageLoss {
    int W, H;
    const float* image;         /* [3,H,W] render */
    const float* alpha;         /* [H,W] render_alpha, or NULL (no mask loss) */
    const float* gt_image;      /* [3,H,W] viewpoint_cam.original_image */
    const float* bkgd_mask;     /* [H,W] float, or NULL (no mask loss) */
    const float* bound_mask;    /* [H,W] float; pixels equal to 1 enter Ll1 / mask_loss (:111-112); NULL = all pixels */
    const int*   rect;          /* DEVICE int[5]: crop x, y, w, h (cv2.boundingRect of bound_mask, :115) and the number of
                                   bound_mask == 1 pixels -- what moss_mask_bounds() writes; constant per camera */
    float w_l1, w_mask, w_ssim, w_s3im;     /* a zero weight skips that term's work */
    void* scratch;              /* >= moss_image_losses_scratch_bytes(W,H), zeroed ONCE by the caller (self-cleaning) */
    size_t scratch_bytes;
} MossImageLoss;
size_t moss_image_losses_scratch_bytes(int W, int H);
size_t moss_sizeof_image_loss(void);
/* losses: DEVICE float[5] = Ll1, mask_loss, ssim, s3im_loss, weighted total.  dL_dimage [3,H,W] and dL_dalpha [H,W] are
 * fully written (zero where no term reaches), either may be NULL. */
int moss_image_losses(void* stream, const MossImageLoss* loss, float* losses, float* dL_dimage, float* dL_dalpha);
/* moss_image_losses() for callers that cannot allocate device memory (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * Same inputs, outputs and results as moss_image_losses(), but the scratch memory
 * (ticket + per-tile partial sums) is a file-scope static __device__ array inside the library instead of a caller
 * buffer: loss->scratch and loss->scratch_bytes are IGNORED (may be NULL / 0), and the call performs no cudaMalloc,
 * cudaFree or any other allocating call.  Limit: at most MOSS_IMAGE_LOSSES_STATIC_MAX_TILES tiles of 32x32 pixels,
 * ceil(W/32) * ceil(H/32) <= 4096 (e.g. 2048 x 2048); a larger image returns MOSS_ERR_STATIC_SCRATCH_LIMIT and launches
 * nothing.  Every GPU that loads the library holds the array (MOSS_IMAGE_LOSSES_STATIC_SCRATCH_BYTES = 131 360 bytes: the launch's tile grid is
 * aligned to the crop, one extra tile column and row)
 * until the process exits.  All calls on one device share it: calls that may OVERLAP on the GPU (different streams
 * without an ordering between them) race on it and give undefined results -- serialise them or use moss_image_losses()
 * with one scratch buffer per stream.  The part that must start at zero (the ticket) is cleared on `stream` by every
 * call, so an aborted earlier call leaves nothing behind. */
#define MOSS_IMAGE_LOSSES_STATIC_MAX_TILES 4096
#define MOSS_IMAGE_LOSSES_STATIC_SCRATCH_BYTES (256 + 16 * (2 * MOSS_IMAGE_LOSSES_STATIC_MAX_TILES + 2))
int moss_image_losses_static(void* stream, const MossImageLoss* loss, float* losses, float* dL_dimage, float* dL_dalpha);
/* out: DEVICE int[8]; out[0..4] = x, y, w, h of the bounding rectangle of the non-zero pixels of mask (as
 * cv2.boundingRect(mask.astype(np.uint8)), (0,0,0,0) if none) and the number of pixels equal to 1; out[5..7] scratch. */
int moss_mask_bounds(void* stream, int W, int H, const float* mask, int* out);
/* a[0..na) and b[0..nb) *= *scale_dev unless it is exactly 1: the autograd backward of a loss node, without a host read. */
int moss_scale_gradients(void* stream, const float* scale_dev, float* a, size_t na, float* b, size_t nb);

/* Densification statistics of one iteration, for the Gaussians with radii > 0 (the reference's visibility_filter,
 * gaussian_renderer/__init__.py:128): max_radii2D = max(max_radii2D, radii) (train_ZJU.py:173), xyz_gradient_accum +=
 * |dL_dmean2D[:, :2]|, denom += 1 (GaussianModel.add_densification_stats, scene/gaussian_model.py:816-818).  All
 * float [P] (the reference's [P,1] tensors are the same memory), updated in place; any of the three may be NULL.
 * dL_dmean2D [P,3] is the backward's screen-space gradient (viewspace_point_tensor.grad). */
int moss_densify_stats(void* stream, int P, const int* radii, const float* dL_dmean2D, float* xyz_gradient_accum,
                       float* denom, float* max_radii2D);

/* Fused AdamW step over up to 8 parameter groups in ONE launch.  Replaces `gaussians.optimizer.step()` for the
 * per-Gaussian groups of `torch.optim.AdamW(l, lr=0.0, eps=1e-15)` (scene/gaussian_model.py:203-225, train_ZJU.py:189);
 * same arithmetic and order as torch/optim/adam.py::_single_tensor_adam with decoupled weight decay, amsgrad and
 * maximize off.  `step` is the 1-based step count AFTER this update (torch increments before using it).  All arrays
 * float32 [numel], updated in place (grad is read only). */
typedef struct MossAdamGroup {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    size_t numel;
    double lr;          /* hyper-parameters are doubles, like the Python floats torch derives its scalars from: */
    int step;           /* 1 - beta2 formed from a float beta2 would be off by 1e-5 relative */
} MossAdamGroup;
int moss_adamw_step(void* stream, int n_groups, const MossAdamGroup* groups, double beta1, double beta2, double eps,
                    double weight_decay);
size_t moss_sizeof_adam_group(void);

/* Nearest SMPL vertex of every Gaussian: replaces the knn_cuda call `self.knn(smpl_pts.float(), query_pts.float())`
 * with KNN(k=1, transpose_mode=True) (scene/gaussian_model.py:85, :827 -> vert_ids; :657 -> distance).  Exhaustive
 * search, smallest squared distance, lowest index on ties.  xyz [P,3], verts [V,3] -> vert_ids int32 [P] and,
 * optionally (NULL to skip), dist float [P] (Euclidean).  V >= 1. */
int moss_nearest_vertex(void* stream, int P, const float* xyz, int V, const float* verts, int* vert_ids, float* dist);
/* The same search through a uniform 32^3 grid over the vertices (constant for a run: build once, query every render).
 * Identical answers to moss_nearest_vertex() -- same distance expression, same (distance, index) order, ties included --
 * while a Gaussian near the surface examines a few dozen vertices instead of all V.  `grid`: caller-owned device
 * buffer of >= moss_vertex_grid_bytes(V) bytes, 16-byte aligned, written by the build and read by the queries. */
size_t moss_vertex_grid_bytes(int V);
int moss_vertex_grid_build(void* stream, int V, const float* verts, void* grid, size_t grid_bytes);
int moss_nearest_vertex_grid(void* stream, int P, const float* xyz, int V, const void* grid, size_t grid_bytes,
                             int* vert_ids, float* dist);

/* Data-parallel exchange step (no reference counterpart: MOSS trains on one GPU).  Sum, in
 * place and on every rank, `numel` floats starting `offset_floats` (multiple of 4) into a SYMMETRIC buffer: the same
 * allocation made on each of `world` (<= 8) GPUs of one node and mapped into every process (CUDA IPC / fabric handles,
 * e.g. torch.distributed._symmetric_memory).  buffer_ptrs_dev[p] / signal_ptrs_dev[p] are DEVICE arrays holding rank
 * p's buffer and zero-initialised signal pad (signal_pad_bytes each, >= moss_peer_signal_bytes(world); a larger pad
 * allows more blocks) as mapped in the calling process.
 * One kernel per rank: flag barrier, two-shot all-reduce with loads/stores on the peer pointers (rank order, so the
 * result is bit-identical on every rank and from run to run), flag barrier.  Every rank must call it with the same
 * numel/offset and either the same, strictly increasing, non-zero `epoch`, or always epoch 0: the epoch then lives on
 * the device (error_word_dev[1] counts completed exchanges; the last block to finish bumps it), which makes the call
 * CUDA-graph capturable.  error_word_dev: device uint32[4], zeroed by the caller once; [0] becomes non-zero if a peer did not
 * arrive within ~10 s.  multicast_ptr: the buffer's NVLS multicast mapping or NULL; with it the data phase is one
 * multimem.ld_reduce + one multimem.st per 16 bytes (reduction inside the NVSwitch) instead of world loads + stores. */
size_t moss_peer_signal_bytes(int world);
int moss_peer_allreduce(void* stream, int rank, int world, void* const* buffer_ptrs_dev, void* const* signal_ptrs_dev,
                        size_t signal_pad_bytes, size_t offset_floats, size_t numel, unsigned epoch,
                        unsigned* error_word_dev, void* multicast_ptr);

/* Backward with the data-parallel exchange fused in.  `push` describes a SYMMETRIC packed gradient
 * buffer (same allocation on each of `world` <= 8 GPUs of one node, mapped into every process) plus a symmetric staging
 * area and the signal pads.  Every output pointer of this call that lies inside [grad_base, grad_base + numel) is not
 * stored locally: the kernel that produces the rows (backward preprocess, or the view sum of a batch) stores each
 * element straight into the staging area of the rank owning it (reduce-scatter as posted NVLink stores), and ONE short
 * kernel behind it sums the `world` staged copies of this rank's slice in rank order and stores the result into every
 * rank's buffer (multimem.st through the NVSwitch when `multicast_grad` is given).  When the call's work has finished
 * on all ranks, [grad_base, grad_base + numel) holds the sum over ranks on every rank, bit-identical everywhere and from
 * run to run (given MOSS_FLAG_DETERMINISTIC).  Outputs outside that range stay local.  All ranks must make the call
 * with the same numel.  `views` == NULL with V == 1: the scene's own camera.
 * Staging area: world * moss_push_slice_floats(numel, world) floats per rank, zeroed once by the caller. */
typedef struct MossPush {
    int rank, world;
    float* grad_base;            /* this rank's packed gradient buffer, 16-byte aligned */
    size_t numel;                /* floats summed over the ranks (multiple of 4) */
    void* const* grad_ptrs;      /* HOST array [world]: every rank's grad_base as mapped in this process */
    void* const* stage_ptrs;     /* HOST array [world]: every rank's staging area as mapped in this process */
    void* const* signal_ptrs;    /* HOST array [world]: zero-initialised signal pads (shared with moss_peer_allreduce) */
    size_t signal_pad_bytes;
    void* multicast_grad;        /* NVLS multicast mapping of grad_base, or NULL */
    unsigned* error_word_dev;    /* device uint32[4], zeroed once; [0] != 0: a peer did not arrive; [1]: exchanges done */
} MossPush;
size_t moss_push_slice_floats(size_t numel, int world);
/* The exchange alone, for a packed gradient that something else produced: sums [grad_base, grad_base + numel) over the
 * ranks in place (push to the owners' staging areas, gather), one kernel. */
int moss_push_allreduce(void* stream, const MossPush* push);
size_t moss_sizeof_push(void);
int moss_rasterize_backward_push(void* stream, const MossScene* scene, int V, const MossView* views, unsigned flags,
                                 const int* radii, const float* dL_dpix, const float* dL_ddepth, const float* dL_dalpha,
                                 float* dL_dmean2D, float* dL_dconic, float* dL_dopacity, float* dL_dcolor,
                                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot,
                                 void* tmp_ws, size_t tmp_bytes, const MossPush* push);

/* Internal-state export for parity tests (the reference needs a C++ shim over fromChunk for this,
 * DGR/cuda_rasterizer/rasterizer_impl.cu:155-194).  Copies one state array into `dst` (device) on
 * `stream`; returns the byte count or a negative MOSS_ERR_*. */
#define MOSS_STATE_DEPTHS        0  /* f32[P] */
#define MOSS_STATE_CLAMPED       1  /* u8[3P] */
#define MOSS_STATE_MEANS2D       2  /* f32[2P] */
#define MOSS_STATE_COV3D         3  /* f32[6P] (only when built from scales/rotations) */
#define MOSS_STATE_CONIC_OPACITY 4  /* f32[4P] */
#define MOSS_STATE_RGB           5  /* f32[3P] */
#define MOSS_STATE_TILES_TOUCHED 7  /* u32[P] */
#define MOSS_STATE_KEYS_SORTED   9  /* u64[R] (needs MOSS_FLAG_EXPORT_KEYS) */
#define MOSS_STATE_POINT_LIST    11 /* u32[R] */
#define MOSS_STATE_RANGES        12 /* u32[2*tiles] */
#define MOSS_STATE_N_CONTRIB     13 /* u32[W*H] */
#define MOSS_STATE_FINAL_T       14 /* f32[W*H] */
#define MOSS_STATE_COUNTERS      15 /* u32[4]: num_rendered, overflow flag, capacity, reserved */
long long moss_export_state(void* stream, const MossScene* scene, int which, void* dst, long long max_bytes);
/* The same for the workspaces of a batched call: per-Gaussian arrays are [V*P], per-pixel [V*W*H], ranges [V*tiles] with
 * offsets into the batch's one sorted list (view v's instances are contiguous: MOSS_STATE_VIEW_BASE, u32[V+1]);
 * MOSS_STATE_POINT_LIST holds id' = v*P + i. */
#define MOSS_STATE_VIEW_BASE     16
long long moss_export_state_batch(void* stream, const MossScene* scene, int V, int which, void* dst, long long max_bytes);

/* Instrumentation (no reference counterpart: the reference only has cudaDeviceSynchronize-after-every-launch
 * debugging, DGR/cuda_rasterizer/auxiliary.h:166-173).
 *   moss_kernel_launches(): number of kernels this library has launched in this process (bench.py's gpu_launches).
 *   moss_profile_enable(1): bracket every kernel launch with CUDA events on the launching stream;
 *   moss_profile_read(): synchronise the recorded events, add their durations per stage into ms[0..MOSS_STAGE_COUNT),
 *   launch counts into launches[], and clear the pool.  Returns MOSS_OK or MOSS_ERR_CUDA. */
#define MOSS_STAGE_PREPROCESS      0
#define MOSS_STAGE_TILE_SCAN       1
#define MOSS_STAGE_SCATTER         2
#define MOSS_STAGE_BLEND_FWD       3
#define MOSS_STAGE_BLEND_BWD       4
#define MOSS_STAGE_PREPROCESS_BWD  5
#define MOSS_STAGE_OTHER           6
#define MOSS_STAGE_TILE_SORT       7
#define MOSS_STAGE_SKIN            8
#define MOSS_STAGE_SKIN_BWD        9
#define MOSS_STAGE_PEER_REDUCE     10
#define MOSS_STAGE_NEAREST         11
#define MOSS_STAGE_DENSIFY_STATS   12
#define MOSS_STAGE_ADAMW           13
#define MOSS_STAGE_SUM_VIEWS       14
#define MOSS_STAGE_PUSH_GATHER     15
#define MOSS_STAGE_SMPL_FRAME      16
#define MOSS_STAGE_SMPL_FRAME_BWD  17
#define MOSS_STAGE_IMAGE_LOSSES    18
#define MOSS_STAGE_COUNT           19
/*   moss_nvtx_enable(1) (or MOSS_NVTX=1 in the environment): one NVTX range per C-ABI call, named like the entry point, and
 *   one per kernel launch inside it, named like moss_stage_name(); `ncu --nvtx --nvtx-include "moss_rasterize_backward/"`
 *   or an nsys timeline then attribute every launch.  Off by default; a no-op without a tool. */
unsigned long long moss_kernel_launches(void);
int moss_nvtx_enable(int on);
int moss_profile_enable(int on);
int moss_profile_read(double* ms, unsigned long long* launches);
const char* moss_stage_name(int stage);

int moss_abi_version(void);
size_t moss_sizeof_scene(void);
const char* moss_error_string(int status);
int moss_last_cuda_error(void);

#ifdef __cplusplus
}
#endif
#endif /* MOSS_SPLAT_H_INCLUDED */
