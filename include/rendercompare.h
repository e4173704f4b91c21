/*
 * rendercompare.h -- C ABI of librendercompare.so (sm_100a).
 *
 * Drop-in boundary for the render-and-compare data path of DeepIM
 * (reference: wangg12/mx-DeepIM).  Every entry point is `extern "C"`, takes
 * plain pointers and sizes, and (except the host-pointer drop-in `_flow`)
 * returns an int status.  "dev" entry points take DEVICE pointers plus a
 * CUDA stream handle (cudaStream_t passed as void*; NULL = legacy default
 * stream); they never allocate, never synchronise and never touch the host.
 *
 * Each declaration cites the reference interface it replaces (paths relative
 * to the reference root).
 */
#ifndef RENDERCOMPARE_H_
#define RENDERCOMPARE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ------------------------------------------------------------------ status */
#define RC_OK 0
#define RC_ERR_INVALID_ARG 1  /* null pointer, negative size, bad enum        */
#define RC_ERR_CUDA 2         /* a CUDA runtime call failed; see rc_last_cuda_error */
#define RC_ERR_UNALIGNED 3    /* pointer/width not aligned as the entry point requires */
#define RC_ERR_CAPACITY 4     /* a caller-sized workspace or handle table is too small */

int rc_version(void);                      /* ABI version, currently 1                */
const char *rc_status_string(int status);  /* static string for an RC_* code          */
const char *rc_last_cuda_error(void);      /* cudaGetErrorString of the last failure  */
int rc_device_sm_count(int device_id, int *sm_count);

/* ---------------------------------------------------------------- GT flow
 * Replaces lib/flow_c/gpu_flow_kernel.cu:32-69 (flow_kernel) and :82-148
 * (_flow); bound by lib/flow_c/gpu_flow.pyx:12-16,24-41.
 *
 * Layouts (all float32, C-contiguous):
 *   depth_src [B,1,H,W]  rendered depth (metres, 0 = background)
 *   depth_tgt [B,1,H,W]  observed depth
 *   KT        [B,3,4]    K * [R_rel | t_rel]
 *   Kinv      [3,3]
 *   flow      [B,2,H,W]  channel 0 = delta-row, channel 1 = delta-col
 *   valid     [B,1,H,W]  1.0 where the z-test passed, else 0.0
 */

/* Same signature and HOST-pointer contract as the reference's `_flow`
 * (lib/flow_c/gpu_flow.hpp:1-3): blocking, selects `device_id`, outputs are
 * fully overwritten.  Like the reference it has no return value; unlike the
 * reference a CUDA failure is recorded (rc_last_status) instead of printed. */
/* The library ALSO exports this function under the reference's own C++ linkage
 * (`_Z5_flowPfS_S_S_S_S_iiii`, csrc/flow_cxx.cu), which is what lib/flow_c/gpu_flow.pyx:12-16
 * links against when it is compiled as C++ with the reference's unmodified gpu_flow.hpp. */
void _flow(float *flow, float *valid, float *depth_src, float *depth_tgt,
           float *KT, float *Kinv, int batch_size, int height, int width,
           int device_id);
int rc_last_status(void);

/* Same work, status-returning.  Uses an internal per-device staging arena
 * that grows monotonically (no per-call cudaMalloc/cudaFree).  Re-entrant like
 * the reference's `_flow`: concurrent callers on one device are serialised on
 * the arena, callers on different devices do not interact.  Batches of two or
 * more pairs are walked by up to 8 worker threads (env RC_FLOW_HOST_THREADS;
 * 0 / 1 = plain pageable copies), each with a page-locked slot, a device slot
 * and a stream, so the host-side staging of the caller's pageable buffers runs
 * in parallel and overlaps the transfers; the results are the same kernel's. */
int rc_flow_host(float *flow, float *valid, const float *depth_src,
                 const float *depth_tgt, const float *KT, const float *Kinv,
                 int batch_size, int height, int width, int device_id);

/* Frees the host-pointer path's grow-only staging arena of `device_id` (device buffer,
 * page-locked slots, worker streams).  The reference's _flow mallocs and frees per call
 * (gpu_flow_kernel.cu:86-147); ours keeps the buffers between calls, so a long-lived process
 * that is done with the host-pointer entry points can hand the memory back.  Safe to call
 * at any time; the next _flow / rc_flow_host call re-creates what it needs. */
int rc_flow_host_release(int device_id);

/* Device-pointer fast path.  Requires W % 4 == 0 and 16-byte aligned planes. */
int rc_flow_dev(const float *depth_src, const float *depth_tgt, const float *KT,
                const float *Kinv, float *flow, float *valid, int batch_size,
                int height, int width, void *stream);

/* --------------------------------------------------------------- zoom ops
 * Replace the MXNet CustomOps in deepim/operator_py (zoom_mask.py,
 * zoom_image.py, zoom_image_with_factor.py, zoom_depth.py, zoom_flow.py,
 * zoom_mask_with_factor.py, zoom_trans.py) together with the MXNet built-ins
 * they call (GridGenerator(affine) + BilinearSampler).
 */

/* Crop box + zoom factor.  Replaces the per-sample host block
 * deepim/operator_py/zoom_mask.py:36-119 (mode RC_ZF_MASK) and
 * deepim/operator_py/zoom_image.py:30-118 (mode RC_ZF_IMAGE).
 *
 *   obs, rend   [B,C,H,W] the two inputs the bbox is taken from
 *               RC_ZF_MASK : obs = mask_gt_observed, rend = mask_rendered;
 *                            pixel valid iff sum_c obs > 0.3, resp.
 *                            sum_c (rend > 0.2 ? 1 : 0) > 0.3
 *               RC_ZF_IMAGE: pixel valid iff sum_c (img + pixel_means[c]) > 0.01
 *   src_pose    [B,3,4]
 *   K           [3,3] (device)
 *   pixel_means [C] (device; RC_ZF_IMAGE only, may be NULL for RC_ZF_MASK)
 *   zoom_factor [B,4]  (wx, wy, tx, ty) float32
 *   bbox        [B,8]  int32: obs x0,x1,y0,y1 then rend x0,x1,y0,y1
 *               (x0 > x1 marks an empty set; rend empty -> reference fallback,
 *               obs empty -> zoom_factor row is NaN, the reference raises)
 *   workspace   [B,RC_ZF_WS_INTS] int32 scratch, caller-owned.  Must be ZERO before the first
 *               call; every call leaves it zeroed again (no init launch, no memset).
 */
#define RC_ZF_WS_INTS 10
#define RC_ZF_MASK 0
#define RC_ZF_IMAGE 1
int rc_zoom_factor_dev(const float *obs, const float *rend, const float *src_pose,
                       const float *K, const float *pixel_means, float *zoom_factor,
                       int32_t *bbox, int32_t *workspace, int batch_size,
                       int channels, int height, int width, int mode, void *stream);

/* The same kernel for callers that cannot allocate device memory (CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Arguments and results as rc_zoom_factor_dev, minus
 * `workspace`: the scratch is a static device array inside the library
 * (RC_ZF_NOALLOC_MAX_BATCH x RC_ZF_WS_INTS int32 = 160 KB per GPU, held until the process exits),
 * cleared by a stream-ordered memset in front of every launch.  Calls no cudaMalloc / cudaFree /
 * allocating routine.  batch_size > RC_ZF_NOALLOC_MAX_BATCH returns RC_ERR_CAPACITY ("capacity
 * exceeded") and launches nothing.  Every call on a device shares the array: calls that overlap
 * on the GPU (different streams without an ordering between them) RACE -- order them, or use
 * rc_zoom_factor_dev with one workspace per stream. */
#define RC_ZF_NOALLOC_MAX_BATCH 4096
int rc_zoom_factor_noalloc_dev(const float *obs, const float *rend, const float *src_pose,
                               const float *K, const float *pixel_means, float *zoom_factor,
                               int32_t *bbox, int batch_size, int channels, int height, int width,
                               int mode, void *stream);
int rc_zoom_factor_noalloc_max_batch(void); /* = RC_ZF_NOALLOC_MAX_BATCH */

/* One resampled plane family.  `src`/`dst` point at [B,channels,H,W] tensors;
 * every channel of the family gets the same pre/post operation.            */
typedef struct rc_zoom_group {
  const float *src;
  float *dst;
  int32_t channels;
  int32_t pre_op;   /* RC_PRE_*  applied to each source tap before blending  */
  int32_t post_op;  /* RC_POST_* applied to the blended value                */
  int32_t flags;    /* RC_ZOOM_ACCUMULATE: dst += value (req "add") instead of dst = value;
                       RC_ZOOM_SRC_BROADCAST: src is [B,1,H,W] and feeds every dst channel
                       (np.tile(valid, [1,2,1,1]) of batch_updater_py_multi.py:352 without
                       materialising the tiled tensor) */
  float mean[4];    /* per-channel constant for RC_PRE_ADD_MEAN / RC_POST_*_SUB_MEAN */
} rc_zoom_group;

#define RC_ZOOM_ACCUMULATE 1
#define RC_ZOOM_SRC_BROADCAST 2
#define RC_ZOOM_SRC_FRAME_SPARSE 4  /* set by rc_compare_step_dev only: src is zero outside the renderer's
                                       covered tiles and pre/post map 0 to a constant */

#define RC_PRE_NONE 0
#define RC_PRE_THRESH02 1   /* tap = tap > 0.2 ? 1 : 0   (zoom_mask.py:40-45)          */
#define RC_PRE_ADD_MEAN 2   /* tap = tap + mean[c]       (zoom_image_with_factor.py:46) */

#define RC_POST_NONE 0
#define RC_POST_ROUND 1          /* mx.nd.round, half away from zero (zoom_mask.py:121) */
#define RC_POST_ROUND_M045 2     /* round(v - 0.45)           (zoom_flow.py:74-76)      */
#define RC_POST_DIV_WX 3         /* v / wx_in                 (zoom_flow.py:66)         */
#define RC_POST_MUL_WX 4         /* v * wx_in                 (zoom_flow.py:64)         */
#define RC_POST_SUB_MEAN 5       /* v - mean[c]               (zoom_image_with_factor.py:71-72) */
#define RC_POST_SPOT_SUB_MEAN 6  /* max(v, centre spot) - mean[c]  (:36-45,:67-70)      */

#define RC_ZOOM_MAX_GROUPS 12
#define RC_ZOOM_MAX_PLANES 16

/* Bilinear resampling of up to RC_ZOOM_MAX_PLANES planes that share one
 * sampling grid per sample.  Replaces the GridGenerator loops and the
 * BilinearSampler calls at zoom_mask.py:107-129, zoom_image.py:95-115,
 * zoom_image_with_factor.py:52-72, zoom_depth.py:30-48, zoom_flow.py:33-76,
 * zoom_mask_with_factor.py:41-66.
 *   zoom_factor [B,4] device
 *   inverse     0: grid from (wx,wy,tx,ty) as given
 *               1: inverse zoom (zoom_flow.py:35-44): wx' = 1/wx ...
 *   groups      HOST array of n_groups descriptors (copied into kernel params)
 * Requires W % 4 == 0 and 16-byte aligned planes. */
int rc_zoom_sample_dev(const float *zoom_factor, int inverse,
                       const rc_zoom_group *groups, int n_groups, int batch_size,
                       int height, int width, void *stream);

/* ZoomTrans forward and backward (deepim/operator_py/zoom_trans.py:22-76).
 *   trans_in/out [B,3]; x,y are divided by wx (inverse=0) or multiplied
 *   (inverse=1); z passes through.  float32 scalar arithmetic (np.float32 op
 *   np.float32 in the reference).  Forward: scale_xy = 1.  Backward: pass the
 *   incoming gradient as trans_in and scale_xy = b_zoom_grad (:59-70).     */
int rc_zoom_trans_dev(const float *zoom_factor, const float *trans_in,
                      float *trans_out, int batch_size, int inverse, int scale_xy,
                      int accumulate, void *stream);

/* Zero-fill (every zoom backward: zoom_mask.py:136-140, zoom_flow.py:79-82 ...). */
int rc_fill_zero_dev(float *dst, int64_t count, void *stream);

/* ------------------------------------------------ point-matching loss path
 * Replaces deepim/operator_py/transform3d.py:42-118 (forward), :120-327
 * (backward) and the loss graph deepim/symbols/deepIM_flownet.py:460-499.
 *
 *   points    [B,3,N]  model points (zero padded)
 *   rotation  [B,4]    quaternion w,x,y,z (already L2-normalised upstream)
 *   trans     [B,3]    translation delta (un-zoomed)
 *   pose_src  [B,3,4]
 *   T_means, T_stds  [3] HOST floats (CustomOp kwargs in the reference)
 *   rot_coord  RC_ROT_MODEL / RC_ROT_CAMERA / RC_ROT_CAMERA_NEW / RC_ROT_NAIVE
 */
#define RC_ROT_MODEL 0
#define RC_ROT_CAMERA 1
#define RC_ROT_CAMERA_NEW 2
#define RC_ROT_NAIVE 3

/* Transform3D forward: out[B,3,N] = (R_delta (x) R_src) * points + T_tgt. */
int rc_transform3d_fwd_dev(const float *points, const float *rotation,
                           const float *trans, const float *pose_src,
                           const float *T_means, const float *T_stds, int rot_coord,
                           float *out, int batch_size, int num_points,
                           int accumulate, void *stream);

/* Transform3D backward: grad_rot[B,4], grad_trans[B,3] from out_grad[B,3,N]. */
int rc_transform3d_bwd_dev(const float *out_grad, const float *points,
                           const float *rotation, const float *trans,
                           const float *pose_src, const float *T_means,
                           const float *T_stds, int rot_coord, float *grad_rot,
                           float *grad_trans, int batch_size, int num_points,
                           int accumulate, void *stream);

#define RC_PM_L1 0
#define RC_PM_L2 1
#define RC_PM_SMOOTH_L1 2

/* Fused Transform3D + point-matching loss forward.
 *   points_obs [B,3,N], weights [B,3,N]
 *   loss_map   [B,3,N]  w * f((est-obs)/norm_term)  (may be NULL)
 *   loss_sum   [B]      per-sample sum of loss_map  (may be NULL)
 *   est        [B,3,N]  transformed points          (may be NULL)          */
int rc_pm_loss_fwd_dev(const float *points, const float *points_obs,
                       const float *weights, const float *rotation,
                       const float *trans, const float *pose_src,
                       const float *T_means, const float *T_stds, int rot_coord,
                       int loss_type, float norm_term, float sl1_scalar,
                       float *est, float *loss_map, float *loss_sum,
                       int batch_size, int num_points, void *stream);

/* Fused loss-gradient + Transform3D backward: MakeLoss(grad_scale) head
 * gradient, d f / d est, then transform3d.py:120-191 in one pass.          */
int rc_pm_loss_bwd_dev(const float *points, const float *points_obs,
                       const float *weights, const float *rotation,
                       const float *trans, const float *pose_src,
                       const float *T_means, const float *T_stds, int rot_coord,
                       int loss_type, float norm_term, float sl1_scalar,
                       float grad_scale, float *grad_rot, float *grad_trans,
                       int batch_size, int num_points, void *stream);

/* ------------------------------------------------------ fused compare step
 * One call = one pass of the path between two network evaluations for B
 * pairs: gpu_flow (lib/pair_matching/batch_updater_py_multi.py:331-352), then
 * ZoomMask / ZoomImageWithFactor / ZoomFlow / ZoomTrans
 * (deepim/symbols/deepIM_flownet.py:595-698) and Transform3D + point-matching
 * loss (:446-499).  Replaces six Python-level CustomOp invocations and their
 * device<->host syncs by four kernel launches issued from C on one stream.
 * All pointers are device pointers unless marked host.  Optional outputs may
 * be NULL.  `marks` (host array of cudaEvent_t, or NULL) receives up to
 * RC_STEP_MARKS event records: [0] before flow, [1] after flow, [2] after
 * zoom factor (+ fused ZoomTrans), [3] after the resampler, [4] = [3], [5] after
 * the loss; NULL entries are skipped.  With grad_rot/grad_trans set the step also runs the
 * loss backward (one more launch) and, with grad_zoom_trans, the ZoomTrans backward (one more).
 */
#define RC_STEP_MARKS 6
typedef struct rc_compare_step_args {
  /* inputs */
  const float *image_observed;     /* [B,3,H,W] mean-subtracted */
  const float *image_rendered;     /* [B,3,H,W] */
  const float *depth_gt_observed;  /* [B,1,H,W] */
  const float *depth_rendered;     /* [B,1,H,W] (also the rendered "mask", zoom_mask.py:40-45) */
  const float *mask_observed;      /* [B,1,H,W] */
  const float *mask_gt_observed;   /* [B,1,H,W] */
  const float *src_pose;           /* [B,3,4] */
  const float *KT;                 /* [B,3,4] */
  const float *Kinv;               /* [3,3] */
  const float *K;                  /* [3,3] */
  const float *point_cloud_model;    /* [B,3,N] */
  const float *point_cloud_weights;  /* [B,3,N] */
  const float *point_cloud_observed; /* [B,3,N] */
  const float *rot_est;            /* [B,4] */
  const float *trans_est;          /* [B,3] */
  const float *trans_gt;           /* [B,3] */
  /* outputs */
  float *flow;                     /* [B,2,H,W]; NULL = test phase: no GT flow, no flow planes */
  float *valid;                    /* [B,1,H,W] */
  float *zoom_factor;              /* [B,4] */
  int32_t *bbox;                   /* [B,8] or NULL */
  int32_t *workspace;              /* [B,RC_ZF_WS_INTS], zero before first use */
  float *zoom_mask_observed;       /* [B,1,H,W] */
  float *zoom_mask_gt_observed;    /* [B,1,H,W] */
  float *zoom_mask_rendered;       /* [B,1,H,W] */
  float *zoom_image_observed;      /* [B,3,H,W] */
  float *zoom_image_rendered;      /* [B,3,H,W] */
  float *zoom_flow;                /* [B,2,H,W] */
  float *zoom_flow_weights;        /* [B,2,H,W] */
  float *zoom_trans_gt;            /* [B,3] */
  float *loss_map;                 /* [B,3,N] or NULL */
  float *loss_sum;                 /* [B] */
  /* backward leg (training; transform3d.py:120-327 + MakeLoss grad_scale, zoom_trans.py:50-76);
   * all three NULL = forward only */
  float *grad_rot;                 /* [B,4] d loss / d rot_est */
  float *grad_trans;               /* [B,3] d loss / d trans_est (un-zoomed) */
  float *grad_zoom_trans;          /* [B,3] ZoomTrans(b_inv_zoom=True).backward of grad_trans, or NULL */
  /* constants (host values) */
  float pixel_means[4];            /* per image channel, already in channel order */
  float T_means[3];
  float T_stds[3];
  float norm_term;
  float sl1_scalar;
  float grad_scale;                /* MakeLoss grad_scale = LW_PM / NUM_3D_SAMPLE (deepIM_flownet.py:493-498) */
  int32_t rot_coord;
  int32_t loss_type;
  int32_t zoom_grad;               /* b_zoom_grad of invZoomTrans (deepIM_flownet.py:376-383: False) */
  int32_t batch_size, height, width, num_points;
  void *stream;                    /* cudaStream_t */
  void **marks;                    /* host array of RC_STEP_MARKS cudaEvent_t, or NULL */
  /* Sparse frame (optional; all three zero = dense).  When the caller also drives the
   * rasteriser, the rendered frame is zero outside the 16x16 tiles the renderer's work list
   * names, so the step can skip the background: the rendered crop box (zoom_mask.py:40-62) is
   * reduced over those tiles only and combined with `observed_box`, and the GT flow
   * (gpu_flow_kernel.cu:32-69, zero wherever depth_src <= 1e-3) is computed on those tiles
   * and re-zeroed on the tiles the previous frame covered.  Results are identical to the
   * dense step's. */
  void *frame_renderer;            /* rc_renderer handle whose LAST rc_render_dev call wrote depth_rendered, or NULL */
  const int32_t *observed_box;     /* [B,4] x0,x1,y0,y1 of mask_gt_observed > 0.3 (= bbox[:, :4] of a dense
                                      rc_zoom_factor_dev call; x1 < 0: empty); required with frame_renderer */
  int32_t outputs_track_frames;    /* 1: flow/valid and the zoomed outputs still hold THIS step's output for
                                      the renderer's previous frame (same buffers, one step per render);
                                      0: contents unknown, every plane is written in full */
  uint8_t *zoom_tile_state;        /* [B, ceil(H/8) * ceil(W/128) * 4] bytes or NULL.  With frame_renderer: one
                                      flag per 32 x 8 block of resampler outputs = "the rendered-side planes with a
                                      constant background (zoom_mask_rendered, zoom_flow,
                                      zoom_flow_weights) hold that background here".  A tile whose source
                                      window lies in tiles the frame does not cover, and whose flag says
                                      it already holds the background, is not recomputed for those planes.
                                      Maintained by the step; ignored (but rewritten) when
                                      outputs_track_frames == 0 */
  const float *loss_src_pose;      /* [B,3,4] or NULL (= src_pose).  The pose the point-matching loss and its
                                      backward transform the model points with.  In the reference the loss
                                      belongs to the network pass that PRODUCED rot_est/trans_est, i.e. to the
                                      pose before the updater applied them (deepIM_flownet.py:446-499 runs in
                                      the forward on src_pose_i; batch_updater_py_multi.py:211-312 refines the
                                      pose afterwards), while flow/zoom read the refined pose `src_pose`.  A
                                      refine loop therefore binds the pre-update pose here. */
  /* INPUT_DEPTH graphs (deepIM_flownet.py:669-681, test graph :827-835): ZoomDepth of the two depth
   * maps with the step's zoom factor (zoom_depth.py:24-50), as two more planes of the same
   * resampler launch.  All three NULL = the shipped configs (INPUT_DEPTH: False). */
  const float *depth_observed;     /* [B,1,H,W] or NULL */
  float *zoom_depth_observed;      /* [B,1,H,W] or NULL */
  float *zoom_depth_rendered;      /* [B,1,H,W] or NULL (source: depth_rendered) */
  int32_t phases;                  /* 0 = the whole step in one call; else a mask of RC_STEP_PHASE_*: the
                                      caller issues the phases itself, e.g. the loss phase on a second
                                      stream beside the rasteriser (it reads poses and points only) and
                                      the ZoomTrans backward beside the resampler.  Ordering between the
                                      phases is then the caller's: FRAME before SAMPLE and ZT_BWD, LOSS
                                      before ZT_BWD */
} rc_compare_step_args;
#define RC_STEP_PHASE_LOSS 1    /* Transform3D + point-matching loss (+ backward)             */
#define RC_STEP_PHASE_FRAME 2   /* GT flow, crop boxes, zoom factor (+ ZoomTrans forward)     */
#define RC_STEP_PHASE_SAMPLE 4  /* the resampler launch                                       */
#define RC_STEP_PHASE_ZT_BWD 8  /* invZoomTrans.backward of grad_trans                        */

int rc_compare_step_dev(const rc_compare_step_args *args);

/* CUDA-event helpers so that a host language without a CUDA binding can time
 * kernels on the stream they were launched on. */
int rc_event_create(void **event);
int rc_event_destroy(void *event);
int rc_event_record(void *event, void *stream);
int rc_event_synchronize(void *event);
int rc_event_elapsed_ms(void *start, void *end, float *ms);

/* -------------------------------------------------------------- rasteriser
 * Replaces the OpenGL renderers of lib/render_glumpy:
 *   Render_Py.__init__ / render        render_py_multi.py:49-147 (GLSL :17-47)
 *   Render_Py_Light_ModelNet_Multi     render_py_light_modelnet_multi.py:83-231 (GLSL :14-79)
 * and the post-render block lib/pair_matching/batch_updater_py_multi.py:279-320
 * (BGR->RGB, HWC->CHW, - pixel_means, mask = depth > 0.2), fused into the resolve pass.
 *
 * Geometry convention (derived from my_compute_calib_proj / _get_view_mtx,
 * render_py_multi.py:152-178): pixel (x, y) is covered iff the
 * pinhole projection (fx X/Z + cx, fy Y/Z + cy) of the triangle contains the
 * integer point (x, y); nearest fragment wins (GL_LESS), ties go to the lower
 * triangle index; no face culling; near/far planes at zNear/zFar.
 */
typedef struct rc_mesh_desc {      /* HOST pointers, copied at create time */
  const float *vertices;           /* [V,3] object frame, metres            */
  const float *uv;                 /* [V,2] or NULL                         */
  const float *normals;            /* [V,3] or NULL (needed for RC_RENDER_LAMBERT) */
  const int32_t *triangles;        /* [T,3] vertex indices                  */
  const uint8_t *texture;          /* [tex_h,tex_w,3] RGB, row 0 = top of the image file */
  int32_t n_vertices, n_triangles, tex_h, tex_w;
} rc_mesh_desc;

/* Uploads all meshes and allocates every workspace for up to `max_batch`
 * renders per call on `device_id`.  No allocation happens in rc_render_dev. */
int rc_renderer_create(const rc_mesh_desc *meshes, int n_meshes, int height, int width,
                       int max_batch, int device_id, void **handle);
int rc_renderer_destroy(void *handle);

#define RC_RENDER_TEXTURED 0  /* gl_FragColor = texture2D(u_texture, v_texcoord)          */
#define RC_RENDER_LAMBERT 1   /* ... * ((1-r) + r*brightness) * light_intensity (ModelNet) */
#define RC_RENDER_LIT_LINEMOD 2 /* ... * ((1-r) + r*brightness * light_intensity): the LINEMOD
                                   light shaders, render_py_light.py:36-79 and
                                   render_py_light_multi_program.py:38-80 */

typedef struct rc_render_args {
  const int32_t *cls_idx;     /* dev [B] mesh index per render                         */
  const float *poses;         /* dev [B,3,4] object-to-camera [R|t]                    */
  const float *light_pos;     /* dev [B,3] (LAMBERT; GL eye coordinates) or NULL       */
  const float *light_int;     /* dev [B,3] (LAMBERT) or NULL                           */
  float *depth;               /* dev [B,1,H,W] metres, 0 = background                  */
  float *image_chw;           /* dev [B,3,H,W] RGB - pixel_means, or NULL (batch_updater :279-284) */
  float *bgr_hwc;             /* dev [B,H,W,3] BGR 0..255 as render() returns it, or NULL */
  float *mask;                /* dev [B,1,H,W] depth > mask_thresh ? 1 : 0, or NULL (:316-319) */
  int32_t *tri_id;            /* dev [B,H,W] winning triangle, -1 = background, or NULL */
  int32_t *status;            /* dev [1] set non-zero on tile-list overflow, or NULL   */
  float K[9];                 /* intrinsics (host values)                              */
  float pixel_means[3];       /* subtracted from image_chw channels R,G,B              */
  float znear, zfar, mask_thresh, brightness_ratio;
  int32_t mode;               /* RC_RENDER_*                                           */
  int32_t batch_size;
  int32_t incremental;        /* != 0: the caller guarantees that every output buffer still holds,
                                 unmodified, what the previous rc_render_dev call on this handle
                                 wrote to the same pointers (same batch size, same pixel_means);
                                 the background is then re-written only where that frame had
                                 geometry instead of streaming 20 B/px.  Falls back to the full
                                 background pass whenever pointers / batch / means differ.   */
  void *stream;               /* cudaStream_t                                          */
  void **marks;               /* host array of 5 cudaEvent_t (before setup, after setup, after
                                 scan, after fill, after raster) or NULL               */
} rc_render_args;

int rc_render_dev(void *handle, const rc_render_args *args);

/* Tells the handle that `n_renders` rc_render_dev calls captured into a CUDA graph have just been
 * REPLAYED (a replay executes the kernels but not the host side of rc_render_dev, which is what
 * tracks the ping-pong half of the incremental background's item lists).  Only the parity of
 * `n_renders` matters.  Call it after every replay of a graph that holds an odd number of
 * renders; harmless otherwise. */
int rc_renderer_note_replay(void *handle, int n_renders);

/* Counters of the last rc_render_dev call on this handle's workspaces (debug / bench):
 * out[0] = triangles that survived culling, out[1] = tile-list entries, out[2] = how many of
 * out[0] were micro-triangles merged straight into the z-winner blocks (no record, no list
 * entry).  `out` holds 3 values.  Synchronises the stream. */
int rc_render_stats(void *handle, int batch_size, void *stream, int64_t *out);

/* ------------------------------------------------------ pose update (host driver math)
 * The per-pair block of lib/pair_matching/batch_updater_py_multi.py:211-312, batched:
 *   refined_pose = RT_transform(src_pose, rot_est, trans_est)     (RT_transform.py:135-161)
 *   rot_res, trans_res = calc_RT_delta(refined_pose, tgt_pose, rot_type="QUAT")   (:16-48)
 *   KT = K . calc_se3(refined_pose, tgt_pose)                     (:186-197, updater :306-312)
 * All arrays device float32: src_pose, tgt_pose, refined_pose, KT [B,3,4]; rot_est, rot_res
 * [B,4] (w,x,y,z); trans_est, trans_res [B,3]; K [3,3].  T_means/T_stds host [3].
 * refined_pose must not alias src_pose.  rot_coord: MODEL / CAMERA / CAMERA_NEW. */
int rc_pose_update_dev(const float *src_pose, const float *rot_est, const float *trans_est,
                       const float *tgt_pose, const float *K, const float *T_means,
                       const float *T_stds, int rot_coord, float *refined_pose, float *rot_res,
                       float *trans_res, float *KT, int batch_size, void *stream);

/* calc_RT_delta(pose_src, pose_tgt, T_means, T_stds, rot_coord, rot_type="QUAT")
 * (lib/pair_matching/RT_transform.py:16-48) batched on the device: the regression labels of
 * iteration 0, which the reference's loaders compute per pair on the host.  src_pose, tgt_pose
 * [B,3,4]; rot_res [B,4] (w,x,y,z, w >= 0); trans_res [B,3]. */
int rc_pose_delta_dev(const float *src_pose, const float *tgt_pose, const float *T_means,
                      const float *T_stds, int rot_coord, float *rot_res, float *trans_res,
                      int batch_size, void *stream);

/* Synthetic-workload utility (not a reference interface): the network stand-in of SURVEY.md 8(d)
 * for refine loops without the CNN -- rot_est = normalise((1-a) * identity + a * rot_gt + noise_q),
 * trans_est = a * trans_gt + noise_t; all arrays device float32 ([B,4] / [B,3]). */
int rc_standin_predict_dev(const float *rot_gt, const float *trans_gt, const float *noise_q,
                           const float *noise_t, float damping, float *rot_est, float *trans_est,
                           int batch_size, void *stream);

/* rc_pose_update_dev for refine loops: optionally with the stand-in at the head of the same kernel
 * (one launch instead of two on the loop's chain of dependent launches), and optionally in two
 * parts, so that a caller can run the expensive half beside the rasteriser:
 *   parts = RC_POSE_PART_POSE    refined_pose only -- all the render waits for;
 *           RC_POSE_PART_LABELS  rot_res / trans_res / KT only (the float64 eigen-solve of
 *                                mat2quat is most of the kernel's latency);
 *           both                 everything, as rc_pose_update_dev.
 * rot_gt != NULL: the stand-in's prediction is made from (rot_gt, trans_gt, noise) first, WRITTEN
 * to rot_est / trans_est (the loss and the caller read them) and used as the update; rot_gt /
 * trans_gt may alias rot_res / trans_res (labels in, new labels out).  rot_gt == NULL: rot_est /
 * trans_est are the update as given.  Not a reference interface. */
#define RC_POSE_PART_POSE 1
#define RC_POSE_PART_LABELS 2
int rc_pose_update_standin_dev(const float *src_pose, const float *rot_gt, const float *trans_gt,
                               const float *noise_q, const float *noise_t, float damping,
                               float *rot_est, float *trans_est, const float *tgt_pose,
                               const float *K, const float *T_means, const float *T_stds,
                               int rot_coord, float *refined_pose, float *rot_res,
                               float *trans_res, float *KT, int parts, int batch_size,
                               void *stream);

/* ------------------------------------------------ test-time mask policies (8(f) rank 1)
 * The "box_rendered" / "box_observed" branches of update_data_batch,
 * lib/pair_matching/data_pair.py:101-131 (fed by deepim/core/tester.py:571-594): the new
 * observed mask is the bounding rectangle of the non-zero pixels of `mask` (the rendered
 * mask for box_rendered, the observed one for box_observed),
 *     out[y_start:y_end, x_start:x_end] = 1, 0 elsewhere
 * with the reference's exclusive slice end (the last row and column of the box stay 0).
 *   mask, out  dev [B,1,H,W] float32 (non-negative masks); may not alias
 *   bbox       dev [B,4] int32: x_start, x_end, y_start, y_end (np.min / np.max of the
 *              non-zero columns / rows); x_start > x_end marks an empty mask, for which the
 *              reference raises and `out` is all zero
 *   workspace  dev [B,RC_BOX_WS_INTS] int32, zero before the first call; left zeroed.
 * Requires W % 4 == 0 and 16-byte aligned planes. */
#define RC_BOX_WS_INTS 5
int rc_mask_box_dev(const float *mask, float *out, int32_t *bbox, int32_t *workspace,
                    int batch_size, int height, int width, void *stream);

/* -------------------------------------------- iteration-0 input preparation (8(f) rank 2)
 * image.transform, lib/utils/image.py:709-720: BGR HWC uint8 -> [B,3,H,W] float32 RGB with
 * out[:, i] = im[:, :, 2 - i] - pixel_means[2 - i] (float64 arithmetic, float32 store).
 *   bgr_hwc dev [B,H,W,3] uint8 (4-byte aligned)   pixel_means_bgr HOST [3] float64 (B,G,R order) */
int rc_prep_image_dev(const uint8_t *bgr_hwc, float *out_chw, const double *pixel_means_bgr,
                      int batch_size, int height, int width, void *stream);

/* Depth decode, lib/utils/image.py:200-206 / :229-260: uint16 -> float32, optional
 * `depth *= (mask == mask_idx)` (:244), then / depth_factor in float32.
 *   raw dev [B,H,W] uint16 (8-byte aligned); mask dev [B,H,W] uint8 + mask_idx dev [B] int32,
 *   both NULL to skip; out dev [B,1,H,W]. */
int rc_prep_depth_dev(const uint16_t *raw, const uint8_t *mask, const int32_t *mask_idx,
                      float *out, float depth_factor, int batch_size, int height, int width,
                      void *stream);

/* Mask decode, lib/utils/image.py:300-304 (get_pair_mask): out = (mask_png == mask_idx) ? 1 : 0.
 *   mask dev [B,H,W] uint8 (4-byte aligned); mask_idx dev [B] int32; out dev [B,1,H,W]. */
int rc_prep_mask_dev(const uint8_t *mask, const int32_t *mask_idx, float *out, int batch_size,
                     int height, int width, void *stream);

/* Mask from a decoded depth map (metres): out = depth > thresh ? 1 : (keep_small ? depth : 0).
 * keep_small = 1 is mask_rendered of get_pair_mask, lib/utils/image.py:487-492
 * (`depth_rendered[depth_rendered > 0.2] = 1`, smaller values stay); keep_small = 0 is the
 * binary mask the init policies box from (:335-337, :446-447).  count % 4 == 0, 16-byte
 * aligned device pointers; in place allowed. */
int rc_mask_from_depth_dev(const float *depth, float *out, float thresh, int keep_small,
                           int64_t count, void *stream);

/* mask_dilate, lib/utils/mask_dilate.py:10-58, with the random draws made by the caller:
 * thickness dev [B,4] int32 = thickness of the four growth blocks in the reference's order
 * (rows shifted down :22-29, up :30-37, columns shifted right :38-45, left :46-53), 0 where
 * the drawn `direction` skips the block.  mask / out dev [B,1,H,W] float32, out != mask. */
int rc_mask_dilate_dev(const float *mask, const int32_t *thickness, float *out, int batch_size,
                       int height, int width, void *stream);

/* get_point_cloud_observed, lib/utils/image.py:592-596: out = R . X + T per sample
 * (float64 accumulation, float32 store).  points/out dev [B,3,N]; pose dev [B,3,4]. */
int rc_points_transform_dev(const float *points, const float *pose, float *out, int batch_size,
                            int num_points, void *stream);

/* calc_flow, lib/pair_matching/flow.py:12-81 (+ backproject_camera, lib/utils/projection.py:
 * 47-67): the float64 numpy rule used for iteration-0 labels and test-time ground truth
 * (np.round half-to-even, bound test on the rounded pixel, clamped look-up, visible needs
 * depth_src != 0, |d_tgt - z| < thresh and |d_tgt| > 1e-10) -- NOT the CUDA rule of
 * rc_flow_dev.
 *   depth_src, depth_tgt dev [B,1,H,W] float32
 *   transform  dev [B,3,4] float64 = K . se3_mul(pose_tgt, se3_inverse(pose_src)) (:31)
 *   Kinv_host  HOST [9] float64 = inv(K)
 *   flow       dev [B,2,H,W] float32; channel order (dh, dw), or (dw, dh) if standard_rep
 *   visible    dev [B,1,H,W] float32 0/1
 *   x_valid    dev [B,3,capacity] float32 or NULL: back-projected source points of the
 *              visible pixels in pixel order (flow.py:80); counts dev [B] int32 receives
 *              the number per sample (points beyond `capacity` are dropped);
 *   workspace  dev [B * rc_calc_flow_blocks(H, W)] int32 scratch (needed with x_valid). */
int rc_calc_flow_blocks(int height, int width);
int rc_calc_flow_dev(const float *depth_src, const float *depth_tgt, const double *transform,
                     const double *Kinv_host, double thresh, int standard_rep, float *flow,
                     float *visible, float *x_valid, int32_t *counts, int32_t *workspace,
                     int capacity, int batch_size, int height, int width, void *stream);

/* ------------------------------------------------ evaluation reductions (8(f) rank 4)
 * Pose errors of lib/utils/pose_error.py for B (estimate, ground truth) pose pairs:
 *   out [B,5] float64: add (:71-86), adi (:89-109, nearest neighbour by exact search),
 *                      arp_2d (:55-68), re in degrees (:126-131), te (:134-145)
 *   pose_est, pose_gt dev [B,3,4] float32; points dev [B,3,N] float32, or [3,N] shared by all
 *   pairs when points_shared != 0; n_valid dev [B] int32 (points beyond it are padding) or
 *   NULL; K dev [3,3] float32; workspace dev [B * rc_pose_error_chunks(N) * 3] float64.
 *   want_adi = 0 skips the O(N^2) search (adi column = 0). */
int rc_pose_error_chunks(int num_points);
int rc_pose_error_dev(const float *pose_est, const float *pose_gt, const float *points,
                      int points_shared, const int32_t *n_valid, const float *K, int want_adi,
                      double *out, double *workspace, int batch_size, int num_points,
                      void *stream);

/* Visibility masks of lib/utils/visibility.py:7-33 over `count` pixels (any batch shape):
 *   out = d_test > 0 && d_model > 0 && d_model - d_test <= delta           (estimate_visib_mask)
 *   with visib_gt != NULL additionally  || (visib_gt && d_model > 0)    (estimate_visib_mask_est)
 * float32 0/1 output; count % 4 == 0 and 16-byte aligned pointers. */
int rc_visib_mask_dev(const float *d_test, const float *d_model, const float *visib_gt,
                      float delta, float *out, int64_t count, void *stream);

/* Flow end-point errors, deepim/core/tester.py:719-736 (calc_EPE_one_pair):
 *   out [B,6] float64: epe_all, num_all, epe_viz, num_viz, epe_vizbg, num_vizbg
 *   flow_pred, flow_gt dev [B,2,H,W]; visible dev [B,1,H,W];
 *   bg dev [B,1,H,W] or NULL -> visible == 0 && depth_rendered == 0 (tester.py:712) when
 *   depth_rendered [B,1,H,W] is given, else empty;
 *   workspace dev [B * rc_flow_epe_chunks() * 5] float64. */
int rc_flow_epe_chunks(void);
int rc_flow_epe_dev(const float *flow_pred, const float *flow_gt, const float *visible,
                    const float *bg, const float *depth_rendered, double *out, double *workspace,
                    int batch_size, int height, int width, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* RENDERCOMPARE_H_ */
