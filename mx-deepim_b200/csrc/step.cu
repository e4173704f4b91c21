// step.cu -- the fused compare step: every launch of one pass of the path
// between two network evaluations, issued from C on one stream.
//
// Stands in for the host driver
//   /root/reference/lib/pair_matching/batch_updater_py_multi.py:331-352 (gpu_flow + tile)
// and the CustomOp invocations the MXNet executor makes per pass
//   /root/reference/deepim/symbols/deepIM_flownet.py:595-698, :446-499.
//
// B200 design: at batch 16 a step is ~100 us of HBM time; issuing its
// launches from Python (ctypes + tensor bookkeeping per call) would leave the
// GPU idle between kernels.  One C call enqueues all four kernels (flow; bbox + zoom
// factor + ZoomTrans; resampler; point-matching loss) in a few microseconds, so the host
// runs ahead of the device.  Flow, the bbox reduction and the loss are independent of each
// other and each is too small to fill 148 SMs (latency-bound, 15-25 us), so they are issued
// back to back with programmatic stream serialisation (loss -> bbox -> flow; the first two
// call griddepcontrol.launch_dependents at once) and run concurrently; the resampler is a
// normal launch and waits for all three.  Optional event marks bracket each kernel for roofline timing
// on the launching stream; asking for the flow/bbox/loss marks serialises the step so that
// the per-kernel times are clean.
#include "common.cuh"

namespace rc {
static inline int mark(void **marks, int i, cudaStream_t s) {
  if (marks && marks[i]) RC_CUDA_TRY(record_mark(static_cast<cudaEvent_t>(marks[i]), s));
  return RC_OK;
}
}  // namespace rc

extern "C" {

int rc_compare_step_dev(const rc_compare_step_args *a) {
  RC_REQUIRE(a != nullptr, RC_ERR_INVALID_ARG);
  const int B = a->batch_size, H = a->height, W = a->width, N = a->num_points;
  RC_REQUIRE(B >= 0 && H > 1 && W > 1 && N >= 0, RC_ERR_INVALID_ARG);
  if (B == 0) return rc::note_status(RC_OK);
  cudaStream_t s = static_cast<cudaStream_t>(a->stream);
  int st;
  // phases: 0 = everything in this call; else only the named parts (the caller orders them,
  // e.g. the loss on a second stream beside the rasteriser: it reads poses and points only)
  RC_REQUIRE(a->phases >= 0 && a->phases < 16, RC_ERR_INVALID_ARG);
  const int ph = a->phases ? a->phases : 15;
  const bool ph_loss = (ph & RC_STEP_PHASE_LOSS) != 0, ph_frame = (ph & RC_STEP_PHASE_FRAME) != 0,
             ph_sample = (ph & RC_STEP_PHASE_SAMPLE) != 0, ph_ztb = (ph & RC_STEP_PHASE_ZT_BWD) != 0;
  const bool do_loss = ph_loss && N > 0 && a->loss_sum;
  // training backward (cfg 5): d loss / d (rot_est, trans_est) with MakeLoss's grad_scale
  const bool do_bwd = ph_loss && N > 0 && a->grad_rot && a->grad_trans;
  // test phase (deepIM_flownet.py:778-838): no GT flow, no loss -- flow == NULL skips both the
  // flow kernel and the flow planes of the resampler
  const bool do_flow = a->flow != nullptr && a->valid != nullptr;
  // the loss belongs to the pass that produced rot_est / trans_est: pre-update pose if given
  const float *loss_pose = a->loss_src_pose ? a->loss_src_pose : a->src_pose;
  // overlap unless the caller wants per-kernel marks of the overlapped kernels
  const bool serial = a->marks && (a->marks[0] || a->marks[1] || a->marks[4] || a->marks[5]);
  auto loss = [&]() -> int {
    return rc_pm_loss_fwd_dev(a->point_cloud_model, a->point_cloud_observed,
                              a->point_cloud_weights, a->rot_est, a->trans_est, loss_pose,
                              a->T_means, a->T_stds, a->rot_coord, a->loss_type, a->norm_term,
                              a->sl1_scalar, nullptr, a->loss_map, a->loss_sum, B, N, a->stream);
  };
  auto loss_bwd = [&]() -> int {
    return rc_pm_loss_bwd_dev(a->point_cloud_model, a->point_cloud_observed,
                              a->point_cloud_weights, a->rot_est, a->trans_est, loss_pose,
                              a->T_means, a->T_stds, a->rot_coord, a->loss_type, a->norm_term,
                              a->sl1_scalar, a->grad_scale, a->grad_rot, a->grad_trans, B, N,
                              a->stream);
  };
  // crop boxes + zoom factor (ZoomMask, first half); ZoomTrans of the GT translation label
  // rides in the finalising block of the same kernel.
  const bool zt = a->zoom_trans_gt && a->trans_gt;
  auto bbox = [&]() -> int {
    return rc::launch_zoom_factor(a->mask_gt_observed, a->depth_rendered, a->src_pose, a->K,
                                  nullptr, a->zoom_factor, a->bbox, a->workspace,
                                  zt ? a->trans_gt : nullptr, zt ? a->zoom_trans_gt : nullptr, B,
                                  1, H, W, RC_ZF_MASK, a->stream);
  };
  // sparse frame: the rendered box and the GT flow from the renderer's tile items (frame.cu)
  const bool sparse = a->frame_renderer != nullptr;
  rc::FrameItems F;
  bool flow_tiles = false;  // flow on the tiles only (else the dense kernel writes the planes)
  if (sparse && (ph_frame || ph_sample)) {
    RC_REQUIRE(a->observed_box != nullptr, RC_ERR_INVALID_ARG);
    if ((st = rc::renderer_frame_items(a->frame_renderer, &F)) != RC_OK) return st;
    RC_REQUIRE(F.H == H && F.W == W && B <= F.max_batch, RC_ERR_INVALID_ARG);
    flow_tiles = do_flow && a->outputs_track_frames != 0 && F.prev_list != nullptr;
  }
  auto dense_flow = [&]() -> int {
    return rc_flow_dev(a->depth_rendered, a->depth_gt_observed, a->KT, a->Kinv, a->flow, a->valid,
                       B, H, W, a->stream);
  };
  auto tiles = [&](bool with_boxes) -> int {  // rendered box (+ flow) over the frame's tiles
    rc::FrameFinal fin = {a->observed_box, a->src_pose, a->K, a->zoom_factor, a->bbox,
                          zt ? a->trans_gt : nullptr, zt ? a->zoom_trans_gt : nullptr};
    return rc::launch_frame_tiles(F, a->depth_rendered, a->depth_gt_observed, a->KT, a->Kinv,
                                  flow_tiles ? a->flow : nullptr, flow_tiles ? a->valid : nullptr,
                                  a->workspace, B, H, W, flow_tiles, with_boxes ? &fin : nullptr,
                                  a->stream);
  };
  auto boxes = [&]() -> int {  // zoom factor (+ ZoomTrans) from the two boxes
    return rc::launch_zoom_factor_boxes(a->observed_box, a->src_pose, a->K, a->zoom_factor,
                                        a->bbox, a->workspace, zt ? a->trans_gt : nullptr,
                                        zt ? a->zoom_trans_gt : nullptr, B, H, W, a->stream);
  };
  struct PdlScope {  // launches inside the scope may start before their predecessor ends
    explicit PdlScope(bool on) { rc::tl_pdl = on; }
    ~PdlScope() { rc::tl_pdl = false; }
  };
  if (!serial) {
    // loss (normal launch: waits for everything before it) -> bbox -> flow, the last two
    // programmatically serialised, so all three are resident together; the resampler is a
    // normal launch again and therefore waits for all of them.
    // backward first (normal launch, releases its dependents at once), then the forward loss
    // beside it: neither reads the other's output
    if (do_bwd && (st = loss_bwd()) != RC_OK) return st;
    if (do_loss) {
      PdlScope scope(do_bwd);
      if ((st = loss()) != RC_OK) return st;
    }
    if (!ph_frame) {
      // (the caller issues the frame phase)
    } else if (!sparse) {
      {
        PdlScope scope(do_loss || do_bwd);
        if ((st = bbox()) != RC_OK) return st;
      }
      if (do_flow) {
        PdlScope scope(true);
        if ((st = dense_flow()) != RC_OK) return st;
      }
    } else {
      // [dense flow when the planes' contents are unknown] -> tile pass -> zoom factor; the
      // zoom-factor kernel waits (griddepcontrol.wait) for the tile pass right before it
      bool first = !(do_loss || do_bwd);
      if (do_flow && !flow_tiles) {
        PdlScope scope(!first);
        if ((st = dense_flow()) != RC_OK) return st;
        first = false;
      }
      // small batches: the tile pass's last block also turns the boxes into zoom factors
      // (+ ZoomTrans) -- one launch less on a chain that is pure latency there (32 pairs: 200 ->
      // 195 us per iteration).  Larger batches keep the separate one-thread-per-pair kernel: a
      // single block working through 1024 pairs costs more (15 us) than that launch (5 us).
      const bool fused_boxes = B <= 64;
      {
        PdlScope scope(!first);
        if ((st = tiles(fused_boxes)) != RC_OK) return st;
      }
      if (!fused_boxes) {
        PdlScope scope(true);
        if ((st = boxes()) != RC_OK) return st;
      }
    }
  } else {
    if ((st = rc::mark(a->marks, 0, s)) != RC_OK) return st;
    if (ph_frame && do_flow && !flow_tiles && (st = dense_flow()) != RC_OK) return st;
    if (ph_frame && sparse && (st = tiles(false)) != RC_OK) return st;
    if ((st = rc::mark(a->marks, 1, s)) != RC_OK) return st;
    if (ph_frame && (st = sparse ? boxes() : bbox()) != RC_OK) return st;
  }
  if ((st = rc::mark(a->marks, 2, s)) != RC_OK) return st;
  if (ph_sample) {
  // (3) one resampler launch: 3 masks + 6 image channels + 2 flow + 2 flow-weight planes
  rc_zoom_group g[9];
  auto set = [&](int i, const float *src, float *dst, int ch, int pre, int post, int flags) {
    g[i].src = src;
    g[i].dst = dst;
    g[i].channels = ch;
    g[i].pre_op = pre;
    g[i].post_op = post;
    g[i].flags = flags;
    for (int c = 0; c < 4; ++c) g[i].mean[c] = a->pixel_means[c];
  };
  int n = 0;
  if (a->zoom_mask_observed)
    set(n++, a->mask_observed, a->zoom_mask_observed, 1, RC_PRE_NONE, RC_POST_ROUND, 0);
  if (a->zoom_mask_gt_observed)
    set(n++, a->mask_gt_observed, a->zoom_mask_gt_observed, 1, RC_PRE_NONE, RC_POST_ROUND, 0);
  // rendered-side planes whose background maps to a constant: skippable per output tile
  const int sp = (sparse && a->zoom_tile_state) ? RC_ZOOM_SRC_FRAME_SPARSE : 0;
  if (a->zoom_mask_rendered)
    set(n++, a->depth_rendered, a->zoom_mask_rendered, 1, RC_PRE_THRESH02, RC_POST_ROUND, sp);
  if (a->zoom_image_observed)
    set(n++, a->image_observed, a->zoom_image_observed, 3, RC_PRE_ADD_MEAN, RC_POST_SUB_MEAN, 0);
  if (a->zoom_image_rendered)
    set(n++, a->image_rendered, a->zoom_image_rendered, 3, RC_PRE_ADD_MEAN, RC_POST_SUB_MEAN, 0);
  if (do_flow && a->zoom_flow) set(n++, a->flow, a->zoom_flow, 2, RC_PRE_NONE, RC_POST_DIV_WX, sp);
  if (do_flow && a->zoom_flow_weights)
    set(n++, a->valid, a->zoom_flow_weights, 2, RC_PRE_NONE, RC_POST_ROUND_M045,
        RC_ZOOM_SRC_BROADCAST | sp);
  // INPUT_DEPTH (deepIM_flownet.py:669-681): ZoomDepth's two planes ride in the same launch
  if (a->depth_observed && a->zoom_depth_observed)
    set(n++, a->depth_observed, a->zoom_depth_observed, 1, RC_PRE_NONE, RC_POST_NONE, 0);
  if (a->zoom_depth_rendered)
    set(n++, a->depth_rendered, a->zoom_depth_rendered, 1, RC_PRE_NONE, RC_POST_NONE, sp);
  rc::ZoomSparse zs;
  if (sp) {
    zs.tile_count = F.tile_count;
    zs.tile_state = a->zoom_tile_state;
    zs.tiles_x = F.tiles_x;
    zs.tiles_y = F.tiles_y;
    zs.state_valid = a->outputs_track_frames != 0;
  }
  st = rc::launch_zoom_sample(a->zoom_factor, 0, g, n, B, H, W, sp ? &zs : nullptr, a->stream);
  if (st != RC_OK) return st;
  }  // ph_sample
  if ((st = rc::mark(a->marks, 3, s)) != RC_OK) return st;
  if ((st = rc::mark(a->marks, 4, s)) != RC_OK) return st;
  // Transform3D + point-matching loss (already issued when overlapped)
  if (serial && do_loss && (st = loss()) != RC_OK) return st;
  if (serial && do_bwd && (st = loss_bwd()) != RC_OK) return st;
  // invZoomTrans.backward (zoom_trans.py:50-76): needs this step's zoom factor
  if (ph_ztb && N > 0 && a->grad_rot && a->grad_trans && a->grad_zoom_trans) {
    st = rc_zoom_trans_dev(a->zoom_factor, a->grad_trans, a->grad_zoom_trans, B, 1,
                           a->zoom_grad ? 1 : 0, 0, a->stream);
    if (st != RC_OK) return st;
  }
  if ((st = rc::mark(a->marks, 5, s)) != RC_OK) return st;
  return rc::note_status(RC_OK);
}

int rc_event_create(void **event) {
  RC_REQUIRE(event != nullptr, RC_ERR_INVALID_ARG);
  cudaEvent_t e;
  RC_CUDA_TRY(cudaEventCreate(&e));
  *event = e;
  return rc::note_status(RC_OK);
}

int rc_event_destroy(void *event) {
  if (event) RC_CUDA_TRY(cudaEventDestroy(static_cast<cudaEvent_t>(event)));
  return rc::note_status(RC_OK);
}

int rc_event_record(void *event, void *stream) {
  RC_REQUIRE(event != nullptr, RC_ERR_INVALID_ARG);
  RC_CUDA_TRY(rc::record_mark(static_cast<cudaEvent_t>(event), static_cast<cudaStream_t>(stream)));
  return rc::note_status(RC_OK);
}

int rc_event_synchronize(void *event) {
  RC_REQUIRE(event != nullptr, RC_ERR_INVALID_ARG);
  RC_CUDA_TRY(cudaEventSynchronize(static_cast<cudaEvent_t>(event)));
  return rc::note_status(RC_OK);
}

int rc_event_elapsed_ms(void *start, void *end, float *ms) {
  RC_REQUIRE(start && end && ms, RC_ERR_INVALID_ARG);
  RC_CUDA_TRY(cudaEventElapsedTime(ms, static_cast<cudaEvent_t>(start),
                                   static_cast<cudaEvent_t>(end)));
  return rc::note_status(RC_OK);
}

}  // extern "C"
