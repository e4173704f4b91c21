// zoom_finalize.cuh -- the crop-box workspace layout and the zoom factor computed from the two
// boxes (deepim/operator_py/zoom_mask.py:63-119), shared by zoom.cu (dense crop-box kernel, the
// boxes-only kernel) and frame.cu (the sparse frame pass finalises in its last block).
#pragma once
#include <limits.h>

#include "common.cuh"

namespace rc {

// workspace layout per sample (int32 x 10):
//   [0..3] obs x0,x1,y0,y1   [4..7] rend x0,x1,y0,y1   [8] ticket  [9] unused
// Entries are kept in an all-zero-when-idle encoding so that no init launch is needed:
// minima are stored as INT_MAX - x and maxima as x + 1, both merged with atomicMax
// (0 = empty set); the finalising block reads them with atomicExch(.., 0), which leaves
// the workspace zeroed for the next call.
#define RC_ZF_WS_STRIDE RC_ZF_WS_INTS

// Zoom factor from the two boxes; float64 scalar arithmetic as numpy 1.x
// evaluates zoom_mask.py:63-105 (np.float32 scalar op Python/int64 scalar ->
// float64), stored as float32 (:116-119).
__device__ __forceinline__ void zoom_factor_finalize(const int32_t *w, const float *__restrict__ src_pose,
                                     const float *__restrict__ K, float *__restrict__ zf,
                                     int32_t *__restrict__ bbox, int H, int W) {
  const int ox0 = w[0], ox1 = w[1], oy0 = w[2], oy1 = w[3];
  int rx0 = w[4], rx1 = w[5], ry0 = w[6], ry1 = w[7];
  if (bbox) {
#pragma unroll
    for (int k = 0; k < 8; ++k) bbox[k] = w[k];
  }
  if (ox1 < 0) {  // observed set empty: the reference raises (np.min of empty), zoom_mask.py:59
    const float nanv = __int_as_float(0x7fc00000);
    zf[0] = zf[1] = zf[2] = zf[3] = nanv;
    return;
  }
  // np.dot(K, src_pose[:, 3]) in float32 (zoom_mask.py:72)
  const float t0 = src_pose[3], t1 = src_pose[7], t2 = src_pose[11];
  const float c0 = __fadd_rn(__fadd_rn(__fmul_rn(K[0], t0), __fmul_rn(K[1], t1)), __fmul_rn(K[2], t2));
  const float c1 = __fadd_rn(__fadd_rn(__fmul_rn(K[3], t0), __fmul_rn(K[4], t1)), __fmul_rn(K[5], t2));
  const float c2 = __fadd_rn(__fadd_rn(__fmul_rn(K[6], t0), __fmul_rn(K[7], t1)), __fmul_rn(K[8], t2));
  double zcx, zcy;
  if (rx1 < 0) {  // "NO POINT VALID IN MASK rendered" fallback, zoom_mask.py:74-81
    rx0 = ox0;
    rx1 = ox1;
    ry0 = oy0;
    ry1 = oy1;
    zcx = (double)(ox0 + ox1) * 0.5;
    zcy = (double)(oy0 + oy1) * 0.5;
  } else {
    zcx = (double)__fdiv_rn(c0, c2);  // :73
    zcy = (double)__fdiv_rn(c1, c2);
  }
  const double left = fmax(zcx - rx0, zcx - ox0);
  const double right = fmax(rx1 - zcx, ox1 - zcx);
  const double up = fmax(zcy - ry0, zcy - oy0);
  const double down = fmax(oy1 - zcy, ry1 - zcy);
  const double m = fmax(fmax(__dmul_rn(0.75, right), __dmul_rn(0.75, left)), fmax(up, down));
  const double crop_h = __dmul_rn(__dmul_rn(m, 1.4), 2.0);  // :96-100
  const double wx = crop_h / (double)H;                     // :102-103 (both / height)
  const double tx = __dadd_rn(__dmul_rn(zcx / (double)W, 2.0), -1.0);  // :104
  const double ty = __dadd_rn(__dmul_rn(zcy / (double)H, 2.0), -1.0);  // :105
  zf[0] = (float)wx;
  zf[1] = (float)wx;
  zf[2] = (float)tx;
  zf[3] = (float)ty;
}


// One pair's zoom factor from the observed box and the rendered box left in the workspace entries
// 4..7 by atomicMax merges (read with atomicExch, which re-arms them), + the fused ZoomTrans
// forward (zoom_trans.py:30-44) when trans_out is given.
__device__ __forceinline__ void zoom_factor_from_boxes(int b, const int32_t *__restrict__ observed_box,
                                                       int32_t *__restrict__ ws,
                                                       const float *__restrict__ src_pose,
                                                       const float *__restrict__ K,
                                                       float *__restrict__ zoom_factor,
                                                       int32_t *__restrict__ bbox,
                                                       const float *__restrict__ trans_in,
                                                       float *__restrict__ trans_out, int H, int W) {
  int32_t *w = ws + b * RC_ZF_WS_STRIDE;
  int32_t fin[8];
#pragma unroll
  for (int k = 0; k < 4; ++k) fin[k] = observed_box[b * 4 + k];
#pragma unroll
  for (int k = 4; k < 8; ++k) {
    const int e = atomicExch(w + k, 0);
    fin[k] = (k & 1) ? e - 1 : INT_MAX - e;
  }
  zoom_factor_finalize(fin, src_pose + b * 12, K, zoom_factor + b * 4,
                       bbox ? bbox + b * 8 : nullptr, H, W);
  if (trans_out) {
    const float wx = zoom_factor[b * 4];
    trans_out[b * 3 + 0] = __fdiv_rn(trans_in[b * 3 + 0], wx);
    trans_out[b * 3 + 1] = __fdiv_rn(trans_in[b * 3 + 1], wx);
    trans_out[b * 3 + 2] = trans_in[b * 3 + 2];
  }
}

}  // namespace rc
