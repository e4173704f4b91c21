// flow_device.cuh -- the per-pixel GT flow rule shared by flow.cu (whole planes) and frame.cu
// (the rendered frame's tiles only).
#pragma once
#include "common.cuh"

namespace rc {

struct FlowCoef {
  float kt[12];
  float ki[6];
};

__device__ __forceinline__ void flow_pixel(float d, float fw, float fh, const FlowCoef &c,
                                           const float *__restrict__ tgt_plane, int H,
                                           int W, float &f0, float &f1, float &v) {
  f0 = 0.f;
  f1 = 0.f;
  v = 0.f;
  if ((double)d > 1E-3) {  // gpu_flow_kernel.cu:45 (double compare)
    const float ax = __fadd_rn(__fmaf_rn(fw, c.ki[0], __fmul_rn(fh, c.ki[1])), c.ki[2]);
    const float ay = __fadd_rn(__fmaf_rn(fw, c.ki[3], __fmul_rn(fh, c.ki[4])), c.ki[5]);
    const float x = __fmul_rn(ax, d);  // :42
    const float y = __fmul_rn(ay, d);  // :43
    const float z = d;                 // :44
    const float xp = __fadd_rn(
        __fmaf_rn(z, c.kt[2], __fmaf_rn(x, c.kt[0], __fmul_rn(y, c.kt[1]))), c.kt[3]);
    const float yp = __fadd_rn(
        __fmaf_rn(z, c.kt[6], __fmaf_rn(x, c.kt[4], __fmul_rn(y, c.kt[5]))), c.kt[7]);
    const float zs = __fadd_rn(
        __fmaf_rn(z, c.kt[10], __fmaf_rn(x, c.kt[8], __fmul_rn(y, c.kt[9]))), c.kt[11]);
    const float zp = (float)((double)zs + 1E-15);  // :49
    const float wp = __fdiv_rn(xp, zp);            // :50
    const float hp = __fdiv_rn(yp, zp);            // :51
    if (wp >= 0.f && wp <= (float)(W - 1) && hp >= 0.f && hp <= (float)(H - 1)) {  // :54
      const int wi = (int)roundf(wp);  // :52
      const int hi = (int)roundf(hp);  // :53
      const float dt = __ldg(tgt_plane + hi * W + wi);  // :55
      if ((double)fabsf(__fsub_rn(zp, dt)) < 3E-3) {    // :56
        f0 = __fsub_rn(hp, fh);  // :57  channel 0 = delta row
        f1 = __fsub_rn(wp, fw);  // :58  channel 1 = delta col
        v = 1.f;                 // :59
      }
    }
  }
}

__device__ __forceinline__ void load_coef(FlowCoef &c, const float *__restrict__ KT,
                                          const float *__restrict__ Kinv, int b) {
#pragma unroll
  for (int i = 0; i < 12; ++i) c.kt[i] = __ldg(KT + b * 12 + i);
#pragma unroll
  for (int i = 0; i < 6; ++i) c.ki[i] = __ldg(Kinv + i);
}

}  // namespace rc
