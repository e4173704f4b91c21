/*
 * oracle/flow_oracle.c -- TEST INFRASTRUCTURE ONLY (never on the product path).
 *
 * CPU restatement of the reference GT-flow kernel
 *   /root/reference/lib/flow_c/gpu_flow_kernel.cu:32-69  (flow_kernel)
 * with the floating-point association that nvcc (default -fmad=true,
 * -prec-div=true) emits for it on sm_100a.  The association was read from the
 * SASS of the unmodified reference source (cuobjdump -sass, see DESIGN.md
 * "flow: expression tree"):
 *
 *   a   = fma(w, Ki0, h*Ki1) + Ki2            (FMUL, FFMA, FADD)
 *   x   = a * d                               (FMUL)
 *   xp  = fma(z, KT2, fma(x, KT0, y*KT1)) + KT3
 *   zp  = (float)((double)(fma(...)+KT11) + 1e-15)
 *   wp  = xp / zp   (IEEE correctly rounded division)
 *   round() on a float argument resolves to roundf (half away from zero)
 *   d > 1E-3 and |zp-dt| < 3E-3 are compared in double.
 *
 * Build with -ffp-contract=off so that the compiler adds no contraction of its
 * own: every fused operation below is an explicit fmaf().
 *
 * Parity pinning: the reference ships no golden vectors for this path
 * (SURVEY.md section 4).  This restatement is pinned (a) on the GPU box
 * against the reference .cu itself compiled for sm_100a (oracle/_ref,
 * tests/test_flow_gpu.py::test_flow_matches_reference_kernel, bit-exact) and
 * (b) against the reference numpy twin calc_flow through committed fixtures
 * (tests/golden/flow_calc_flow_*.npz) on pixels where the two rules agree.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#endif

/* one pixel, reference line numbers in gpu_flow_kernel.cu */
static inline void flow_pixel(int w, int h, int b, int height, int width,
                              const float *depth_src, const float *depth_tgt,
                              const float *KT_all, const float *Kinv,
                              float *flow, float *valid) {
  const int index = (b * height + h) * width + w;
  const float d_src = depth_src[index];                          /* :41 */
  const float fw = (float)w, fh = (float)h;
  const float ax = fmaf(fw, Kinv[0], fh * Kinv[1]) + Kinv[2];    /* :42 */
  const float ay = fmaf(fw, Kinv[3], fh * Kinv[4]) + Kinv[5];    /* :43 */
  const float x = ax * d_src;
  const float y = ay * d_src;
  const float z = d_src;                                         /* :44 */
  float f0 = 0.f, f1 = 0.f, v = 0.f;
  if ((double)d_src > 1E-3) {                                    /* :45 */
    const float *KT = KT_all + b * 12;                           /* :46 */
    const float x_proj = fmaf(z, KT[2], fmaf(x, KT[0], y * KT[1])) + KT[3];   /* :47 */
    const float y_proj = fmaf(z, KT[6], fmaf(x, KT[4], y * KT[5])) + KT[7];   /* :48 */
    const float z_sum = fmaf(z, KT[10], fmaf(x, KT[8], y * KT[9])) + KT[11];
    const float z_proj = (float)((double)z_sum + 1E-15);         /* :49 */
    const float w_proj = x_proj / z_proj;                        /* :50 */
    const float h_proj = y_proj / z_proj;                        /* :51 */
    const int w_proj_i = (int)roundf(w_proj);                    /* :52 */
    const int h_proj_i = (int)roundf(h_proj);                    /* :53 */
    if (w_proj >= 0 && w_proj <= (float)(width - 1) && h_proj >= 0 &&
        h_proj <= (float)(height - 1)) {                         /* :54 */
      const float d_tgt = depth_tgt[(b * height + h_proj_i) * width + w_proj_i]; /* :55 */
      if ((double)fabsf(z_proj - d_tgt) < 3E-3) {                /* :56 */
        f0 = h_proj - fh;                                        /* :57 */
        f1 = w_proj - fw;                                        /* :58 */
        v = 1.f;                                                 /* :59 */
      }
    }
  }
  flow[((b * 2 + 0) * height + h) * width + w] = f0;             /* :57,:64 */
  flow[((b * 2 + 1) * height + h) * width + w] = f1;             /* :58,:65 */
  valid[index] = v;                                              /* :59,:66 */
}

/* Same argument order as the reference host entry point
 *   void _flow(float* flow, float* valid, float* depth_src, float* depth_tgt,
 *              float* KT, float* Kinv, int batch_size, int height, int width,
 *              int device_id)            (lib/flow_c/gpu_flow.hpp:1-3)
 * minus device_id, plus a thread count (0 = all). */
void oracle_flow(float *flow, float *valid, const float *depth_src,
                 const float *depth_tgt, const float *KT, const float *Kinv,
                 int batch_size, int height, int width, int num_threads) {
#ifdef _OPENMP
  if (num_threads > 0) omp_set_num_threads(num_threads);
#endif
  const int rows = batch_size * height;
#pragma omp parallel for schedule(static)
  for (int r = 0; r < rows; ++r) {
    const int b = r / height, h = r % height;
    for (int w = 0; w < width; ++w)
      flow_pixel(w, h, b, height, width, depth_src, depth_tgt, KT, Kinv, flow, valid);
  }
}

int oracle_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
