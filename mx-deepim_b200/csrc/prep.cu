// prep.cu -- the callers either side of the compare step:
// test-time mask policies of update_data_batch and the iteration-0 input preparation.
//
//   /root/reference/lib/pair_matching/data_pair.py:101-131   box_rendered / box_observed masks
//   /root/reference/lib/utils/image.py                       image transform, depth decode
//
// All kernels are streaming passes (128-bit accesses, one resident wave or a plain grid);
// nothing here allocates or synchronises.
#include <limits.h>

#include "common.cuh"

namespace rc {

// ------------------------------------------------------------- mask box policy
// workspace per sample (int32 x RC_BOX_WS_INTS): [0] INT_MAX - x_min, [1] x_max + 1,
// [2] INT_MAX - y_min, [3] y_max + 1 (all merged with atomicMax, 0 = empty), [4] ticket.
// Same all-zero-when-idle encoding as the zoom-factor workspace: the finalising block reads
// with atomicExch(.., 0), so no init launch or memset is ever needed.
__global__ void __launch_bounds__(256)
mask_bbox_kernel(const float *__restrict__ mask, int32_t *__restrict__ bbox,
                 int32_t *__restrict__ ws, int H, int W) {
  const int b = blockIdx.y;
  const int quads = (H * W) >> 2;
  const float4 *m = reinterpret_cast<const float4 *>(mask + (size_t)b * H * W);
  int x0 = INT_MAX, x1 = -1, y0 = INT_MAX, y1 = -1;
  const int stride = gridDim.x * blockDim.x;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < quads; q += stride) {
    const float4 v = ld_stream_f4(m + q);
    // np.nonzero of the column / row maxima (data_pair.py:107-110): a column counts when
    // its maximum is non-zero; for the non-negative masks of this path that is "any pixel
    // non-zero", and negative-only columns (maximum 0 or below) are excluded like numpy does
    // only when the maximum is exactly zero -- masks are 0/1 here.
    const unsigned nz = (v.x != 0.f ? 1u : 0u) | (v.y != 0.f ? 2u : 0u) | (v.z != 0.f ? 4u : 0u) |
                        (v.w != 0.f ? 8u : 0u);
    if (nz) {
      const int p = q << 2;
      const int y = p / W, x = p - y * W;
      x0 = min(x0, x + __ffs(nz) - 1);
      x1 = max(x1, x + 31 - __clz(nz));
      y0 = min(y0, y);
      y1 = max(y1, y);
    }
  }
  x0 = warp_min(x0);
  x1 = warp_max(x1);
  y0 = warp_min(y0);
  y1 = warp_max(y1);
  __shared__ int sm[8][4];
  __shared__ int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    sm[warp][0] = x0;
    sm[warp][1] = x1;
    sm[warp][2] = y0;
    sm[warp][3] = y1;
  }
  __syncthreads();
  int32_t *w = ws + b * RC_BOX_WS_INTS;
  if (threadIdx.x < 4) {
    const int k = threadIdx.x;
    int acc = sm[0][k];
    for (int i = 1; i < (int)(blockDim.x >> 5); ++i)
      acc = (k & 1) ? max(acc, sm[i][k]) : min(acc, sm[i][k]);
    if (k & 1) {
      if (acc >= 0) atomicMax(w + k, acc + 1);
    } else {
      if (acc != INT_MAX) atomicMax(w + k, INT_MAX - acc);
    }
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(w + 4, 1) == (int)gridDim.x - 1);
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    int e[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) e[k] = atomicExch(w + k, 0);
    atomicExch(w + 4, 0);
    // x_start, x_end, y_start, y_end as np.min / np.max of the non-zero indices; an empty
    // mask is reported as x_start > x_end (the reference raises there)
    bbox[b * 4 + 0] = e[0] ? INT_MAX - e[0] : INT_MAX;
    bbox[b * 4 + 1] = e[1] - 1;
    bbox[b * 4 + 2] = e[2] ? INT_MAX - e[2] : INT_MAX;
    bbox[b * 4 + 3] = e[3] - 1;
  }
}

// mask_observed[y_start:y_end, x_start:x_end] = 1.0 -- the slice excludes the last row and
// column of the box, exactly as the reference writes it (data_pair.py:115, :130).
__global__ void __launch_bounds__(256)
mask_box_fill_kernel(const int32_t *__restrict__ bbox, float *__restrict__ out, int H, int W) {
  const int b = blockIdx.y;
  const int quads = (H * W) >> 2;
  const int xs = bbox[b * 4 + 0], xe = bbox[b * 4 + 1], ys = bbox[b * 4 + 2], ye = bbox[b * 4 + 3];
  float4 *o = reinterpret_cast<float4 *>(out + (size_t)b * H * W);
  const int stride = gridDim.x * blockDim.x;
  for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < quads; q += stride) {
    const int p = q << 2;
    const int y = p / W, x = p - y * W;
    const bool yin = (y >= ys) && (y < ye);
    float4 v;
    v.x = (yin && x + 0 >= xs && x + 0 < xe) ? 1.f : 0.f;
    v.y = (yin && x + 1 >= xs && x + 1 < xe) ? 1.f : 0.f;
    v.z = (yin && x + 2 >= xs && x + 2 < xe) ? 1.f : 0.f;
    v.w = (yin && x + 3 >= xs && x + 3 < xe) ? 1.f : 0.f;
    st_f4(o + q, v);
  }
}


// ------------------------------------------------------------- input preparation
// image.transform (lib/utils/image.py:709-720): BGR HWC uint8 -> RGB CHW, minus PIXEL_MEANS
// (given in B,G,R order as the config holds them): out[i] = im[:, :, 2 - i] - means[2 - i],
// evaluated in float64 like the reference's np.zeros((1,3,H,W)) tensor, stored float32.
// One thread = 4 pixels: three 32-bit loads (12 bytes), three 128-bit stores.
__global__ void __launch_bounds__(256)
prep_image_kernel(const uint8_t *__restrict__ bgr, float *__restrict__ out, double m0, double m1,
                  double m2, int plane_quads) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= plane_quads) return;
  const uint32_t *src = reinterpret_cast<const uint32_t *>(bgr + (size_t)b * plane_quads * 12) + (size_t)q * 3;
  const uint32_t w0 = __ldg(src), w1 = __ldg(src + 1), w2 = __ldg(src + 2);
  // bytes: B0 G0 R0 B1 | G1 R1 B2 G2 | R2 B3 G3 R3
  const int B_[4] = {(int)(w0 & 255u), (int)(w0 >> 24), (int)((w1 >> 16) & 255u), (int)((w2 >> 8) & 255u)};
  const int G_[4] = {(int)((w0 >> 8) & 255u), (int)(w1 & 255u), (int)(w1 >> 24), (int)((w2 >> 16) & 255u)};
  const int R_[4] = {(int)((w0 >> 16) & 255u), (int)((w1 >> 8) & 255u), (int)(w2 & 255u), (int)(w2 >> 24)};
  float4 *o = reinterpret_cast<float4 *>(out + (size_t)b * 3 * plane_quads * 4) + q;
  // channel 0 = R - means[2], 1 = G - means[1], 2 = B - means[0]
  st_f4(o, make_float4((float)((double)R_[0] - m2), (float)((double)R_[1] - m2),
                       (float)((double)R_[2] - m2), (float)((double)R_[3] - m2)));
  st_f4(o + plane_quads, make_float4((float)((double)G_[0] - m1), (float)((double)G_[1] - m1),
                                     (float)((double)G_[2] - m1), (float)((double)G_[3] - m1)));
  st_f4(o + 2 * (size_t)plane_quads, make_float4((float)((double)B_[0] - m0), (float)((double)B_[1] - m0),
                                                 (float)((double)B_[2] - m0), (float)((double)B_[3] - m0)));
}

// Depth decode (lib/utils/image.py:200-206, :229-260): uint16 PNG -> float32, optionally
// zeroed outside the instance (depth *= mask == mask_idx, :244), then / DEPTH_FACTOR in
// float32.  One thread = 4 pixels.
__global__ void __launch_bounds__(256)
prep_depth_kernel(const uint16_t *__restrict__ raw, const uint8_t *__restrict__ mask,
                  const int32_t *__restrict__ mask_idx, float *__restrict__ out, float factor,
                  int plane_quads) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= plane_quads) return;
  const uint2 w = __ldg(reinterpret_cast<const uint2 *>(raw + (size_t)b * plane_quads * 4) + q);
  float v[4] = {(float)(w.x & 0xffffu), (float)(w.x >> 16), (float)(w.y & 0xffffu), (float)(w.y >> 16)};
  if (mask) {
    const uint32_t mw = __ldg(reinterpret_cast<const uint32_t *>(mask + (size_t)b * plane_quads * 4) + q);
    const int want = mask_idx[b];
#pragma unroll
    for (int k = 0; k < 4; ++k) v[k] = __fmul_rn(v[k], ((int)((mw >> (8 * k)) & 255u) == want) ? 1.f : 0.f);
  }
  st_f4(reinterpret_cast<float4 *>(out + (size_t)b * plane_quads * 4) + q,
        make_float4(__fdiv_rn(v[0], factor), __fdiv_rn(v[1], factor), __fdiv_rn(v[2], factor),
                    __fdiv_rn(v[3], factor)));
}

// Mask decode (lib/utils/image.py:300-304): cur_mask[mask_png == mask_idx] = 1.0, else 0.
__global__ void __launch_bounds__(256)
prep_mask_kernel(const uint8_t *__restrict__ mask, const int32_t *__restrict__ mask_idx,
                 float *__restrict__ out, int plane_quads) {
  const int b = blockIdx.y;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= plane_quads) return;
  const uint32_t mw = __ldg(reinterpret_cast<const uint32_t *>(mask + (size_t)b * plane_quads * 4) + q);
  const int want = mask_idx[b];
  st_f4(reinterpret_cast<float4 *>(out + (size_t)b * plane_quads * 4) + q,
        make_float4((int)(mw & 255u) == want ? 1.f : 0.f, (int)((mw >> 8) & 255u) == want ? 1.f : 0.f,
                    (int)((mw >> 16) & 255u) == want ? 1.f : 0.f, (int)(mw >> 24) == want ? 1.f : 0.f));
}

// Mask from a decoded depth map: mask_rendered of get_pair_mask (lib/utils/image.py:487-492:
// `depth[depth > 0.2] = 1`, smaller values stay) or the binary test-time init mask (:446-447).
__global__ void __launch_bounds__(256)
mask_from_depth_kernel(const float *depth, float *out, float thresh,
                       int keep_small, long long quads) {
  const long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= quads) return;
  const float4 d = reinterpret_cast<const float4 *>(depth)[q];  // plain load: `out` may alias `depth`
  float4 o;
  o.x = d.x > thresh ? 1.f : (keep_small ? d.x : 0.f);
  o.y = d.y > thresh ? 1.f : (keep_small ? d.y : 0.f);
  o.z = d.z > thresh ? 1.f : (keep_small ? d.z : 0.f);
  o.w = d.w > thresh ? 1.f : (keep_small ? d.w : 0.f);
  st_f4(reinterpret_cast<float4 *>(out) + q, o);
}

// mask_dilate (lib/utils/mask_dilate.py:10-58) with the caller's random draws: t[4] per sample =
// thickness of the down / up / right / left growth (0 = that block is not taken).  Each block
// adds 1 where the pixel is 0 and the pixel `t` away in the opposite direction is not, and the
// result is clamped at 1; a non-zero pixel keeps its value (clamped at 1).
__global__ void __launch_bounds__(256)
mask_dilate_kernel(const float *__restrict__ mask, const int32_t *__restrict__ t4,
                   float *__restrict__ out, int H, int W) {
  const int b = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= H * W) return;
  const float *m = mask + (size_t)b * H * W;
  const int y = p / W, x = p - y * W;
  const int td = t4[b * 4 + 0], tu = t4[b * 4 + 1], tr = t4[b * 4 + 2], tl = t4[b * 4 + 3];
  const float v = __ldg(m + p);
  float o;
  if (v != 0.f) {
    o = v > 1.f ? 1.f : v;
  } else {
    bool grow = false;
    if (td > 0 && y >= td) grow |= __ldg(m + p - td * W) != 0.f;
    if (tu > 0 && y < H - tu) grow |= __ldg(m + p + tu * W) != 0.f;
    if (tr > 0 && x >= tr) grow |= __ldg(m + p - tr) != 0.f;
    if (tl > 0 && x < W - tl) grow |= __ldg(m + p + tl) != 0.f;
    o = grow ? 1.f : 0.f;
  }
  out[(size_t)b * H * W + p] = o;
}

// get_point_cloud_observed (lib/utils/image.py:592-596): R . X + T in float64, stored float32.
__global__ void __launch_bounds__(256)
points_transform_kernel(const float *__restrict__ pts, const float *__restrict__ pose,
                        float *__restrict__ out, int N) {
  const int b = blockIdx.y;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const float *P = pose + b * 12;
  const double x = pts[(size_t)b * 3 * N + n], y = pts[(size_t)b * 3 * N + N + n],
               z = pts[(size_t)b * 3 * N + 2 * N + n];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const double v = fma((double)P[r * 4 + 2], z, fma((double)P[r * 4 + 1], y, (double)P[r * 4] * x)) +
                     (double)P[r * 4 + 3];
    out[(size_t)b * 3 * N + r * N + n] = (float)v;
  }
}


// ------------------------------------------------------------- calc_flow (numpy rule)
// lib/pair_matching/flow.py:12-81 with backproject_camera (lib/utils/projection.py:47-67):
// float64 throughout, np.round (half to even) on the projected pixel, clamped look-up of the
// target depth, visible = depth_src != 0 && raw pixel inside && |d_tgt - z| < thresh &&
// |d_tgt| > 1e-10.  Dot products are accumulated left to right with FMA.
//   Kinv  [9]  float64 (np.linalg.inv(K))      transform [B,12] float64 (K . se3)
struct CalcFlowParams {
  double Kinv[9];
  double thresh;
  int standard_rep;
};

__device__ __forceinline__ void calc_flow_pixel(double d, int w, int h, const CalcFlowParams &P,
                                                const double *__restrict__ T,
                                                const float *__restrict__ depth_tgt, int H, int W,
                                                float &f0, float &f1, float &vis, double (&X)[3]) {
  const double xw = (double)w, yh = (double)h;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double r = fma(P.Kinv[i * 3 + 2], 1.0, fma(P.Kinv[i * 3 + 1], yh, P.Kinv[i * 3] * xw));
    X[i] = d * r;
  }
  double Xp[3];
#pragma unroll
  for (int i = 0; i < 3; ++i)
    Xp[i] = fma(T[i * 4 + 3], 1.0, fma(T[i * 4 + 2], X[2], fma(T[i * 4 + 1], X[1], T[i * 4] * X[0])));
  const double pz = Xp[2] + 1e-15;
  const double pw = Xp[0] / pz, ph = Xp[1] / pz;
  bool v = d != 0.0;
  const double rw = rint(pw), rh = rint(ph);  // np.round: half to even
  const bool inside = (rw >= 0.0) && (rw < (double)W) && (rh >= 0.0) && (rh < (double)H);
  const int cw = (int)fmin(fmax(rw, 0.0), (double)(W - 1));
  const int ch = (int)fmin(fmax(rh, 0.0), (double)(H - 1));
  double dt = 0.0;
  if (v) dt = (double)__ldg(depth_tgt + (size_t)ch * W + cw);
  v = v && inside && (fabs(dt - pz) < P.thresh) && (fabs(dt) > 1e-10);
  const double fw = pw - xw, fh = ph - yh;
  f0 = v ? (float)(P.standard_rep ? fw : fh) : 0.f;
  f1 = v ? (float)(P.standard_rep ? fh : fw) : 0.f;
  vis = v ? 1.f : 0.f;
}

// grid = (ceil(HW/256), B).  Also counts the visible pixels of every block (for X_valid).
__global__ void __launch_bounds__(256)
calc_flow_kernel(const float *__restrict__ depth_src, const float *__restrict__ depth_tgt,
                 const double *__restrict__ transform, CalcFlowParams P, float *__restrict__ flow,
                 float *__restrict__ visible, int32_t *__restrict__ block_counts, int H, int W) {
  const int b = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t plane = (size_t)H * W;
  float f0 = 0.f, f1 = 0.f, vis = 0.f;
  if (p < H * W) {
    const int h = p / W, w = p - h * W;
    double X[3];
    calc_flow_pixel((double)__ldg(depth_src + b * plane + p), w, h, P, transform + b * 12,
                    depth_tgt + b * plane, H, W, f0, f1, vis, X);
    flow[(size_t)b * 2 * plane + p] = f0;
    flow[(size_t)b * 2 * plane + plane + p] = f1;
    visible[b * plane + p] = vis;
  }
  if (block_counts) {
    const int c = __syncthreads_count(vis != 0.f);
    if (threadIdx.x == 0) block_counts[(size_t)b * gridDim.x + blockIdx.x] = c;
  }
}

// Exclusive scan of the per-block counts of one sample (one block per sample) -> offsets and
// the sample's total.
__global__ void __launch_bounds__(1024)
calc_flow_scan_kernel(int32_t *__restrict__ block_counts, int32_t *__restrict__ totals, int nblocks) {
  __shared__ int warp_sums[32];
  __shared__ int carry;
  const int b = blockIdx.x;
  int32_t *c = block_counts + (size_t)b * nblocks;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < nblocks; base += 1024) {
    const int i = base + threadIdx.x;
    const int v = i < nblocks ? c[i] : 0;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if ((threadIdx.x & 31) >= o) x += y;
    }
    if ((threadIdx.x & 31) == 31) warp_sums[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
      int s = warp_sums[threadIdx.x];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, s, o);
        if (threadIdx.x >= o) s += y;
      }
      warp_sums[threadIdx.x] = s;
    }
    __syncthreads();
    const int before = carry + ((threadIdx.x >> 5) ? warp_sums[(threadIdx.x >> 5) - 1] : 0);
    if (i < nblocks) c[i] = before + x - v;
    __syncthreads();
    if (threadIdx.x == 1023) carry = before + x;
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[b] = carry;
}

// X_valid (flow.py:80): the back-projected source points of the visible pixels, in pixel
// order, written as float32 [B,3,capacity].
__global__ void __launch_bounds__(256)
calc_flow_compact_kernel(const float *__restrict__ depth_src, const float *__restrict__ visible,
                         const int32_t *__restrict__ block_offsets, CalcFlowParams P,
                         float *__restrict__ x_valid, int capacity, int H, int W) {
  const int b = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t plane = (size_t)H * W;
  const bool v = p < H * W && visible[b * plane + p] != 0.f;
  const unsigned ballot = __ballot_sync(0xffffffffu, v);
  __shared__ int wbase[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) wbase[warp] = __popc(ballot);
  __syncthreads();
  int base = block_offsets[(size_t)b * gridDim.x + blockIdx.x];
  for (int i = 0; i < warp; ++i) base += wbase[i];
  if (v) {
    const int pos = base + __popc(ballot & ((1u << lane) - 1u));
    if (pos < capacity) {
      const int h = p / W, w = p - h * W;
      const double d = (double)depth_src[b * plane + p];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const double r = fma(P.Kinv[i * 3 + 2], 1.0,
                             fma(P.Kinv[i * 3 + 1], (double)h, P.Kinv[i * 3] * (double)w));
        x_valid[((size_t)b * 3 + i) * capacity + pos] = (float)(d * r);
      }
    }
  }
}

}  // namespace rc

extern "C" {

int rc_mask_box_dev(const float *mask, float *out, int32_t *bbox, int32_t *workspace,
                    int batch_size, int height, int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(mask && out && bbox && workspace, RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && rc::aligned16(mask) && rc::aligned16(out), RC_ERR_UNALIGNED);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int quads = height * width / 4;
  int bps = (quads + 256 * 8 - 1) / (256 * 8);  // eight quads per thread
  if (bps < 1) bps = 1;
  dim3 grid(bps, batch_size);
  rc::mask_bbox_kernel<<<grid, 256, 0, s>>>(mask, bbox, workspace, height, width);
  RC_CUDA_TRY(cudaGetLastError());
  rc::mask_box_fill_kernel<<<grid, 256, 0, s>>>(bbox, out, height, width);
  return rc::note_status(rc::check_launch());
}


int rc_prep_image_dev(const uint8_t *bgr_hwc, float *out_chw, const double *pixel_means_bgr,
                      int batch_size, int height, int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(bgr_hwc && out_chw && pixel_means_bgr, RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && rc::aligned16(out_chw) && ((uintptr_t)bgr_hwc & 3u) == 0,
             RC_ERR_UNALIGNED);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  const int quads = height * width / 4;
  dim3 grid((quads + 255) / 256, batch_size);
  rc::prep_image_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      bgr_hwc, out_chw, pixel_means_bgr[0], pixel_means_bgr[1], pixel_means_bgr[2], quads);
  return rc::note_status(rc::check_launch());
}

int rc_prep_depth_dev(const uint16_t *raw, const uint8_t *mask, const int32_t *mask_idx,
                      float *out, float depth_factor, int batch_size, int height, int width,
                      void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0 && depth_factor != 0.f, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(raw && out && ((mask == nullptr) == (mask_idx == nullptr)), RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && rc::aligned16(out) && ((uintptr_t)raw & 7u) == 0 &&
                 ((uintptr_t)mask & 3u) == 0,
             RC_ERR_UNALIGNED);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  const int quads = height * width / 4;
  dim3 grid((quads + 255) / 256, batch_size);
  rc::prep_depth_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(
      raw, mask, mask_idx, out, depth_factor, quads);
  return rc::note_status(rc::check_launch());
}

int rc_prep_mask_dev(const uint8_t *mask, const int32_t *mask_idx, float *out, int batch_size,
                     int height, int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(mask && mask_idx && out, RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && rc::aligned16(out) && ((uintptr_t)mask & 3u) == 0, RC_ERR_UNALIGNED);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  const int quads = height * width / 4;
  dim3 grid((quads + 255) / 256, batch_size);
  rc::prep_mask_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(mask, mask_idx, out, quads);
  return rc::note_status(rc::check_launch());
}

int rc_mask_from_depth_dev(const float *depth, float *out, float thresh, int keep_small,
                           int64_t count, void *stream) {
  RC_REQUIRE(count >= 0, RC_ERR_INVALID_ARG);
  if (count == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(depth && out, RC_ERR_INVALID_ARG);
  RC_REQUIRE(count % 4 == 0 && rc::aligned16(depth) && rc::aligned16(out), RC_ERR_UNALIGNED);
  const long long quads = count / 4;
  rc::mask_from_depth_kernel<<<(unsigned)((quads + 255) / 256), 256, 0,
                               static_cast<cudaStream_t>(stream)>>>(depth, out, thresh,
                                                                    keep_small ? 1 : 0, quads);
  return rc::note_status(rc::check_launch());
}

int rc_mask_dilate_dev(const float *mask, const int32_t *thickness, float *out, int batch_size,
                       int height, int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(mask && thickness && out && mask != out, RC_ERR_INVALID_ARG);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  dim3 grid((height * width + 255) / 256, batch_size);
  rc::mask_dilate_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(mask, thickness, out,
                                                                             height, width);
  return rc::note_status(rc::check_launch());
}

int rc_points_transform_dev(const float *points, const float *pose, float *out, int batch_size,
                            int num_points, void *stream) {
  RC_REQUIRE(batch_size >= 0 && num_points >= 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0 || num_points == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(points && pose && out, RC_ERR_INVALID_ARG);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  dim3 grid((num_points + 255) / 256, batch_size);
  rc::points_transform_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(points, pose, out,
                                                                                  num_points);
  return rc::note_status(rc::check_launch());
}

int rc_calc_flow_blocks(int height, int width) { return (height * width + 255) / 256; }

int rc_calc_flow_dev(const float *depth_src, const float *depth_tgt, const double *transform,
                     const double *Kinv_host, double thresh, int standard_rep, float *flow,
                     float *visible, float *x_valid, int32_t *counts, int32_t *workspace,
                     int capacity, int batch_size, int height, int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0 && capacity >= 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(depth_src && depth_tgt && transform && Kinv_host && flow && visible, RC_ERR_INVALID_ARG);
  const bool compact = x_valid != nullptr;
  RC_REQUIRE(!compact || (counts && workspace), RC_ERR_INVALID_ARG);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  rc::CalcFlowParams P;
  for (int i = 0; i < 9; ++i) P.Kinv[i] = Kinv_host[i];
  P.thresh = thresh;
  P.standard_rep = standard_rep ? 1 : 0;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int nblocks = rc_calc_flow_blocks(height, width);
  dim3 grid(nblocks, batch_size);
  rc::calc_flow_kernel<<<grid, 256, 0, s>>>(depth_src, depth_tgt, transform, P, flow, visible,
                                            compact ? workspace : nullptr, height, width);
  RC_CUDA_TRY(cudaGetLastError());
  if (compact) {
    rc::calc_flow_scan_kernel<<<batch_size, 1024, 0, s>>>(workspace, counts, nblocks);
    RC_CUDA_TRY(cudaGetLastError());
    rc::calc_flow_compact_kernel<<<grid, 256, 0, s>>>(depth_src, visible, workspace, P, x_valid,
                                                      capacity, height, width);
  }
  return rc::note_status(rc::check_launch());
}

}  // extern "C"
