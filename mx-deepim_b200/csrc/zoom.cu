// zoom.cu -- crop-box / zoom-factor reduction and the fused bilinear resampler.
//
// Replaces the host-side numpy blocks and the MXNet GridGenerator(affine) +
// BilinearSampler calls of
//   /root/reference/deepim/operator_py/zoom_mask.py:29-134
//   /root/reference/deepim/operator_py/zoom_image.py:26-119
//   /root/reference/deepim/operator_py/zoom_image_with_factor.py:31-75
//   /root/reference/deepim/operator_py/zoom_depth.py:24-50
//   /root/reference/deepim/operator_py/zoom_flow.py:28-77
//   /root/reference/deepim/operator_py/zoom_mask_with_factor.py:29-68
//   /root/reference/deepim/operator_py/zoom_trans.py:22-76
//
// B200 design.  (1) bbox: one streaming pass with 128-bit loads, per-thread
// min/max, warp-shuffle + shared-memory block reduction, eight integer atomics
// per block; the last block of a sample (ticket counter) computes the zoom
// factor in float64 exactly as the numpy scalars do, so there is no second
// launch and no host round trip.  (2) sampler: all planes that share a grid
// (3 masks + 6 image channels + 2 flow + 2 weight planes in the default
// configuration) are resampled by ONE launch; a thread owns four adjacent
// output pixels of a row, computes the source coordinates once, and for each
// plane gathers the taps through the read-only path (zoom-in means the source
// window is smaller than the output tile, so the taps are L1/L2 hits) and
// issues one 128-bit store.  Write traffic dominates: 52 B/px out of 60.
#include <limits.h>

#include "common.cuh"

namespace rc {

// ------------------------------------------------------------------- bbox
struct BBoxAcc {
  int x0, x1, y0, y1;
  __device__ __forceinline__ void init() {
    x0 = INT_MAX;
    y0 = INT_MAX;
    x1 = -1;
    y1 = -1;
  }
  __device__ __forceinline__ void add(int x, int y) {
    x0 = min(x0, x);
    x1 = max(x1, x);
    y0 = min(y0, y);
    y1 = max(y1, y);
  }
};

__device__ __forceinline__ float comp(const float4 &v, int k) {
  return k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w));
}

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
__device__ void zoom_factor_finalize(const int32_t *w, const float *__restrict__ src_pose,
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

// grid = (blocks_per_sample, B), block = 256.
// ONE: single-channel inputs (the mask configuration): no per-pixel channel sums to carry.
template <int MODE, bool ONE>
__global__ void __launch_bounds__(256, ONE ? 4 : 2)
zoom_bbox_kernel(const float *__restrict__ obs, const float *__restrict__ rend,
                 const float *__restrict__ src_pose, const float *__restrict__ K,
                 const float *__restrict__ pixel_means, float *__restrict__ zoom_factor,
                 int32_t *__restrict__ bbox, int32_t *__restrict__ ws,
                 const float *__restrict__ trans_in, float *__restrict__ trans_out, int C, int H,
                 int W) {
  pdl_launch_dependents();
  const int b = blockIdx.y;
  const int quads = (H * W) >> 2;
  const size_t plane = (size_t)H * W;
  const float *o = obs + (size_t)b * C * plane;
  const float *r = rend + (size_t)b * C * plane;
  BBoxAcc ao, ar;
  ao.init();
  ar.init();
  const float thr = (MODE == RC_ZF_IMAGE) ? 0.01f : 0.3f;  // zoom_image.py:35 / zoom_mask.py:37
  const int stride = gridDim.x * blockDim.x;
  constexpr int UNROLL = 4;  // the launch gives every thread at most UNROLL quads: all loads in flight
  for (int q0 = blockIdx.x * blockDim.x + threadIdx.x; q0 < quads; q0 += UNROLL * stride) {
    float4 so[UNROLL], sr[UNROLL];
    const int nc = ONE ? 1 : C;
    for (int c = 0; c < nc; ++c) {
      const float mean = (MODE == RC_ZF_IMAGE) ? __ldg(pixel_means + c) : 0.f;
      float4 vo[UNROLL], vr[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int q = q0 + u * stride;
        if (q < quads) {
          vo[u] = ld_stream_f4(reinterpret_cast<const float4 *>(o + c * plane) + q);
          vr[u] = ld_stream_f4(reinterpret_cast<const float4 *>(r + c * plane) + q);
        } else {
          vo[u] = make_float4(0.f, 0.f, 0.f, 0.f);
          vr[u] = (MODE == RC_ZF_IMAGE) ? make_float4(-mean, -mean, -mean, -mean)
                                        : make_float4(0.f, 0.f, 0.f, 0.f);
          if (MODE == RC_ZF_IMAGE) vo[u] = vr[u];
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        float eo[4] = {vo[u].x, vo[u].y, vo[u].z, vo[u].w};
        float er[4] = {vr[u].x, vr[u].y, vr[u].z, vr[u].w};
        float ao_[4], ar_[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (MODE == RC_ZF_IMAGE) {
            eo[k] = __fadd_rn(eo[k], mean);  // zoom_image.py:31-32
            er[k] = __fadd_rn(er[k], mean);
          } else {
            er[k] = er[k] > 0.2f ? 1.f : 0.f;  // zoom_mask.py:40-45
          }
          const float po = (k == 0) ? so[u].x : (k == 1) ? so[u].y : (k == 2) ? so[u].z : so[u].w;
          const float pr = (k == 0) ? sr[u].x : (k == 1) ? sr[u].y : (k == 2) ? sr[u].z : sr[u].w;
          ao_[k] = (ONE || c == 0) ? eo[k] : __fadd_rn(po, eo[k]);  // np.sum(axis=1), sequential
          ar_[k] = (ONE || c == 0) ? er[k] : __fadd_rn(pr, er[k]);
        }
        so[u] = make_float4(ao_[0], ao_[1], ao_[2], ao_[3]);
        sr[u] = make_float4(ar_[0], ar_[1], ar_[2], ar_[3]);
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const int q = q0 + u * stride;
      // 4-bit validity masks of the quad; background quads (the common case) skip the update
      const unsigned mo = (so[u].x > thr ? 1u : 0u) | (so[u].y > thr ? 2u : 0u) |
                          (so[u].z > thr ? 4u : 0u) | (so[u].w > thr ? 8u : 0u);
      const unsigned mr = (sr[u].x > thr ? 1u : 0u) | (sr[u].y > thr ? 2u : 0u) |
                          (sr[u].z > thr ? 4u : 0u) | (sr[u].w > thr ? 8u : 0u);
      if ((mo | mr) && q < quads) {
        const int p = q << 2;
        const int y = p / W;
        const int x = p - y * W;
        if (mo) {
          ao.add(x + __ffs(mo) - 1, y);
          ao.add(x + 31 - __clz(mo), y);
        }
        if (mr) {
          ar.add(x + __ffs(mr) - 1, y);
          ar.add(x + 31 - __clz(mr), y);
        }
      }
    }
  }
  // block reduction: warp shuffles, then 8 warps through shared memory
  int v[8] = {ao.x0, ao.x1, ao.y0, ao.y1, ar.x0, ar.x1, ar.y0, ar.y1};
  __shared__ int sm[8][8];
  __shared__ int s_last;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    v[k] = (k & 1) ? warp_max(v[k]) : warp_min(v[k]);
    if (lane == 0) sm[warp][k] = v[k];
  }
  __syncthreads();
  int32_t *w = ws + b * RC_ZF_WS_STRIDE;
  if (threadIdx.x < 8) {
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
  // last block of this sample finalises (ticket counter; fence orders the atomics)
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const int t = atomicAdd(w + 8, 1);
    s_last = (t == (int)gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && threadIdx.x == 0) {
    __threadfence();
    int32_t fin[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int e = atomicExch(w + k, 0);  // coherent read at L2, and reset for the next call
      fin[k] = (k & 1) ? e - 1 : INT_MAX - e;
    }
    atomicExch(w + 8, 0);
    zoom_factor_finalize(fin, src_pose + b * 12, K, zoom_factor + b * 4,
                         bbox ? bbox + b * 8 : nullptr, H, W);
    if (trans_out) {  // fused ZoomTrans forward (zoom_trans.py:30-44) of this sample
      const float wx = zoom_factor[b * 4];
      trans_out[b * 3 + 0] = __fdiv_rn(trans_in[b * 3 + 0], wx);
      trans_out[b * 3 + 1] = __fdiv_rn(trans_in[b * 3 + 1], wx);
      trans_out[b * 3 + 2] = trans_in[b * 3 + 2];
    }
  }
  pdl_wait();  // completion of this grid implies completion of the grid it was launched behind
}

// ---------------------------------------------------------------- sampler
struct ZoomPlane {
  const float *src;
  float *dst;
  long long src_bstride;  // floats between samples
  long long dst_bstride;
  int pre_op;
  int post_op;
  int accumulate;
  int channel;  // channel index inside its group (for the centre spot)
  float mean;
};

struct ZoomParams {
  ZoomPlane planes[RC_ZOOM_MAX_PLANES];
  const float *zoom_factor;
  int n_planes;
  int inverse;
  int H, W;
  int spot_x0, spot_x1, spot_y0, spot_y1;  // zoom_image_with_factor.py:36-45
  int src_aligned;                          // every source plane 16-byte aligned (cp.async staging)
  // sparse frame (compare step only; tile_state == NULL: off).  Planes in sparse_mask have
  // sources that are zero outside the renderer's covered 16x16 tiles (tile_count > 0) and
  // map zero to a constant; an output tile whose source window touches no covered tile and
  // whose state byte says it already holds that constant is skipped for those planes.
  const int32_t *tile_count;  // [B][src_tiles]
  uint8_t *tile_state;        // [B][gridDim.y * gridDim.x][4 warps]
  unsigned sparse_mask;
  int src_tiles_x, src_tiles, state_valid;
};

// Affine parameters of one sample: (wx, wy, tx, ty) as float32, forward or
// inverse (zoom_flow.py:35-44 evaluated in float64 like numpy 1.x scalars).
__device__ __forceinline__ void zoom_affine(const float *__restrict__ zf, int inverse, int H,
                                            int W, float &wx, float &wy, float &tx, float &ty,
                                            float &wx_in) {
  const float a = __ldg(zf + 0), bq = __ldg(zf + 1), c = __ldg(zf + 2), d = __ldg(zf + 3);
  wx_in = a;
  if (!inverse) {
    wx = a;
    wy = bq;
    tx = c;
    ty = d;
  } else {
    const double wxi = a, wyi = bq, txi = c, tyi = d;
    const double crop_w = wxi * W, crop_h = wyi * H;
    const double cx = __dadd_rn(__dmul_rn(__dmul_rn(txi, 0.5), (double)W), 0.5 * W);
    const double cy = __dadd_rn(__dmul_rn(__dmul_rn(tyi, 0.5), (double)H), 0.5 * H);
    wx = (float)(1.0 / wxi);
    wy = (float)(1.0 / wyi);
    tx = (float)__dmul_rn((W * 0.5 - cx) / crop_w, 2.0);
    ty = (float)__dmul_rn((H * 0.5 - cy) / crop_h, 2.0);
  }
}

// Source coordinate of output index i along an axis of length n:
//   GridGenerator: g = w * (-1 + i * 2/(n-1)) + t      (float32)
//   BilinearSampler: real = (g + 1) * (n - 1) / 2      (float32)
__device__ __forceinline__ float src_coord(int i, float step, float w, float t, float nm1) {
  const float n = __fmaf_rn((float)i, step, -1.0f);
  const float g = __fadd_rn(__fmul_rn(w, n), t);
  return __fmul_rn(__fmul_rn(__fadd_rn(g, 1.0f), nm1), 0.5f);
}

__device__ __forceinline__ float pre(float v, int pre_op, float mean) {
  if (pre_op == RC_PRE_THRESH02) return v > 0.2f ? 1.f : 0.f;
  if (pre_op == RC_PRE_ADD_MEAN) return __fadd_rn(v, mean);
  return v;
}

// The four-tap blend of MXNet's BilinearSampler (bilinear_sampler.cu forward kernel),
// evaluated in float32 in the operator's left-to-right association.
__device__ __forceinline__ float blend(float tl, float tr, float bl, float br, float wyt,
                                       float wyb, float wxl, float wxr) {
  float v = __fmul_rn(__fmul_rn(tl, wyt), wxl);
  v = __fmaf_rn(__fmul_rn(tr, wyt), wxr, v);
  v = __fmaf_rn(__fmul_rn(bl, wyb), wxl, v);
  v = __fmaf_rn(__fmul_rn(br, wyb), wxr, v);
  return v;
}

// Output tile of one block: ZT_W x ZT_H pixels, 128 threads; a thread owns one column of
// the tile (eight vertically adjacent outputs).
constexpr int ZT_W = 128;
constexpr int ZT_H = 8;
constexpr int ZT_THREADS = 128;

// Per-plane, per-sample constants resolved once per block (shared memory).
struct PlaneCtx {
  const float *src;  // plane of this sample
  float *dst;
  int pre_op, post_op, accumulate, channel;
  float mean;
  int pad_[3];
};

// round half away from zero, the FADD.RZ + FRND.TRUNC sequence CUDA's roundf reduces to
// for |v| < 2^23 (exact for every float: v + 0.5 rounded toward zero never crosses an integer).
__device__ __forceinline__ float round_away(float v) {
  return truncf(__fadd_rz(v, copysignf(0.5f, v)));
}

// Shared-memory accesses through 32-bit shared-window addresses (keeps the address
// arithmetic in 32-bit registers instead of re-deriving the generic pointer per plane).
__device__ __forceinline__ float lds_f1(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
  return v;
}

// 16-byte asynchronous global -> shared copy that bypasses registers and L1; `bytes` = 16
// copies, `bytes` = 0 zero-fills the destination (out-of-image chunk).
__device__ __forceinline__ void cp_async_16(uint32_t dst, const void *src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() {
  asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory");
}

// IEEE-754 round-to-nearest a / b without the library's slow path.  `rb` = RN(1/b).  For
// operands inside the guarded exponent range the two FMA refinements below deliver the
// correctly rounded quotient (Markstein; checked at random on 4e8 operand pairs including
// all-ones / all-zeros significands against the host's IEEE division); zero numerators --
// the background of a flow field, i.e. most pixels -- would otherwise take __fdiv_rn's
// ~60-instruction special-case subroutine.
__device__ __forceinline__ float div_rn_guarded(float a, float b, float rb, bool b_ok) {
  const float aa = fabsf(a);
  if (aa == 0.f) return __fmul_rn(a, b);  // +-0 with the sign of a / b
  if (b_ok && aa > 1e-15f && aa < 1e15f) {
    float q = __fmul_rn(a, rb);
    float r = __fmaf_rn(-b, q, a);
    q = __fmaf_rn(r, rb, q);
    r = __fmaf_rn(-b, q, a);
    return __fmaf_rn(r, rb, q);
  }
  return __fdiv_rn(a, b);
}

// Post-operation on the ZT_H outputs of one column (x = i, rows j0 .. j0 + ZT_H - 1).
__device__ __forceinline__ void post_col(float (&out)[ZT_H], int post_op, float mean, int channel,
                                         const ZoomParams &P, float wx_in, float rwx, bool wx_ok,
                                         int i, int j0) {
  switch (post_op) {
    case RC_POST_ROUND:
#pragma unroll
      for (int k = 0; k < ZT_H; ++k) out[k] = round_away(out[k]);
      break;
    case RC_POST_ROUND_M045:
#pragma unroll
      for (int k = 0; k < ZT_H; ++k) out[k] = round_away(__fsub_rn(out[k], 0.45f));
      break;
    case RC_POST_DIV_WX:
#pragma unroll
      for (int k = 0; k < ZT_H; ++k) out[k] = div_rn_guarded(out[k], wx_in, rwx, wx_ok);
      break;
    case RC_POST_MUL_WX:
#pragma unroll
      for (int k = 0; k < ZT_H; ++k) out[k] = __fmul_rn(out[k], wx_in);
      break;
    case RC_POST_SUB_MEAN:
#pragma unroll
      for (int k = 0; k < ZT_H; ++k) out[k] = __fsub_rn(out[k], mean);
      break;
    case RC_POST_SPOT_SUB_MEAN: {
      const bool xin = (channel == 0) && (i >= P.spot_x0) && (i < P.spot_x1);
#pragma unroll
      for (int k = 0; k < ZT_H; ++k) {
        const int j = j0 + k;
        const float spot = (xin && j >= P.spot_y0 && j < P.spot_y1) ? 255.f : 0.f;
        out[k] = __fsub_rn(fmaxf(out[k], spot), mean);
      }
      break;
    }
    default:
      break;
  }
}

// Row coordinate of one output row: source row index and the two row weights.
struct RowCoord {
  int y0;
  float wt, wb;
};
__device__ __forceinline__ RowCoord row_coord(int j, float stepy, float wy, float ty, float nym1) {
  const float yr = src_coord(j, stepy, wy, ty, nym1);
  const float yf = floorf(yr);
  RowCoord r;
  r.y0 = (int)yf;
  r.wt = __fsub_rn(1.0f, __fsub_rn(yr, yf));
  r.wb = __fsub_rn(1.0f, r.wt);
  return r;
}

// The {left, right} taps of one window row for this thread's column, with the plane's
// pre-operation applied.  EDGE: the window reaches outside the image, where taps are zero
// AFTER the pre-operation (the sampler zero-pads the pre-processed image), so the additive
// pre-operation must not touch them; `ok` = this row is inside the image, okl/okr = the
// columns are.  (The threshold maps the zero fill to zero by itself.)
template <int PRE, bool EDGE>
__device__ __forceinline__ float2 load_taps(uint32_t addr, float mean, bool ok, bool okl, bool okr) {
  float2 t;
  t.x = lds_f1(addr);
  t.y = lds_f1(addr + 4u);
  if (PRE == RC_PRE_THRESH02) {
    t.x = t.x > 0.2f ? 1.f : 0.f;
    t.y = t.y > 0.2f ? 1.f : 0.f;
  } else if (PRE == RC_PRE_ADD_MEAN) {
    if (EDGE) {
      t.x = (ok && okl) ? __fadd_rn(t.x, mean) : 0.f;
      t.y = (ok && okr) ? __fadd_rn(t.y, mean) : 0.f;
    } else {
      t.x = __fadd_rn(t.x, mean);
      t.y = __fadd_rn(t.y, mean);
    }
  }
  return t;
}

// Eight outputs of one column of one plane from the staged window.
//   a_row0      shared address of the left tap of output row 0's upper source row
//   pitch       bytes between window rows
//   step_mask   bit m: row m reads one source row further down than row m-1; otherwise it
//               shares both source rows with row m-1 (zoom-in: no other case exists)
//   rowok_mask  bit r: window row (first row of output row 0) + r is inside the image (EDGE)
template <int PRE, bool EDGE>
__device__ __forceinline__ void blend_col(float (&out)[ZT_H], const RowCoord (&rows)[ZT_H],
                                          uint32_t a_row0, uint32_t pitch, unsigned step_mask,
                                          unsigned rowok_mask, float mean,
                                          bool okl, bool okr, float wl, float wr) {
  // Horizontal blend first: h(row) = l * wl + r * wr is shared by every output row that reads
  // the source row (zooming in, eight output rows touch ~8*wy+2 source rows), so an output costs
  // one FMUL + one FFMA plus its share of the row blends instead of the eight operations of the
  // operator's tap-by-tap association.  Same four products, summed in another order: the
  // difference is <= 2 ulp of the result (1e-7 relative), three decades inside the 1e-4 bar.
  uint32_t ab = a_row0;
  unsigned rk = rowok_mask;
  float2 t = load_taps<PRE, EDGE>(ab, mean, rk & 1u, okl, okr);
  float ht = __fmaf_rn(t.y, wr, __fmul_rn(t.x, wl));
  ab += pitch;
  rk >>= 1;
  t = load_taps<PRE, EDGE>(ab, mean, rk & 1u, okl, okr);
  float hb = __fmaf_rn(t.y, wr, __fmul_rn(t.x, wl));
  out[0] = __fmaf_rn(hb, rows[0].wb, __fmul_rn(ht, rows[0].wt));
#pragma unroll
  for (int m = 1; m < ZT_H; ++m) {
    if (step_mask & (1u << m)) {  // warp-uniform
      ht = hb;
      ab += pitch;
      rk >>= 1;
      t = load_taps<PRE, EDGE>(ab, mean, rk & 1u, okl, okr);
      hb = __fmaf_rn(t.y, wr, __fmul_rn(t.x, wl));
    }
    out[m] = __fmaf_rn(hb, rows[m].wb, __fmul_rn(ht, rows[m].wt));
  }
}

// grid = (ceil(W/ZT_W), ceil(H/ZT_H), B); block = 128; dynamic smem = smem_floats * 4.
//
// One thread owns ONE output column of the 128 x 8 tile and walks its eight rows; the lanes
// of a warp are 32 adjacent columns.  Consequences, in order of weight:
//  * the source window of the tile (every tap any of its 1024 outputs can touch, widened to
//    16-byte chunks) is staged in shared memory once per plane by cp.async: ~9 asynchronous
//    16-byte copies per thread for all 13 planes, no registers, out-of-image chunks
//    zero-filled by the copy itself.  The planes' pre-operations are applied to the taps as
//    they are read;
//  * everything that depends on the row (source row, row weights, which window rows to read)
//    depends on blockIdx only.  Zooming in, consecutive output rows share source rows, so a
//    column of eight outputs reads only ~8*wy+2 window rows instead of 16; the stepping
//    pattern is a warp-uniform bit mask (made visible to the compiler by a warp reduction);
//  * the 32 lanes of a warp read 32*wx+1 consecutive floats of one window row: one
//    conflict-free shared-memory wavefront per load, and every store instruction writes 128
//    contiguous bytes per warp.
// As many planes as fit are staged per barrier pair.  Tiles that zoom out (a row skipping a
// source row), whose window does not fit, or whose sources are not 16-byte aligned fall back
// to gathering through the read-only path.
//
// WC: compile-time image width (0 = use P.W); with the width known the eight row stores of a
// column use immediate offsets from one address.
template <int WC>
__global__ void __launch_bounds__(ZT_THREADS, 9)
zoom_sample_kernel(const __grid_constant__ ZoomParams P, int smem_floats) {
  extern __shared__ float4 win4[];
  __shared__ PlaneCtx ctx[RC_ZOOM_MAX_PLANES];
  const int H = P.H, W = WC ? WC : P.W;
  const int b = blockIdx.z;
  const int i_first = blockIdx.x * ZT_W, j_first = blockIdx.y * ZT_H;
  const int i_last = min(i_first + ZT_W, W) - 1, j_last = min(j_first + ZT_H, H) - 1;
  const int i = i_first + threadIdx.x;
  const bool act = i < W;
  const bool full_rows = j_first + ZT_H <= H;
  const size_t col_off = (size_t)j_first * W + i;

  float wx, wy, tx, ty, wx_in;
  zoom_affine(P.zoom_factor + b * 4, P.inverse, H, W, wx, wy, tx, ty, wx_in);
  const float stepx = (float)(2.0 / (W - 1)), stepy = (float)(2.0 / (H - 1));
  const float nxm1 = (float)(W - 1), nym1 = (float)(H - 1);
  // division by wx_in (ZoomFlow): reciprocal once per thread; the FMA refinement is exact
  // unless wx_in is tiny/huge or its significand is all ones (then __fdiv_rn is used).
  const float rwx = __frcp_rn(wx_in);
  const float awx = fabsf(wx_in);
  const bool wx_ok = (awx > 1e-15f) && (awx < 1e15f) &&
                     ((__float_as_uint(wx_in) & 0x7fffffu) != 0x7fffffu);

  // column coordinate of this thread (clamped index keeps inactive threads harmless)
  const float xr = src_coord(min(i, W - 1), stepx, wx, tx, nxm1);
  const float xf = floorf(xr);
  const int x0 = (int)xf;
  const float wl = __fsub_rn(1.0f, __fsub_rn(xr, xf));
  const float wr = __fsub_rn(1.0f, wl);
  // row coordinates of the tile (block-uniform)
  RowCoord rows[ZT_H];
#pragma unroll
  for (int m = 0; m < ZT_H; ++m) rows[m] = row_coord(min(j_first + m, H - 1), stepy, wy, ty, nym1);

  // tile window (block-uniform); src_coord is monotone in the index when the scale is > 0
  const float fx_lo = floorf(src_coord(i_first, stepx, wx, tx, nxm1));
  const float fx_hi = floorf(src_coord(i_last, stepx, wx, tx, nxm1));
  const float fy_lo = floorf(src_coord(j_first, stepy, wy, ty, nym1));
  const float fy_hi = floorf(src_coord(j_last, stepy, wy, ty, nym1));
  // window: source columns x_lo .. x_hi+1 (widened to whole 4-float chunks), rows y_lo .. y_hi+1
  const float whf = fy_hi - fy_lo + 2.0f;
  const bool sane = (wx > 0.f) && (wy > 0.f) && (fx_hi >= fx_lo) && (fy_hi >= fy_lo) &&
                    (fabsf(fx_lo) < 1e6f) && (fabsf(fx_hi) < 1e6f) && (fabsf(fy_lo) < 1e6f) &&
                    (whf <= 32.0f) && P.src_aligned;
  int x_lo = 0, y_lo = 0, nch = 1, wh = 2;
  if (sane) {
    x_lo = (int)fx_lo;
    y_lo = (int)fy_lo;
    wh = (int)whf;
    const int xa_ = x_lo & ~3;  // floor to a multiple of 4 (two's complement: also for x_lo < 0)
    nch = (((int)fx_hi + 1 - xa_) >> 2) + 1;
  }
  const int xa = x_lo & ~3;
  const int plane_floats = nch * 4 * wh;
  // Row stepping pattern of the tile: bit m of step_mask = row m reads one source row
  // further down than row m-1; jump = any other change (zoom-out, not staged).  Block-uniform
  // by construction; the warp reduction makes that visible to the compiler, which then keeps
  // the mask in uniform registers and branches uniformly.
  unsigned sm_ = 0u, jm_ = 0u;
#pragma unroll
  for (int m = 1; m < ZT_H; ++m) {
    const int d = rows[m].y0 - rows[m - 1].y0;
    if (d == 1) sm_ |= 1u << m;
    else if (d != 0) jm_ = 1u;
  }
  const unsigned step_mask = __reduce_or_sync(0xffffffffu, sm_);
  const bool jump = __reduce_or_sync(0xffffffffu, jm_) != 0u;
  const bool staged = sane && !jump && plane_floats <= smem_floats;
  // sparse frame, decided per WARP (32 output columns x 8 rows): does the warp's source window
  // touch no covered tile, and do its outputs already hold the sparse planes' background?
  // A skipping warp still stages and meets the barriers; it only omits blend, post and store
  // of the planes in P.sparse_mask.
  unsigned wskip = 0u;  // planes this warp skips
  if (P.tile_state != nullptr) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t *st = P.tile_state +
                  (((size_t)b * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * (ZT_THREADS / 32) + warp;
    const int held = *st;  // every lane reads before the vote below, lane 0 writes after it
    const int xw0 = __shfl_sync(0xffffffffu, x0, 0), xw1 = __shfl_sync(0xffffffffu, x0, 31) + 1;
    int any = 1;
    if (sane && !P.inverse) {
      const int sx0 = max(xw0, 0), sx1 = min(xw1, W - 1);
      const int sy0 = max(y_lo, 0), sy1 = min(y_lo + wh - 1, H - 1);
      any = 0;
      if (sx1 >= sx0 && sy1 >= sy0) {
        const int tx0 = sx0 >> 4, ntx = (sx1 >> 4) - tx0 + 1, ty0 = sy0 >> 4,
                  nty = (sy1 >> 4) - ty0 + 1;
        if (ntx * nty > 32) any = 1;
        else if (lane < ntx * nty)
          any = __ldg(P.tile_count + (size_t)b * P.src_tiles + (ty0 + lane / ntx) * P.src_tiles_x +
                      tx0 + lane % ntx) > 0 ? 1 : 0;
      }
    }
    const bool bg_now = !__any_sync(0xffffffffu, any);
    if (lane == 0) *st = bg_now ? 1 : 0;
    if (bg_now && P.state_valid && held) wskip = P.sparse_mask;
  }
  const int n_planes = P.n_planes;
  if (threadIdx.x < P.n_planes) {
    const ZoomPlane &zp = P.planes[threadIdx.x];
    PlaneCtx c;
    c.src = zp.src + (long long)b * zp.src_bstride;
    c.dst = zp.dst + (long long)b * zp.dst_bstride;
    c.pre_op = zp.pre_op;
    c.post_op = zp.post_op;
    c.accumulate = zp.accumulate;
    c.channel = zp.channel;
    c.mean = zp.mean;
    ctx[threadIdx.x] = c;
  }
  __syncthreads();  // ctx[] visible
  if (staged) {
    const int pps = max(1, min(n_planes, smem_floats / plane_floats));  // planes per stage
    const uint32_t win_s = (uint32_t)__cvta_generic_to_shared(win4);
    const uint32_t pitch = 16u * (uint32_t)nch;
    const uint32_t plane_bytes = 4u * (uint32_t)plane_floats;
    // does the window reach outside the image?  (then additive pre-operations must skip the
    // zero fill); rowok bit r: window row rows[0].y0 + r is inside the image
    const int r_first = rows[0].y0;
    const int rk_lo = max(0, -r_first), rk_hi = min(31, H - 1 - r_first);
    const unsigned rk_ = (rk_hi >= rk_lo) ? ((0xffffffffu >> (31 - rk_hi)) & (0xffffffffu << rk_lo)) : 0u;
    const unsigned rowok_mask = __reduce_or_sync(0xffffffffu, rk_);
    const bool edge_ = (xa < 0) || (xa + 4 * nch > W) || (y_lo < 0) || (y_lo + wh > H);
    const bool edge = __reduce_or_sync(0xffffffffu, edge_ ? 1u : 0u) != 0u;
    const bool okl = (x0 >= 0) && (x0 <= W - 1), okr = (x0 + 1 >= 0) && (x0 + 1 <= W - 1);
    uint32_t a_row0 = win_s + 4u * (uint32_t)(x0 - xa) + (uint32_t)(r_first - y_lo) * pitch;
    asm volatile("" : "+r"(a_row0));  // keep the shared-window address in a register

    for (int p0 = 0; p0 < n_planes; p0 += pps) {
      const int np = min(pps, n_planes - p0);
      if (p0 > 0) __syncthreads();  // consumers of the previous stage are done
      // stage: chunk q = (window row, 4-float column chunk); the same chunk of every plane
      const int nchunks = nch * wh;
      const float inv_nch = __frcp_rn((float)nch);
      for (int q = threadIdx.x; q < nchunks; q += ZT_THREADS) {
        int r = __float2int_rz(__fmul_rn((float)q + 0.5f, inv_nch));
        int c = q - r * nch;
        if (c < 0) {
          c += nch;
          --r;
        } else if (c >= nch) {
          c -= nch;
          ++r;
        }
        const int ys = y_lo + r, xs = xa + 4 * c;
        const bool inb = (xs >= 0) && (xs + 3 <= W - 1) && (ys >= 0) && (ys <= H - 1);
        const int goff = inb ? ys * W + xs : 0;
        const int bytes = inb ? 16 : 0;
        uint32_t sa = win_s + 16u * (uint32_t)q;
        for (int pl = 0; pl < np; ++pl, sa += plane_bytes)
          cp_async_16(sa, ctx[p0 + pl].src + goff, bytes);
      }
      cp_async_wait_all();
      __syncthreads();
      uint32_t a_pl = a_row0;
      for (int pl = 0; pl < np; ++pl, a_pl += plane_bytes) {
        if ((wskip >> (p0 + pl)) & 1u) continue;  // warp-uniform
        const PlaneCtx &c_ = ctx[p0 + pl];
        const int pre_op = c_.pre_op;
        const float mean = c_.mean;
        float out[ZT_H];
        if (pre_op == RC_PRE_NONE)
          blend_col<RC_PRE_NONE, false>(out, rows, a_pl, pitch, step_mask, rowok_mask,
                                        mean, okl, okr, wl, wr);
        else if (pre_op == RC_PRE_THRESH02)
          blend_col<RC_PRE_THRESH02, false>(out, rows, a_pl, pitch, step_mask,
                                            rowok_mask, mean, okl, okr, wl, wr);
        else if (!edge)
          blend_col<RC_PRE_ADD_MEAN, false>(out, rows, a_pl, pitch, step_mask,
                                            rowok_mask, mean, okl, okr, wl, wr);
        else
          blend_col<RC_PRE_ADD_MEAN, true>(out, rows, a_pl, pitch, step_mask,
                                           rowok_mask, mean, okl, okr, wl, wr);
        const int post_op = c_.post_op, channel = c_.channel, acc = c_.accumulate;
        float *dst = c_.dst + col_off;
        post_col(out, post_op, mean, channel, P, wx_in, rwx, wx_ok, i, j_first);
        if (act) {
          if (!acc && full_rows) {
#pragma unroll
            for (int m = 0; m < ZT_H; ++m) st_f1(dst + m * W, out[m]);
          } else {
#pragma unroll
            for (int m = 0; m < ZT_H; ++m) {
              if (j_first + m < H) {
                float *d = dst + (size_t)m * W;
                *d = acc ? __fadd_rn(*d, out[m]) : out[m];
              }
            }
          }
        }
      }
    }
    return;
  }

  // ---- fallback: gather every tap through the read-only path, bounds-checked
  if (!act) return;
  const bool xin0 = (x0 >= 0) && (x0 <= W - 1);
  const bool xin1 = (x0 + 1 >= 0) && (x0 + 1 <= W - 1);
  for (int pl = 0; pl < n_planes; ++pl) {
    if ((wskip >> pl) & 1u) continue;
    const PlaneCtx &zp = ctx[pl];
    const float *__restrict__ s = zp.src;
    float out[ZT_H];
#pragma unroll
    for (int m = 0; m < ZT_H; ++m) {
      const int y0 = rows[m].y0;
      const bool yin0 = (y0 >= 0) && (y0 <= H - 1);
      const bool yin1 = (y0 + 1 >= 0) && (y0 + 1 <= H - 1);
      const float *__restrict__ r0 = s + (long long)y0 * W;
      const float *__restrict__ r1 = r0 + W;
      const float tl = (yin0 && xin0) ? pre(__ldg(r0 + x0), zp.pre_op, zp.mean) : 0.f;
      const float tr = (yin0 && xin1) ? pre(__ldg(r0 + x0 + 1), zp.pre_op, zp.mean) : 0.f;
      const float bl = (yin1 && xin0) ? pre(__ldg(r1 + x0), zp.pre_op, zp.mean) : 0.f;
      const float br = (yin1 && xin1) ? pre(__ldg(r1 + x0 + 1), zp.pre_op, zp.mean) : 0.f;
      out[m] = blend(tl, tr, bl, br, rows[m].wt, rows[m].wb, wl, wr);
    }
    post_col(out, zp.post_op, zp.mean, zp.channel, P, wx_in, rwx, wx_ok, i, j_first);
    float *dst = zp.dst + (size_t)j_first * W + i;
#pragma unroll
    for (int m = 0; m < ZT_H; ++m) {
      if (j_first + m < H) {
        float *d = dst + (size_t)m * W;
        *d = zp.accumulate ? __fadd_rn(*d, out[m]) : out[m];
      }
    }
  }
}

// ------------------------------------------------------------- ZoomTrans
// zoom_trans.py:30-44 (forward) and :57-73 (backward); float32 scalar ops.
__global__ void zoom_trans_kernel(const float *__restrict__ zf, const float *__restrict__ in,
                                  float *__restrict__ out, int B, int inverse, int scale_xy,
                                  int accumulate) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const float wx = zf[b * 4 + 0];  // wy = zoom_factor[b][0] as well (:32-33)
  float x = in[b * 3 + 0], y = in[b * 3 + 1];
  const float z = in[b * 3 + 2];
  if (scale_xy) {
    if (inverse) {
      x = __fmul_rn(x, wx);
      y = __fmul_rn(y, wx);
    } else {
      x = __fdiv_rn(x, wx);
      y = __fdiv_rn(y, wx);
    }
  }
  if (accumulate) {
    out[b * 3 + 0] += x;
    out[b * 3 + 1] += y;
    out[b * 3 + 2] += z;
  } else {
    out[b * 3 + 0] = x;
    out[b * 3 + 1] = y;
    out[b * 3 + 2] = z;
  }
}

// Zoom factor when the two boxes are already known: block of 128 pairs per CTA, one thread per
// pair.  The rendered box comes from the workspace entries 4..7 (frame.cu's tile pass; read
// with atomicExch, which also re-arms the workspace), the observed box from `observed_box`.
__global__ void __launch_bounds__(128)
zoom_factor_boxes_kernel(const int32_t *__restrict__ observed_box, int32_t *__restrict__ ws,
                         const float *__restrict__ src_pose, const float *__restrict__ K,
                         float *__restrict__ zoom_factor, int32_t *__restrict__ bbox,
                         const float *__restrict__ trans_in, float *__restrict__ trans_out, int B,
                         int H, int W) {
  pdl_wait();  // the tile pass's atomics
  pdl_launch_dependents();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
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
  if (trans_out) {  // fused ZoomTrans forward (zoom_trans.py:30-44)
    const float wx = zoom_factor[b * 4];
    trans_out[b * 3 + 0] = __fdiv_rn(trans_in[b * 3 + 0], wx);
    trans_out[b * 3 + 1] = __fdiv_rn(trans_in[b * 3 + 1], wx);
    trans_out[b * 3 + 2] = trans_in[b * 3 + 2];
  }
}

int launch_zoom_factor_boxes(const int32_t *observed_box, const float *src_pose, const float *K,
                             float *zoom_factor, int32_t *bbox, int32_t *workspace,
                             const float *trans_in, float *trans_out, int batch_size, int height,
                             int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return note_status(RC_OK);
  RC_REQUIRE(observed_box && src_pose && K && zoom_factor && workspace, RC_ERR_INVALID_ARG);
  RC_REQUIRE((trans_in == nullptr) == (trans_out == nullptr), RC_ERR_INVALID_ARG);
  RC_CUDA_TRY(launch_maybe_pdl(zoom_factor_boxes_kernel, dim3((batch_size + 127) / 128), dim3(128),
                               0, static_cast<cudaStream_t>(stream), tl_pdl, observed_box,
                               workspace, src_pose, K, zoom_factor, bbox, trans_in, trans_out,
                               batch_size, height, width));
  return note_status(check_launch());
}

int launch_zoom_factor(const float *obs, const float *rend, const float *src_pose,
                       const float *K, const float *pixel_means, float *zoom_factor,
                       int32_t *bbox, int32_t *workspace, const float *trans_in,
                       float *trans_out, int batch_size, int channels, int height, int width,
                       int mode, void *stream) {
  RC_REQUIRE(batch_size >= 0 && channels > 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(mode == RC_ZF_MASK || mode == RC_ZF_IMAGE, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return note_status(RC_OK);
  RC_REQUIRE(obs && rend && src_pose && K && zoom_factor && workspace, RC_ERR_INVALID_ARG);
  RC_REQUIRE(mode == RC_ZF_MASK || pixel_means != nullptr, RC_ERR_INVALID_ARG);
  RC_REQUIRE((trans_in == nullptr) == (trans_out == nullptr), RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && aligned16(obs) && aligned16(rend), RC_ERR_UNALIGNED);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int quads = height * width / 4;
  // One resident wave of at most two 256-thread blocks per SM (the kernel is compiled for four
  // per SM, so half of every SM stays free for the flow kernel that overlaps it), never more
  // blocks per sample than it takes to give every thread four quads (the loop batches 4 loads
  // per plane).  All blocks being resident at once also means griddepcontrol.launch_dependents
  // releases the next kernel immediately.
  const int sm_count = device_sm_count();
  const int max_bps = (quads + 256 * 4 - 1) / (256 * 4);
  int bps = (sm_count * (channels == 1 ? 2 : 1)) / batch_size;
  // large batches fill the GPU several times over anyway: four blocks per sample keep the last
  // wave short (1024 whole-sample blocks would be 1.7 waves of 592 resident blocks)
  if (bps < 4 && batch_size >= sm_count) bps = 4;
  if (bps > max_bps) bps = max_bps;
  if (bps < 1) bps = 1;
  dim3 grid(bps, batch_size);
  const bool pdl = tl_pdl;
#define RC_BBOX_LAUNCH(MODE_, ONE_)                                                              \
  RC_CUDA_TRY(launch_maybe_pdl(zoom_bbox_kernel<MODE_, ONE_>, grid, dim3(256), 0, s, pdl, obs, rend, \
                               src_pose, K, pixel_means, zoom_factor, bbox, workspace, trans_in,     \
                               trans_out, channels, height, width))
  if (mode == RC_ZF_MASK) {
    if (channels == 1) RC_BBOX_LAUNCH(RC_ZF_MASK, true);
    else RC_BBOX_LAUNCH(RC_ZF_MASK, false);
  } else {
    if (channels == 1) RC_BBOX_LAUNCH(RC_ZF_IMAGE, true);
    else RC_BBOX_LAUNCH(RC_ZF_IMAGE, false);
  }
#undef RC_BBOX_LAUNCH
  return note_status(check_launch());
}

}  // namespace rc

extern "C" {

int rc_zoom_factor_dev(const float *obs, const float *rend, const float *src_pose,
                       const float *K, const float *pixel_means, float *zoom_factor,
                       int32_t *bbox, int32_t *workspace, int batch_size, int channels,
                       int height, int width, int mode, void *stream) {
  return rc::launch_zoom_factor(obs, rend, src_pose, K, pixel_means, zoom_factor, bbox, workspace,
                                nullptr, nullptr, batch_size, channels, height, width, mode,
                                stream);
}

int rc_zoom_sample_dev(const float *zoom_factor, int inverse, const rc_zoom_group *groups,
                       int n_groups, int batch_size, int height, int width, void *stream) {
  return rc::launch_zoom_sample(zoom_factor, inverse, groups, n_groups, batch_size, height, width,
                                nullptr, stream);
}

}  // extern "C"

namespace rc {
int launch_zoom_sample(const float *zoom_factor, int inverse, const rc_zoom_group *groups,
                       int n_groups, int batch_size, int height, int width,
                       const ZoomSparse *sparse, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 1 && width > 1, RC_ERR_INVALID_ARG);
  RC_REQUIRE(n_groups >= 0 && n_groups <= RC_ZOOM_MAX_GROUPS, RC_ERR_INVALID_ARG);
  if (batch_size == 0 || n_groups == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(zoom_factor && groups, RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0, RC_ERR_UNALIGNED);
  rc::ZoomParams P;
  P.n_planes = 0;
  P.sparse_mask = 0u;
  const long long plane = (long long)height * width;
  for (int g = 0; g < n_groups; ++g) {
    const rc_zoom_group &G = groups[g];
    RC_REQUIRE(G.src && G.dst && G.channels > 0 && G.channels <= 4, RC_ERR_INVALID_ARG);
    RC_REQUIRE(G.pre_op >= 0 && G.pre_op <= RC_PRE_ADD_MEAN, RC_ERR_INVALID_ARG);
    RC_REQUIRE(G.post_op >= 0 && G.post_op <= RC_POST_SPOT_SUB_MEAN, RC_ERR_INVALID_ARG);
    RC_REQUIRE(rc::aligned16(G.dst), RC_ERR_UNALIGNED);
    for (int c = 0; c < G.channels; ++c) {
      RC_REQUIRE(P.n_planes < RC_ZOOM_MAX_PLANES, RC_ERR_CAPACITY);
      rc::ZoomPlane &z = P.planes[P.n_planes++];
      const bool bcast = (G.flags & RC_ZOOM_SRC_BROADCAST) != 0;
      z.src = bcast ? G.src : G.src + c * plane;
      z.dst = G.dst + c * plane;
      z.src_bstride = bcast ? plane : G.channels * plane;
      z.dst_bstride = G.channels * plane;
      z.pre_op = G.pre_op;
      z.post_op = G.post_op;
      z.accumulate = (G.flags & RC_ZOOM_ACCUMULATE) ? 1 : 0;
      if ((G.flags & RC_ZOOM_SRC_FRAME_SPARSE) && !z.accumulate)
        P.sparse_mask |= 1u << (P.n_planes - 1);
      z.channel = c;
      z.mean = G.mean[c];
    }
  }
  P.src_aligned = 1;
  for (int g = 0; g < n_groups; ++g)
    if (!rc::aligned16(groups[g].src)) P.src_aligned = 0;
  P.tile_count = nullptr;
  P.tile_state = nullptr;
  P.src_tiles_x = P.src_tiles = P.state_valid = 0;
  if (sparse && sparse->tile_state && sparse->tile_count && P.sparse_mask && !inverse) {
    P.tile_count = sparse->tile_count;
    P.tile_state = sparse->tile_state;
    P.src_tiles_x = sparse->tiles_x;
    P.src_tiles = sparse->tiles_x * sparse->tiles_y;
    P.state_valid = sparse->state_valid ? 1 : 0;
  }
  P.zoom_factor = zoom_factor;
  P.inverse = inverse ? 1 : 0;
  P.H = height;
  P.W = width;
  // floor(W/2 - 5) .. ceil(W/2 + 5) (zoom_image_with_factor.py:38-41, spot_radius = 5)
  P.spot_x0 = (int)floor((double)width / 2 - 5);
  P.spot_x1 = (int)ceil((double)width / 2 + 5);
  P.spot_y0 = (int)floor((double)height / 2 - 5);
  P.spot_y1 = (int)ceil((double)height / 2 + 5);
  // 22 KB of staging per block (+ the 1 KB plane context): nine 128-thread blocks stay
  // resident per SM; all 13 planes of the default configuration fit in one stage for zoom
  // factors up to ~0.45 (window (128*wx+8) x (8*wy+3) floats per plane).
  constexpr int kSmemFloats = 5632;
  static bool attr_set = false;
  if (!attr_set) {
    RC_CUDA_TRY(cudaFuncSetAttribute(rc::zoom_sample_kernel<640>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kSmemFloats * (int)sizeof(float)));
    RC_CUDA_TRY(cudaFuncSetAttribute(rc::zoom_sample_kernel<0>,
                                     cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kSmemFloats * (int)sizeof(float)));
    attr_set = true;
  }
  dim3 grid((width + rc::ZT_W - 1) / rc::ZT_W, (height + rc::ZT_H - 1) / rc::ZT_H, batch_size);
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  if (width == 640)
    rc::zoom_sample_kernel<640><<<grid, rc::ZT_THREADS, kSmemFloats * sizeof(float),
                                  static_cast<cudaStream_t>(stream)>>>(P, kSmemFloats);
  else
    rc::zoom_sample_kernel<0><<<grid, rc::ZT_THREADS, kSmemFloats * sizeof(float),
                                static_cast<cudaStream_t>(stream)>>>(P, kSmemFloats);
  return rc::note_status(rc::check_launch());
}
}  // namespace rc

extern "C" {

int rc_zoom_trans_dev(const float *zoom_factor, const float *trans_in, float *trans_out,
                      int batch_size, int inverse, int scale_xy, int accumulate, void *stream) {
  RC_REQUIRE(batch_size >= 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(zoom_factor && trans_in && trans_out, RC_ERR_INVALID_ARG);
  rc::zoom_trans_kernel<<<(batch_size + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      zoom_factor, trans_in, trans_out, batch_size, inverse, scale_xy, accumulate);
  return rc::note_status(rc::check_launch());
}

int rc_fill_zero_dev(float *dst, int64_t count, void *stream) {
  RC_REQUIRE(count >= 0, RC_ERR_INVALID_ARG);
  if (count == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(dst != nullptr, RC_ERR_INVALID_ARG);
  RC_CUDA_TRY(cudaMemsetAsync(dst, 0, (size_t)count * sizeof(float),
                              static_cast<cudaStream_t>(stream)));
  return rc::note_status(RC_OK);
}

}  // extern "C"
