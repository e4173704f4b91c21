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
#include <stdlib.h>

#include "common.cuh"
#include "zoom_finalize.cuh"

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

// (workspace layout and zoom_factor_finalize: zoom_finalize.cuh, shared with frame.cu)

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
  int dup_second;  // version 2: same source and operations as the previous plane (computed once)
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

// ===================================================================== resampler, version 2
//
// zoom_sample2_kernel: the same operator pair as version 1 above, reorganised after ncu showed
// version 1 bound by instruction issue (5.8e7 warp instructions for 6.4e7 outputs, SM 73 %,
// DRAM 50 %), not by HBM.  Measured on B200 (round 2, profiles/r2_zoom_*): default step
// (B = 16, zoom 0.32) 70.4 -> 65.9 us, cfg 5 (B = 1024, zoom ~0.17) 3450 -> 2720 us per launch =
// 98 % of the measured HBM peak on the bytes the launch has to move.
//
//  * ONE WARP owns a 32-column x 16-row output tile; the four warps of a block share the staged
//    source window (one barrier pair per stage) and nothing else.
//  * The blend is separable and is evaluated, plane by plane, in two phases that use two
//    DIFFERENT lane mappings, exchanged through a 2 KB per-warp shared-memory tile:
//      A (lane = output column): for every source row the tile touches (16 wy + 2 of them when
//        zooming in) the horizontal blend h = l * wl + r * wr of that row at this column.  The
//        taps come from the block's source window, staged in shared memory by 16-byte cp.async
//        exactly as in version 1 (as many planes per stage as fit; out-of-image chunks are
//        zero-filled by the copy itself), so a tap is an LDS with a 32-bit address.  The
//        plane's pre-operation is applied to the taps;
//      B (lane = 4 adjacent columns x one of 4 rows, four passes): out = ht * wt + hb * wb for
//        FOUR columns at once: two 128-bit shared loads, the sm_100 packed-fp32 pipe
//        (fma.rn.f32x2 = FFMA2: the same IEEE results as scalar operations, half the issue
//        slots), the plane's post-operation, one 128-bit store.  A warp-wide store instruction
//        writes four 128-byte row segments.
//  * Everything that depends on the tile only (row weights, exchange addresses, tap address) is
//    computed once per tile and reused by all 13 planes; the pre / post operations are template
//    instantiations selected once per plane (no per-value switches).
//  * Same sampling coordinates, weights, tap selection and blend association as version 1
//    ((l*wl + r*wr) per source row, then ht*wt + hb*wb), i.e. bit-identical outputs for float
//    planes; rounded 0/1 planes after RC_PRE_THRESH02 use the comparison v >= 0.5 (identical to
//    roundf for v in [0, 1.5)).
//  * Tiles that are not zoom-in shaped (the 16 rows touch more than 16 source rows: scale > ~0.9,
//    inverse zooms), unaligned sources and degenerate factors: every thread gathers its own
//    outputs (z2_gather_tile).
//
// What the measurements taught (each a separate ncu capture):
//   - taps straight from global memory through the read-only path instead of the staged window:
//     64-bit address arithmetic per load tripled phase A's instruction count and the loads'
//     latency was exposed (107 us);
//   - runs of three planes per phase with one template instantiation per (pre, post, count): 14 k
//     SASS instructions, instruction-cache misses and > 128 live registers (99 us);
//   - `asm volatile` shared loads: a volatile asm keeps its order against every other volatile
//     asm, so "load, wait, blend, store" ran pass after pass with the shared-memory latency
//     exposed each time; the loads are now ordinary (schedulable) asm statements ordered against
//     the barriers by a token operand (73.8 -> 69.4 us);
//   - any value spilled by the register allocator is reloaded at the top of EVERY plane, right
//     where it is needed (15 % of the stall samples were one 4-byte reload): the state carried
//     through the plane loop is kept small and the kernel is bounded to six blocks per SM
//     (80 registers, no spill) rather than seven with spills (69.4 -> 65.9 us).
constexpr int Z2_TR = 16;     // output rows per warp tile
constexpr int Z2_WARPS = 4;   // warps per block: the block covers 128 x 16 outputs
constexpr int Z2_SLOTS = 16;  // rows of 32 floats in a warp's exchange tile (2 KB)

__device__ __forceinline__ unsigned long long pack2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void unpack2(unsigned long long v, float &lo, float &hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ unsigned long long mul2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b,
                                                   unsigned long long c) {
  unsigned long long r;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
  return r;
}
__device__ __forceinline__ unsigned long long add2(unsigned long long a, unsigned long long b) {
  unsigned long long r;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
  return r;
}
// Shared loads that the compiler may SCHEDULE (hoist above independent stores, interleave the
// passes): not volatile -- a volatile asm keeps its order relative to every other volatile asm,
// which would serialise "load, wait, compute, store" pass after pass.  Ordering against the
// writers of the shared tile comes from `tok`: an otherwise unused operand that is re-defined by
// an opaque statement after every barrier (z2_fence), so a load can move neither above the
// barrier that publishes its data nor below the stores that consume it.
__device__ __forceinline__ float4 lds_f4_t(uint32_t addr, int tok) {
  float4 v;
  asm("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4]; // %5"
      : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
      : "r"(addr), "r"(tok));
  return v;
}
__device__ __forceinline__ float lds_f1_t(uint32_t addr, int tok) {
  float v;
  asm("ld.shared.f32 %0, [%1]; // %2" : "=f"(v) : "r"(addr), "r"(tok));
  return v;
}
__device__ __forceinline__ void z2_fence(int &tok) { asm volatile("" : "+r"(tok)::"memory"); }
__device__ __forceinline__ void sts_f1(uint32_t addr, float v) {
  asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ float ldg_nc(const float *p) {
  float v;
  asm volatile("ld.global.nc.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// Phase A of ONE plane: window rows 0 .. wh-1 of the staged plane (this lane's left tap of row 0
// at shared address `a`, rows `pitch` bytes apart) -> exchange rows 0 .. wh-1 (`st` = this lane's
// slot in exchange row 0).  EDGE: the window reaches outside the image, where the staged taps are
// zero and must stay zero under an additive pre-operation (the sampler zero-pads the
// PRE-PROCESSED image); okl / okr / y_lo / H describe which taps those are.
template <int PRE, bool EDGE>
__device__ __forceinline__ void z2_phase_a(uint32_t a, uint32_t pitch, uint32_t st, float wl, float wr,
                                           float mean, int wh, bool okl, bool okr, int y_lo, int H,
                                           int tok) {
  // Rows in groups of Z2_RG, each group fully unrolled with per-row predicates: all taps of a
  // group are in flight before the first blend needs one (a rolled remainder loop would wait for
  // the shared-memory latency row by row; zooming in, a tile has 5 - 8 source rows).
  constexpr int Z2_RG = 4;
  for (int r0 = 0; r0 < wh; r0 += Z2_RG) {
    float tl[Z2_RG], tr[Z2_RG];
#pragma unroll
    for (int k = 0; k < Z2_RG; ++k) {
      tl[k] = 0.f;
      tr[k] = 0.f;
      if (r0 + k < wh) {
        tl[k] = lds_f1_t(a + (uint32_t)k * pitch, tok);
        tr[k] = lds_f1_t(a + (uint32_t)k * pitch + 4u, tok);
      }
    }
#pragma unroll
    for (int k = 0; k < Z2_RG; ++k) {
      if (r0 + k < wh) {
        float l = tl[k], r = tr[k];
        if (PRE == RC_PRE_THRESH02) {
          l = l > 0.2f ? 1.f : 0.f;
          r = r > 0.2f ? 1.f : 0.f;
        } else if (PRE == RC_PRE_ADD_MEAN) {
          if (EDGE) {
            const int ys = y_lo + r0 + k;
            const bool rok = (ys >= 0 && ys <= H - 1);
            l = (rok && okl) ? __fadd_rn(l, mean) : 0.f;
            r = (rok && okr) ? __fadd_rn(r, mean) : 0.f;
          } else {
            l = __fadd_rn(l, mean);
            r = __fadd_rn(r, mean);
          }
        }
        sts_f1(st + 128u * (uint32_t)k, __fmaf_rn(r, wr, __fmul_rn(l, wl)));
      }
    }
    a += Z2_RG * pitch;
    st += 128u * Z2_RG;
  }
}

// Post-operation on four adjacent outputs (columns i .. i+3 of row j) of one plane.
template <int POST>
__device__ __forceinline__ void z2_post(float (&o)[4], float mean, int channel, const ZoomParams &P,
                                        float wx_in, float rwx, bool wx_ok, int i, int j,
                                        bool binary01) {
  if (POST == RC_POST_ROUND) {
    if (binary01) {  // taps are exactly 0 / 1: 0 <= v < 1.5, roundf(v) == (v >= 0.5)
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k] = o[k] >= 0.5f ? 1.f : 0.f;
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k] = round_away(o[k]);
    }
  } else if (POST == RC_POST_ROUND_M045) {
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = round_away(__fsub_rn(o[k], 0.45f));
  } else if (POST == RC_POST_DIV_WX) {
    if (wx_ok) {
      // Markstein: q = a * rb, two FMA refinements -> the correctly rounded quotient whenever
      // nothing over/underflows; a non-finite or overflowing case surfaces as NaN / a huge value
      // and is redone with the IEEE division.  Packed: 5 instructions per pair.
      const unsigned long long b2 = pack2(-wx_in, -wx_in), r2 = pack2(rwx, rwx);
      const unsigned long long a01 = pack2(o[0], o[1]), a23 = pack2(o[2], o[3]);
      unsigned long long q01 = mul2(a01, r2), q23 = mul2(a23, r2);
      unsigned long long e01 = fma2(b2, q01, a01), e23 = fma2(b2, q23, a23);
      q01 = fma2(e01, r2, q01);
      q23 = fma2(e23, r2, q23);
      e01 = fma2(b2, q01, a01);
      e23 = fma2(b2, q23, a23);
      q01 = fma2(e01, r2, q01);
      q23 = fma2(e23, r2, q23);
      float q[4];
      unpack2(q01, q[0], q[1]);
      unpack2(q23, q[2], q[3]);
      const float chk = fmaxf(fmaxf(fabsf(q[0]), fabsf(q[1])), fmaxf(fabsf(q[2]), fabsf(q[3])));
      const float sum = __fadd_rn(__fadd_rn(q[0], q[1]), __fadd_rn(q[2], q[3]));
      if (!(chk <= 1e30f) || sum != sum) {  // rare: redo exactly
#pragma unroll
        for (int k = 0; k < 4; ++k) q[k] = __fdiv_rn(o[k], wx_in);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k] = q[k];
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k) o[k] = __fdiv_rn(o[k], wx_in);
    }
  } else if (POST == RC_POST_MUL_WX) {
#pragma unroll
    for (int k = 0; k < 4; ++k) o[k] = __fmul_rn(o[k], wx_in);
  } else if (POST == RC_POST_SUB_MEAN) {
    const unsigned long long m2 = pack2(-mean, -mean);  // v - m == v + (-m), same rounding
    unpack2(add2(pack2(o[0], o[1]), m2), o[0], o[1]);
    unpack2(add2(pack2(o[2], o[3]), m2), o[2], o[3]);
  } else if (POST == RC_POST_SPOT_SUB_MEAN) {
    const bool yin = (channel == 0) && (j >= P.spot_y0) && (j < P.spot_y1);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float spot = (yin && (i + k) >= P.spot_x0 && (i + k) < P.spot_x1) ? 255.f : 0.f;
      o[k] = __fsub_rn(fmaxf(o[k], spot), mean);
    }
  }
}

// Phase B of ONE plane: the four passes (rows j0 + 4 ps + q) of this thread's four columns.
//   ba[ps]   shared address of this thread's quad in the exchange row of pass ps's UPPER source
//            row (the lower one is the next exchange row)
//   FULL     the tile lies inside the image: no per-row / per-column store predicates
//   DUP      the result also goes to dst2 (RC_ZOOM_SRC_BROADCAST pairs)
template <int POST, int WC, bool FULL, bool DUP>
__device__ __forceinline__ void z2_phase_b(const ZoomParams &P, const uint32_t (&ba)[4],
                                           const float (&wt)[4], const float (&wb)[4], float *dst,
                                           float *dst2, float mean, int channel, bool binary01,
                                           int b, bool col_ok, int tok) {
  const int W = WC ? WC : P.W;
  float wx_in = 1.f, rwx = 1.f;
  bool wx_ok = false;
  if (POST == RC_POST_DIV_WX || POST == RC_POST_MUL_WX) {
    // the sample's own factor (zoom_flow.py:62-66); reloaded here (two of thirteen planes) rather
    // than carried in registers through the plane loop
    wx_in = __ldg(P.zoom_factor + b * 4);
    rwx = __frcp_rn(wx_in);
    const float awx = fabsf(wx_in);
    wx_ok = (awx > 1e-15f) && (awx < 1e15f) && ((__float_as_uint(wx_in) & 0x7fffffu) != 0x7fffffu);
  }
  // output coordinates of this thread's quad (needed by the centre spot and the edge predicates)
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (Z2_WARPS * 32) + (threadIdx.x & ~31) + 4 * (lane & 7);
  const int jq = blockIdx.y * Z2_TR + (lane >> 3);
#pragma unroll
  for (int ps = 0; ps < 4; ++ps) {
    const int j = jq + 4 * ps;
    const unsigned long long wt2 = pack2(wt[ps], wt[ps]), wb2 = pack2(wb[ps], wb[ps]);
    const float4 ht = lds_f4_t(ba[ps], tok);
    const float4 hb = lds_f4_t(ba[ps] + 128u, tok);
    float o[4];
    unpack2(fma2(pack2(hb.x, hb.y), wb2, mul2(pack2(ht.x, ht.y), wt2)), o[0], o[1]);
    unpack2(fma2(pack2(hb.z, hb.w), wb2, mul2(pack2(ht.z, ht.w), wt2)), o[2], o[3]);
    z2_post<POST>(o, mean, channel, P, wx_in, rwx, wx_ok, i, j, binary01);
    if (FULL || (col_ok && j < P.H)) {
      const float4 v = make_float4(o[0], o[1], o[2], o[3]);
      st_f4(reinterpret_cast<float4 *>(dst + 4 * ps * W), v);
      if (DUP) st_f4(reinterpret_cast<float4 *>(dst2 + 4 * ps * W), v);
    }
  }
}

// Fallback for tiles that are not zoom-in shaped (the 16 rows touch more source rows than the
// exchange tile holds: scale > ~0.9, inverse zooms), unaligned sources and degenerate factors:
// every thread gathers its own outputs tap by tap, any pre / post operation.  Kept out of line:
// it must not cost the fast path registers.
__device__ __noinline__ void z2_gather_tile(const ZoomParams &P, int first, bool dup, int b, int i,
                                            int j0, int q) {
  const int H = P.H, W = P.W;
  float wx, wy, tx, ty, wx_in;
  zoom_affine(P.zoom_factor + b * 4, P.inverse, H, W, wx, wy, tx, ty, wx_in);
  const float stepx = (float)(2.0 / (W - 1)), stepy = (float)(2.0 / (H - 1));
  const float nxm1 = (float)(W - 1), nym1 = (float)(H - 1);
  const float rwx = __frcp_rn(wx_in);
  if (i >= W) return;
  const ZoomPlane &zp = P.planes[first];
  const float *src = zp.src + (long long)b * zp.src_bstride;
  for (int ps = 0; ps < 4; ++ps) {
    const int j = j0 + 4 * ps + q;
    if (j >= H) continue;
    const RowCoord rc = row_coord(j, stepy, wy, ty, nym1);
    const int y0 = rc.y0;
    const bool yin0 = (y0 >= 0) && (y0 <= H - 1), yin1 = (y0 + 1 >= 0) && (y0 + 1 <= H - 1);
    float o[4];
    for (int k = 0; k < 4; ++k) {
      const float xr = src_coord(i + k, stepx, wx, tx, nxm1);
      const float xf = floorf(xr);
      const int x0 = (fabsf(xf) < 1e8f) ? (int)xf : -(1 << 30);
      const float wl = __fsub_rn(1.0f, __fsub_rn(xr, xf)), wr = __fsub_rn(1.0f, wl);
      const bool xin0 = (x0 >= 0) && (x0 <= W - 1), xin1 = (x0 + 1 >= 0) && (x0 + 1 <= W - 1);
      const float *r0 = src + (long long)y0 * W + x0;
      const float tl = (yin0 && xin0) ? pre(ldg_nc(r0), zp.pre_op, zp.mean) : 0.f;
      const float tr = (yin0 && xin1) ? pre(ldg_nc(r0 + 1), zp.pre_op, zp.mean) : 0.f;
      const float bl = (yin1 && xin0) ? pre(ldg_nc(r0 + W), zp.pre_op, zp.mean) : 0.f;
      const float br = (yin1 && xin1) ? pre(ldg_nc(r0 + W + 1), zp.pre_op, zp.mean) : 0.f;
      const float ht = __fmaf_rn(tr, wr, __fmul_rn(tl, wl));
      const float hb = __fmaf_rn(br, wr, __fmul_rn(bl, wl));
      o[k] = __fmaf_rn(hb, rc.wb, __fmul_rn(ht, rc.wt));
    }
    switch (zp.post_op) {
      case RC_POST_ROUND: z2_post<RC_POST_ROUND>(o, zp.mean, zp.channel, P, wx_in, rwx, false, i, j, false); break;
      case RC_POST_ROUND_M045: z2_post<RC_POST_ROUND_M045>(o, zp.mean, zp.channel, P, wx_in, rwx, false, i, j, false); break;
      case RC_POST_DIV_WX: z2_post<RC_POST_DIV_WX>(o, zp.mean, zp.channel, P, wx_in, rwx, false, i, j, false); break;
      case RC_POST_MUL_WX: z2_post<RC_POST_MUL_WX>(o, zp.mean, zp.channel, P, wx_in, rwx, false, i, j, false); break;
      case RC_POST_SUB_MEAN: z2_post<RC_POST_SUB_MEAN>(o, zp.mean, zp.channel, P, wx_in, rwx, false, i, j, false); break;
      case RC_POST_SPOT_SUB_MEAN: z2_post<RC_POST_SPOT_SUB_MEAN>(o, zp.mean, zp.channel, P, wx_in, rwx, false, i, j, false); break;
      default: break;
    }
    const float4 v = make_float4(o[0], o[1], o[2], o[3]);
    const long long off = (long long)j * W + i;
    st_f4(reinterpret_cast<float4 *>(zp.dst + (long long)b * zp.dst_bstride + off), v);
    if (dup) {
      const ZoomPlane &z2 = P.planes[first + 1];
      st_f4(reinterpret_cast<float4 *>(z2.dst + (long long)b * z2.dst_bstride + off), v);
    }
  }
}

// grid = (ceil(W/128), ceil(H/16), B); block = 128 (4 warps); 8 KB static exchange tiles +
// `smem_floats` * 4 bytes of dynamic staging.
template <int WC>
__global__ void __launch_bounds__(Z2_WARPS * 32, 6)
zoom_sample2_kernel(const __grid_constant__ ZoomParams P, int smem_floats) {
  extern __shared__ float4 win4[];
  __shared__ __align__(16) float xch[Z2_WARPS][Z2_SLOTS][32];
  const int H = P.H, W = WC ? WC : P.W;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int b = blockIdx.z;
  const int ib = blockIdx.x * (Z2_WARPS * 32);  // first column of the block
  const int i0 = ib + warp * 32, j0c = blockIdx.y * Z2_TR;
  const bool warp_in = i0 < W;  // a warp right of the image only helps staging and meets barriers

  // ---- state carried through the plane loop (kept small on purpose: a spilled value costs a
  // local-memory round trip per plane on the critical path)
  uint32_t tap;          // shared address of this lane's left tap in window row 0 of staged plane 0
  uint32_t pitch;        // bytes between window rows
  uint32_t st_lane;      // this lane's slot in row 0 of the warp's exchange tile (phase A)
  float wl, wr;          // phase A column weights
  uint32_t ba[4];        // phase B: exchange address of the upper source row, per pass
  float rwt[4], rwb[4];  // phase B: row weights, per pass
  int y_first, wh_all, x_a;
  bool edge, okl, okr, staged;
  {
    float wx, wy, tx, ty, wx_in;
    zoom_affine(P.zoom_factor + b * 4, P.inverse, H, W, wx, wy, tx, ty, wx_in);
    const float stepx = (float)(2.0 / (W - 1)), stepy = (float)(2.0 / (H - 1));
    const float nxm1 = (float)(W - 1), nym1 = (float)(H - 1);
    // phase A: this lane's column
    const float xr = src_coord(min(i0 + lane, W - 1), stepx, wx, tx, nxm1);
    const float xf = floorf(xr);
    const int x0 = (fabsf(xf) < 1e6f) ? (int)xf : -(1 << 20);
    wl = __fsub_rn(1.0f, __fsub_rn(xr, xf));
    wr = __fsub_rn(1.0f, wl);
    okl = (x0 >= 0) && (x0 <= W - 1);
    okr = (x0 + 1 >= 0) && (x0 + 1 <= W - 1);
    // the block's source window (block-uniform): columns of outputs ib .. ib+127, rows of the tile
    const float fx_lo = floorf(src_coord(ib, stepx, wx, tx, nxm1));
    const float fx_hi = floorf(src_coord(min(ib + Z2_WARPS * 32, W) - 1, stepx, wx, tx, nxm1));
    const float fy_lo = floorf(src_coord(j0c, stepy, wy, ty, nym1));
    const float fy_hi = floorf(src_coord(min(j0c + Z2_TR, H) - 1, stepy, wy, ty, nym1));
    const bool regular = (wx > 0.f) && (wy > 0.f) && (fy_hi >= fy_lo) && (fx_hi >= fx_lo) &&
                         (fabsf(fy_lo) < 1e6f) && (fabsf(fy_hi) < 1e6f) && (fabsf(fx_lo) < 1e6f) &&
                         (fabsf(fx_hi) < 1e6f) && (fy_hi - fy_lo + 2.0f <= (float)Z2_SLOTS) &&
                         P.src_aligned;
    y_first = regular ? (int)fy_lo : 0;
    wh_all = regular ? (int)(fy_hi - fy_lo) + 2 : 2;
    x_a = regular ? ((int)fx_lo & ~3) : 0;  // floor to a multiple of 4 (also for negative values)
    const int nch = regular ? ((((int)fx_hi + 1 - x_a) >> 2) + 1) : 1;
    pitch = 16u * (uint32_t)nch;
    // every decision below that involves a barrier must be block-uniform: the window and the zoom
    // factor are, but make it explicit -- vote
    const bool reg_all = __syncthreads_and(regular ? 1 : 0) != 0;
    staged = reg_all && nch * 4 * wh_all <= smem_floats;
    // any tap of the WARP's tile outside the image?
    const bool edge_ = !okl || !okr || y_first < 0 || y_first + wh_all > H;
    edge = __any_sync(0xffffffffu, edge_);
    const uint32_t hbase = (uint32_t)__cvta_generic_to_shared(&xch[warp][0][0]);
    st_lane = hbase + 4u * (uint32_t)lane;
    tap = (uint32_t)__cvta_generic_to_shared(win4) + 4u * (uint32_t)(x0 - x_a);
    // phase B: this lane's rows (one per pass); exchange addresses are valid for staged tiles,
    // where the row's upper source row lies in [y_first, y_first + wh_all - 2]
#pragma unroll
    for (int ps = 0; ps < 4; ++ps) {
      const RowCoord rc = row_coord(min(j0c + 4 * ps + (lane >> 3), H - 1), stepy, wy, ty, nym1);
      ba[ps] = hbase + 128u * (uint32_t)(rc.y0 - y_first) + 16u * (uint32_t)(lane & 7);
      rwt[ps] = rc.wt;
      rwb[ps] = rc.wb;
    }
  }
  const int i_tile = i0 + 4 * (lane & 7);
  // W % 4 == 0: a quad is inside or outside as a whole; with the compile-time width a multiple of
  // the block's 128 columns every warp and quad is inside
  const bool col_ok = (WC && WC % (Z2_WARPS * 32) == 0) ? true : (warp_in && i_tile < W);
  // warp-uniform "no store predicates needed" (the vote is executed by every lane: no short-circuit)
  const bool all_cols = __all_sync(0xffffffffu, col_ok);
  const bool full = all_cols && (j0c + Z2_TR <= H);

  // sparse frame, per warp tile (see version 1): skip the sparse planes where the source window
  // touches no covered tile and the outputs already hold the background
  unsigned wskip = 0u;
  const bool warp_in_u = __any_sync(0xffffffffu, warp_in);
  if (P.tile_state != nullptr && warp_in_u) {
    uint8_t *st = P.tile_state +
                  (((size_t)b * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x) * Z2_WARPS + warp;
    // (votes make the values provably warp-uniform, so the skip mask can live in a uniform register)
    const bool held = __any_sync(0xffffffffu, *st != 0);
    const int x0 = x_a + (int)((tap - (uint32_t)__cvta_generic_to_shared(win4)) >> 2);
    const int xw0 = __shfl_sync(0xffffffffu, x0, 0), xw1 = __shfl_sync(0xffffffffu, x0, 31) + 1;
    int any = 1;
    if (staged && !P.inverse) {
      const int sx0 = max(xw0, 0), sx1 = min(xw1, W - 1);
      const int sy0 = max(y_first, 0), sy1 = min(y_first + wh_all - 1, H - 1);
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
    __syncwarp();
    if (lane == 0) *st = bg_now ? 1 : 0;
    if (bg_now && P.state_valid && held) wskip = 1u;  // skip the planes of P.sparse_mask
  }
  if (!warp_in_u) wskip = 2u;  // skip every plane
  // The two skip bits ride in the upper bits of wh_all: as a variable of their own they were
  // the value the register allocator chose to spill, and the reload sat on the critical path of
  // every plane (15 % of the stall samples).
  wh_all |= (int)(wskip << 8);

  if (!staged) {
    // not zoom-in shaped / degenerate / unaligned: every thread gathers its own outputs
    if (warp_in)
      for (int pl = 0; pl < P.n_planes; ++pl) {
        if (P.planes[pl].dup_second) continue;
        const bool dup = pl + 1 < P.n_planes && P.planes[pl + 1].dup_second;
        z2_gather_tile(P, pl, dup, b, i_tile, j0c, lane >> 3);
      }
    return;
  }
  int p0 = 0, p1 = 0;  // planes [p0, p1) are staged; slot = plane - p0
  int tok = 0;         // ordering token of the schedulable shared loads (lds_f*_t)
  for (int pl = 0; pl < P.n_planes; ++pl) {
    const ZoomPlane &zp = P.planes[pl];
    if (zp.dup_second) continue;  // written together with the plane before it
    const int wh = wh_all & 0xff;
    const uint32_t plane_bytes = pitch * (uint32_t)wh;
    if (pl >= p1) {  // block-uniform: stage the next planes
      if (p1 > 0) __syncthreads();  // every consumer of the previous stage is done
      p0 = pl;
      p1 = min(P.n_planes, p0 + smem_floats / (int)(plane_bytes >> 2));
      const int nch = (int)(pitch >> 4);
      const int nchunks = nch * wh;
      const float inv_nch = __frcp_rn((float)nch);
      const uint32_t win_s = (uint32_t)__cvta_generic_to_shared(win4);
      for (int qq = threadIdx.x; qq < nchunks; qq += Z2_WARPS * 32) {
        int r = __float2int_rz(__fmul_rn((float)qq + 0.5f, inv_nch));
        int c = qq - r * nch;
        if (c < 0) {
          c += nch;
          --r;
        } else if (c >= nch) {
          c -= nch;
          ++r;
        }
        const int ys = y_first + r, xs = x_a + 4 * c;
        const bool inb = (xs >= 0) && (xs + 3 <= W - 1) && (ys >= 0) && (ys <= H - 1);
        const int goff = inb ? ys * W + xs : 0;
        const int bytes = inb ? 16 : 0;
        uint32_t sa = win_s + 16u * (uint32_t)qq;
        for (int k = p0; k < p1; ++k, sa += plane_bytes) {
          const ZoomPlane &zk = P.planes[k];
          if (zk.dup_second) continue;
          cp_async_16(sa, zk.src + (long long)b * zk.src_bstride + goff, bytes);
        }
      }
      cp_async_wait_all();
      __syncthreads();
      z2_fence(tok);
    }
    if ((wh_all >> 9) | ((wh_all >> 8) & (int)(P.sparse_mask >> pl) & 1)) continue;  // warp-uniform
    // shared address of this lane's first tap of the plane
    uint32_t a_pl = tap + (uint32_t)(pl - p0) * plane_bytes;
    // The phase bodies below are separate template instantiations inlined into this loop; the
    // tile constants are made opaque once per iteration so that the compiler does not hoist
    // every instantiation's derived values (packed weight pairs, addresses) out of the loop.
    asm volatile("" : "+r"(st_lane), "+r"(a_pl));
#pragma unroll
    for (int ps = 0; ps < 4; ++ps) asm volatile("" : "+r"(ba[ps]), "+f"(rwt[ps]), "+f"(rwb[ps]));
    const int pre_op = zp.pre_op, post_op = zp.post_op;
    const float mean = zp.mean;
    if (pre_op == RC_PRE_ADD_MEAN) {
      if (edge) z2_phase_a<RC_PRE_ADD_MEAN, true>(a_pl, pitch, st_lane, wl, wr, mean, wh, okl, okr, y_first, H, tok);
      else z2_phase_a<RC_PRE_ADD_MEAN, false>(a_pl, pitch, st_lane, wl, wr, mean, wh, okl, okr, y_first, H, tok);
    } else if (pre_op == RC_PRE_THRESH02) {
      z2_phase_a<RC_PRE_THRESH02, false>(a_pl, pitch, st_lane, wl, wr, mean, wh, okl, okr, y_first, H, tok);
    } else {
      z2_phase_a<RC_PRE_NONE, false>(a_pl, pitch, st_lane, wl, wr, mean, wh, okl, okr, y_first, H, tok);
    }
    __syncwarp();
    z2_fence(tok);
    // row of pass 0, this thread's columns (recomputed per plane: cheaper than a register held
    // across the loop)
    const int tile_off = (blockIdx.y * Z2_TR + ((threadIdx.x & 31) >> 3)) * W +
                         blockIdx.x * (Z2_WARPS * 32) + (threadIdx.x & ~31) + 4 * (threadIdx.x & 7);
    float *dst = zp.dst + (long long)b * zp.dst_bstride + tile_off;
    const bool dup = pl + 1 < P.n_planes && P.planes[pl + 1].dup_second;
    float *dst2 = dup ? P.planes[pl + 1].dst + (long long)b * P.planes[pl + 1].dst_bstride + tile_off
                      : nullptr;
    const bool bin = pre_op == RC_PRE_THRESH02;
#define Z2_B2(POST_, FULL_, DUP_) \
  z2_phase_b<POST_, WC, FULL_, DUP_>(P, ba, rwt, rwb, dst, dst2, mean, zp.channel, bin, b, col_ok, tok)
#define Z2_B(POST_)                           \
  do {                                        \
    if (dup) Z2_B2(POST_, false, true);       \
    else if (full) Z2_B2(POST_, true, false); \
    else Z2_B2(POST_, false, false);          \
  } while (0)
    switch (post_op) {
      case RC_POST_ROUND: Z2_B(RC_POST_ROUND); break;
      case RC_POST_ROUND_M045: Z2_B(RC_POST_ROUND_M045); break;
      case RC_POST_DIV_WX: Z2_B(RC_POST_DIV_WX); break;
      case RC_POST_MUL_WX: Z2_B(RC_POST_MUL_WX); break;
      case RC_POST_SUB_MEAN: Z2_B(RC_POST_SUB_MEAN); break;
      case RC_POST_SPOT_SUB_MEAN: Z2_B(RC_POST_SPOT_SUB_MEAN); break;
      default: Z2_B(RC_POST_NONE); break;
    }
#undef Z2_B
#undef Z2_B2
    __syncwarp();  // the exchange tile is rewritten by the next plane
    z2_fence(tok);
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
  zoom_factor_from_boxes(b, observed_box, ws, src_pose, K, zoom_factor, bbox, trans_in, trans_out, H,
                         W);
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
      z.dup_second = 0;
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
  RC_REQUIRE(batch_size <= 65535, RC_ERR_INVALID_ARG);
  // ---- version 2 (zoom_sample2_kernel) whenever every plane is a plain write of one of the
  // operation pairs it is instantiated for; RC_ZOOM_V1=1 forces version 1 (A/B measurements)
  {
    static int force_v1 = -1;
    if (force_v1 < 0) {
      const char *e = getenv("RC_ZOOM_V1");
      force_v1 = (e && e[0] == '1') ? 1 : 0;
    }
    bool v2 = !force_v1;
    for (int pl = 0; pl < P.n_planes && v2; ++pl) {
      const rc::ZoomPlane &z = P.planes[pl];
      if (z.accumulate) v2 = false;
      const bool ok_pair =
          (z.pre_op == RC_PRE_ADD_MEAN &&
           (z.post_op == RC_POST_SUB_MEAN || z.post_op == RC_POST_SPOT_SUB_MEAN)) ||
          (z.pre_op == RC_PRE_THRESH02 && z.post_op == RC_POST_ROUND) ||
          (z.pre_op == RC_PRE_NONE &&
           (z.post_op == RC_POST_NONE || z.post_op == RC_POST_ROUND || z.post_op == RC_POST_ROUND_M045 ||
            z.post_op == RC_POST_DIV_WX || z.post_op == RC_POST_MUL_WX));
      if (!ok_pair) v2 = false;
      // phase B writes 128-bit quads: destination planes 16-byte aligned (checked above per group)
    }
    if (v2) {
      // a broadcast pair (RC_ZOOM_SRC_BROADCAST: the next plane resamples the SAME source with the
      // same operations) is computed once and stored twice
      for (int pl = 0; pl + 1 < P.n_planes; ++pl) {
        const rc::ZoomPlane &z0 = P.planes[pl];
        rc::ZoomPlane &z1 = P.planes[pl + 1];
        if (!z0.dup_second && z1.src == z0.src && z1.src_bstride == z0.src_bstride &&
            z1.pre_op == z0.pre_op && z1.post_op == z0.post_op && z1.mean == z0.mean &&
            z0.post_op != RC_POST_SPOT_SUB_MEAN &&
            ((P.sparse_mask >> (pl + 1)) & 1u) == ((P.sparse_mask >> pl) & 1u))
          z1.dup_second = 1;
      }
      // staging: 22 KB per block (all 13 planes of the default step in one stage up to zoom
      // factors ~0.37: window (128 wx + 8) x (16 wy + 3) floats per plane) + 8 KB exchange tiles:
      // seven blocks per SM
      constexpr int kStageFloats = 5632;
      static bool attr2_set = false;
      if (!attr2_set) {
        RC_CUDA_TRY(cudaFuncSetAttribute(rc::zoom_sample2_kernel<640>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kStageFloats * (int)sizeof(float)));
        RC_CUDA_TRY(cudaFuncSetAttribute(rc::zoom_sample2_kernel<0>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kStageFloats * (int)sizeof(float)));
        attr2_set = true;
      }
      dim3 grid2((width + 127) / 128, (height + rc::Z2_TR - 1) / rc::Z2_TR, batch_size);
      if (width == 640)
        rc::zoom_sample2_kernel<640><<<grid2, rc::Z2_WARPS * 32, kStageFloats * sizeof(float),
                                       static_cast<cudaStream_t>(stream)>>>(P, kStageFloats);
      else
        rc::zoom_sample2_kernel<0><<<grid2, rc::Z2_WARPS * 32, kStageFloats * sizeof(float),
                                     static_cast<cudaStream_t>(stream)>>>(P, kStageFloats);
      return rc::note_status(rc::check_launch());
    }
  }
  dim3 grid((width + rc::ZT_W - 1) / rc::ZT_W, (height + rc::ZT_H - 1) / rc::ZT_H, batch_size);
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
