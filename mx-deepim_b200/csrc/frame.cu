// frame.cu -- the compare step's two full-frame passes over the RENDERED side, restricted to
// the tiles the rasteriser says the frame covers.
//
// Replaces, when the caller also drives the rasteriser (pipeline.RefineLoop):
//   * the rendered half of ZoomMask's crop-box reduction
//     (/root/reference/deepim/operator_py/zoom_mask.py:40-62: rows/columns where
//     mask_rendered > 0.2), and
//   * gpu_flow over the whole plane
//     (/root/reference/lib/flow_c/gpu_flow_kernel.cu:32-69), whose output is zero wherever
//     depth_src <= 1e-3, i.e. everywhere outside the rendered object.
//
// B200 design: both passes are pure HBM streams of planes that are ~90 % background, and the
// rasteriser already holds the list of 16x16 tiles that are not background (its work items).
// One persistent-style launch walks that list: 64 threads per tile, one 128-bit load of the
// rendered depth per thread, the reference's per-pixel flow rule (flow_device.cuh, bit-exact
// with flow.cu), three 128-bit stores, and a warp-reduced crop box merged with 4 atomicMax
// per warp into the zoom-factor workspace (same all-zero-when-idle encoding as zoom.cu).
// Tiles the previous frame covered and this one does not are re-zeroed in the same launch, so
// the flow planes always equal what the dense kernel would have written.  8 + 12 B/px of
// traffic over the whole frame become 20 B/px over the covered tiles only.
#include <limits.h>

#include "common.cuh"
#include "flow_device.cuh"
#include "zoom_finalize.cuh"

namespace rc {

constexpr int FT_EDGE = 16;          // the rasteriser's tile edge (raster.cu RT_TW == RT_TH)
constexpr int FT_ITEMS_PER_BLOCK = 4;  // 64 threads (two warps) per tile

// grid-stride over the items of the last frame (compute) followed by the items of the frame
// before (zero if stale); block = 256.
template <bool DO_FLOW>
__global__ void __launch_bounds__(256, 4)
frame_tiles_kernel(FrameItems F, const float *__restrict__ depth_src,
                   const float *__restrict__ depth_tgt, const float *__restrict__ KT,
                   const float *__restrict__ Kinv, float *__restrict__ flow,
                   float *__restrict__ valid, int32_t *__restrict__ ws, int B, int H, int W,
                   int zero_prev, FrameFinal fin) {
  pdl_launch_dependents();  // the zoom-factor kernel behind this one waits for the grid's end
  const int n_cur = *F.cur_count;
  const int n_prev = (DO_FLOW && zero_prev) ? *F.prev_count : 0;
  const int n_tiles = F.tiles_x * F.tiles_y;
  const int sub = threadIdx.x >> 6, quad = threadIdx.x & 63;
  const int ry = quad >> 2, qx = (quad & 3) << 2;
  const size_t plane = (size_t)H * W;
  for (int base = blockIdx.x * FT_ITEMS_PER_BLOCK; base < n_cur + n_prev;
       base += gridDim.x * FT_ITEMS_PER_BLOCK) {
    const int i = base + sub;  // warp-uniform
    if (i >= n_cur + n_prev) continue;
    const bool is_cur = i < n_cur;
    const int item = is_cur ? __ldg(F.cur_list + i) : __ldg(F.prev_list + (i - n_cur));
    const int b = item >> 16, ty = (item >> 8) & 255, tx = item & 255;
    if (b >= B) continue;
    const int y = ty * FT_EDGE + ry, x = tx * FT_EDGE + qx;
    const bool inside = y < H && x < W;  // W % 4 == 0: a quad is inside or outside as a whole
    const size_t off = (size_t)b * plane + (size_t)y * W + x;
    if (!is_cur) {
      // stale tile of the previous frame: the dense kernel would write zeros here
      if (inside && __ldg(F.tile_count + (size_t)b * n_tiles + ty * F.tiles_x + tx) == 0) {
        const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
        float *f = flow + (size_t)b * 2 * plane + (size_t)y * W + x;
        st_f4(reinterpret_cast<float4 *>(f), z);
        st_f4(reinterpret_cast<float4 *>(f + plane), z);
        st_f4(reinterpret_cast<float4 *>(valid + off), z);
      }
      continue;
    }
    int x0 = INT_MAX, x1 = -1, y0 = INT_MAX, y1 = -1;
    if (inside) {
      const float4 d = ld_stream_f4(reinterpret_cast<const float4 *>(depth_src + off));
      // zoom_mask.py:40-45: mask_rendered > 0.2 -> 1, then "> 0.3" of the 0/1 mask
      const unsigned m = (d.x > 0.2f ? 1u : 0u) | (d.y > 0.2f ? 2u : 0u) | (d.z > 0.2f ? 4u : 0u) |
                         (d.w > 0.2f ? 8u : 0u);
      if (m) {
        x0 = x + __ffs(m) - 1;
        x1 = x + 31 - __clz(m);
        y0 = y1 = y;
      }
      if (DO_FLOW) {
        FlowCoef c;
        load_coef(c, KT, Kinv, b);
        const float *tgt = depth_tgt + (size_t)b * plane;
        const float fh = (float)y;
        float4 o0, o1, ov;
        flow_pixel(d.x, (float)(x + 0), fh, c, tgt, H, W, o0.x, o1.x, ov.x);
        flow_pixel(d.y, (float)(x + 1), fh, c, tgt, H, W, o0.y, o1.y, ov.y);
        flow_pixel(d.z, (float)(x + 2), fh, c, tgt, H, W, o0.z, o1.z, ov.z);
        flow_pixel(d.w, (float)(x + 3), fh, c, tgt, H, W, o0.w, o1.w, ov.w);
        float *f = flow + (size_t)b * 2 * plane + (size_t)y * W + x;
        st_f4(reinterpret_cast<float4 *>(f), o0);
        st_f4(reinterpret_cast<float4 *>(f + plane), o1);
        st_f4(reinterpret_cast<float4 *>(valid + off), ov);
      }
    }
    // the warp belongs to one item, hence one pair: reduce, then 4 atomics per warp
    x0 = warp_min(x0);
    x1 = warp_max(x1);
    y0 = warp_min(y0);
    y1 = warp_max(y1);
    if ((threadIdx.x & 31) == 0 && x1 >= 0) {
      int32_t *w = ws + (size_t)b * RC_ZF_WS_INTS;
      atomicMax(w + 4, INT_MAX - x0);
      atomicMax(w + 5, x1 + 1);
      atomicMax(w + 6, INT_MAX - y0);
      atomicMax(w + 7, y1 + 1);
    }
  }
  pdl_wait();  // completion of this grid implies completion of the grid it was launched behind
  if (fin.zoom_factor == nullptr) return;
  // ---- the LAST block to finish turns the boxes into zoom factors (+ ZoomTrans) for every pair:
  // one launch less on the iteration's chain of dependent kernels (was zoom_factor_boxes_kernel
  // behind this one).  Ticket in the unused workspace entry 9 of pair 0; re-armed here.
  __shared__ int s_last;
  __threadfence();  // this block's box merges before its ticket
  __syncthreads();
  if (threadIdx.x == 0) s_last = atomicAdd(ws + 9, 1) == (int)gridDim.x - 1;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x == 0) ws[9] = 0;
  for (int b = threadIdx.x; b < B; b += blockDim.x)
    zoom_factor_from_boxes(b, fin.observed_box, ws, fin.src_pose, fin.K, fin.zoom_factor, fin.bbox,
                           fin.trans_in, fin.trans_out, H, W);
}

int launch_frame_tiles(const FrameItems &F, const float *depth_src, const float *depth_tgt,
                       const float *KT, const float *Kinv, float *flow, float *valid,
                       int32_t *workspace, int batch_size, int height, int width, bool zero_prev,
                       const FrameFinal *final, void *stream) {
  RC_REQUIRE(depth_src && workspace && F.cur_list && F.cur_count && F.tile_count,
             RC_ERR_INVALID_ARG);
  RC_REQUIRE(F.H == height && F.W == width && batch_size <= F.max_batch, RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && aligned16(depth_src), RC_ERR_UNALIGNED);
  const bool do_flow = flow != nullptr;
  if (do_flow) {
    RC_REQUIRE(depth_tgt && KT && Kinv && valid, RC_ERR_INVALID_ARG);
    RC_REQUIRE(aligned16(flow) && aligned16(valid), RC_ERR_UNALIGNED);
    RC_REQUIRE(!zero_prev || (F.prev_list && F.prev_count), RC_ERR_INVALID_ARG);
  }
  const int sm_count = device_sm_count();
  // the item counts live on the device: a fixed grid of a few blocks per SM strides over them
  long long blocks = ((long long)batch_size * F.tiles_x * F.tiles_y + FT_ITEMS_PER_BLOCK - 1) /
                     FT_ITEMS_PER_BLOCK;
  if (blocks > (long long)sm_count * 8) blocks = (long long)sm_count * 8;
  if (blocks < 1) blocks = 1;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int zp = zero_prev ? 1 : 0;
  FrameFinal fin = {};
  if (final) {
    fin = *final;
    RC_REQUIRE(fin.observed_box && fin.src_pose && fin.K && fin.zoom_factor, RC_ERR_INVALID_ARG);
    RC_REQUIRE((fin.trans_in == nullptr) == (fin.trans_out == nullptr), RC_ERR_INVALID_ARG);
  }
  if (do_flow)
    RC_CUDA_TRY(launch_maybe_pdl(frame_tiles_kernel<true>, dim3((unsigned)blocks), dim3(256), 0, s,
                                 tl_pdl, F, depth_src, depth_tgt, KT, Kinv, flow, valid,
                                 workspace, batch_size, height, width, zp, fin));
  else
    RC_CUDA_TRY(launch_maybe_pdl(frame_tiles_kernel<false>, dim3((unsigned)blocks), dim3(256), 0, s,
                                 tl_pdl, F, depth_src, depth_tgt, KT, Kinv, flow, valid,
                                 workspace, batch_size, height, width, zp, fin));
  return note_status(check_launch());
}

}  // namespace rc
