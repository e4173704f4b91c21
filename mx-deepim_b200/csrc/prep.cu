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

}  // extern "C"
