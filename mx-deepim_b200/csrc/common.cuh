// common.cuh -- shared helpers for librendercompare (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "rendercompare.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "librendercompare is written for sm_100a (B200) only"
#endif

namespace rc {

// Recorded by every failing CUDA call so that rc_last_cuda_error() can report it.
void set_cuda_error(cudaError_t e);
int note_status(int status);
// status.cu: SM count of `device` (< 0 = current device), cached per device.
int device_sm_count(int device = -1);

#define RC_CUDA_TRY(expr)                         \
  do {                                            \
    cudaError_t _e = (expr);                      \
    if (_e != cudaSuccess) {                      \
      ::rc::set_cuda_error(_e);                   \
      return ::rc::note_status(RC_ERR_CUDA);      \
    }                                             \
  } while (0)

#define RC_REQUIRE(cond, code)                    \
  do {                                            \
    if (!(cond)) return ::rc::note_status(code);  \
  } while (0)

// zoom.cu: bbox + zoom factor, optionally with ZoomTrans forward fused into the finalising
// block (trans_in/trans_out both NULL or both set).
int launch_zoom_factor(const float *obs, const float *rend, const float *src_pose,
                       const float *K, const float *pixel_means, float *zoom_factor,
                       int32_t *bbox, int32_t *workspace, const float *trans_in,
                       float *trans_out, int batch_size, int channels, int height, int width,
                       int mode, void *stream);

// zoom.cu: zoom factor from precomputed boxes: the observed box from `observed_box` [B,4], the
// rendered box from the workspace entries 4..7 (filled by frame.cu's tile pass, reset here).
int launch_zoom_factor_boxes(const int32_t *observed_box, const float *src_pose, const float *K,
                             float *zoom_factor, int32_t *bbox, int32_t *workspace,
                             const float *trans_in, float *trans_out, int batch_size, int height,
                             int width, void *stream);

// raster.cu: the (tile, pair) items of the renderer's last frame and of the one before.
// Items are packed pair << 16 | tile row << 8 | tile column over 16x16-pixel tiles.
struct FrameItems {
  const int32_t *cur_list, *cur_count;    // tiles the last frame covers (count: device int)
  const int32_t *prev_list, *prev_count;  // tiles the frame before covered; NULL if there was none
  const int32_t *tile_count;              // [max_batch][tiles] > 0 <=> the tile is in cur_list
  int tiles_x, tiles_y, H, W, max_batch;
};
int renderer_frame_items(void *handle, FrameItems *out);

// frame.cu: rendered crop box (into workspace entries 4..7) and, with flow != NULL, GT flow on
// the frame's tiles (+ zeros on the previous frame's stale tiles when zero_prev).
// `final` (or NULL): what the pass's last block needs to turn the boxes into zoom factors
// (+ ZoomTrans forward) itself instead of leaving that to launch_zoom_factor_boxes.
struct FrameFinal {
  const int32_t *observed_box;
  const float *src_pose, *K;
  float *zoom_factor;
  int32_t *bbox;
  const float *trans_in;
  float *trans_out;
};
int launch_frame_tiles(const FrameItems &F, const float *depth_src, const float *depth_tgt,
                       const float *KT, const float *Kinv, float *flow, float *valid,
                       int32_t *workspace, int batch_size, int height, int width, bool zero_prev,
                       const FrameFinal *final, void *stream);

// zoom.cu: rc_zoom_sample_dev with the compare step's sparse-frame information (or NULL).
// One state byte per warp of an output tile (128 x 8 pixels, 4 warps of 32 columns):
// tile_state is [B][ceil(H/8) * ceil(W/128) * 4].
struct ZoomSparse {
  const int32_t *tile_count;  // the renderer's per-(pair, 16x16 tile) record counts of the frame
  uint8_t *tile_state;
  int tiles_x, tiles_y;
  bool state_valid;
};
int launch_zoom_sample(const float *zoom_factor, int inverse, const rc_zoom_group *groups,
                       int n_groups, int batch_size, int height, int width,
                       const ZoomSparse *sparse, void *stream);

// Set by rc_compare_step_dev around launches that may overlap their predecessor in the stream
// (programmatic dependent launch: the kernel does not consume the predecessor's output).
inline thread_local bool tl_pdl = false;

#ifdef __CUDACC__
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_maybe_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block,
                                           size_t smem, cudaStream_t s, bool pdl, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  if (pdl) {
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
  }
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}
#endif

// Event mark on `s`.  While `s` is being captured into a CUDA graph a plain cudaEventRecord would
// only create a dependency edge; cudaEventRecordExternal makes it a real event-record NODE, so
// that every replay of the graph re-records the event and cudaEventElapsedTime works on it --
// this is how bench.py times kernels inside a replayed refine loop.
static inline cudaError_t record_mark(cudaEvent_t e, cudaStream_t s) {
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &cs) == cudaSuccess && cs == cudaStreamCaptureStatusActive)
    return cudaEventRecordWithFlags(e, s, cudaEventRecordExternal);
  return cudaEventRecord(e, s);
}

static inline bool aligned16(const void *p) {
  return (reinterpret_cast<uintptr_t>(p) & 15u) == 0;
}

static inline int check_launch() {
  cudaError_t e = cudaPeekAtLastError();
  if (e != cudaSuccess) {
    (void)cudaGetLastError();
    set_cuda_error(e);
    return note_status(RC_ERR_CUDA);
  }
  return RC_OK;
}

// ---------------------------------------------------------------- device side
#ifdef __CUDACC__

// 128-bit read-only load that does not allocate in L1 (streamed once).
__device__ __forceinline__ float4 ld_stream_f4(const float4 *p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(p));
  return r;
}

// 128-bit store; default L2 policy so that a consumer kernel can still hit L2.
__device__ __forceinline__ void st_f4(float4 *p, const float4 &v) {
  asm volatile("st.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// Lets the next kernel in the stream start (if it was launched with programmatic stream
// serialization); a no-op otherwise.
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// Blocks until the grids this one was programmatically launched after have completed and
// their writes are visible; a no-op for a normal launch.
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__device__ __forceinline__ void st_f1(float *p, float v) {
  asm volatile("st.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int warp_min(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ int warp_max(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

#endif  // __CUDACC__

}  // namespace rc
