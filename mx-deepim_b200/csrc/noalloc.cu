// noalloc.cu -- rc_zoom_factor_noalloc_dev: the crop-box / zoom-factor kernel for callers that
// cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  No counterpart in the reference.
//
// rc_zoom_factor_dev (zoom.cu; deepim/operator_py/zoom_mask.py:36-119, zoom_image.py:30-118)
// reduces the two crop boxes through a caller-owned int32 workspace [B, RC_ZF_WS_INTS].  Here
// that scratch is a file-scope device array of RC_ZF_NOALLOC_MAX_BATCH rows (160 KB, held by
// every GPU that loads the library until the process exits).  The array's NAME is only a
// host-side stand-in in host code, so its device address is taken once per device with
// cudaGetSymbolAddress (that device current) and handed to the same launcher the allocating
// entry point uses.  Nothing here calls cudaMalloc / cudaFree or any allocating routine.
//
// All calls on a device share the one array: it is cleared (stream-ordered memset of the rows
// the call uses) in front of every launch, and two calls that overlap on the GPU (different
// streams, no ordering between them) race on it -- order them, or use rc_zoom_factor_dev with
// one workspace per stream.
#include <mutex>

#include "common.cuh"

namespace rc {

static __device__ __align__(256) int32_t g_zf_scratch[RC_ZF_NOALLOC_MAX_BATCH * RC_ZF_WS_INTS];

constexpr int kMaxDevices = 64;
static int32_t *g_zf_scratch_addr[kMaxDevices];  // device address of g_zf_scratch, per device
static std::mutex g_zf_scratch_mu;

// Device address of the scratch array on the CURRENT device (looked up once per device).
static int zf_scratch_ptr(int32_t **out) {
  int dev = -1;
  RC_CUDA_TRY(cudaGetDevice(&dev));
  RC_REQUIRE(dev >= 0 && dev < kMaxDevices, RC_ERR_INVALID_ARG);
  std::lock_guard<std::mutex> lock(g_zf_scratch_mu);
  if (!g_zf_scratch_addr[dev]) {
    void *p = nullptr;
    RC_CUDA_TRY(cudaGetSymbolAddress(&p, g_zf_scratch));
    RC_REQUIRE(p != nullptr, RC_ERR_CUDA);
    g_zf_scratch_addr[dev] = static_cast<int32_t *>(p);
  }
  *out = g_zf_scratch_addr[dev];
  return RC_OK;
}

}  // namespace rc

extern "C" {

int rc_zoom_factor_noalloc_max_batch(void) { return RC_ZF_NOALLOC_MAX_BATCH; }

int rc_zoom_factor_noalloc_dev(const float *obs, const float *rend, const float *src_pose,
                               const float *K, const float *pixel_means, float *zoom_factor,
                               int32_t *bbox, int batch_size, int channels, int height, int width,
                               int mode, void *stream) {
  RC_REQUIRE(batch_size >= 0, RC_ERR_INVALID_ARG);
  // "batch_size exceeds RC_ZF_NOALLOC_MAX_BATCH": the static scratch cannot grow
  RC_REQUIRE(batch_size <= RC_ZF_NOALLOC_MAX_BATCH, RC_ERR_CAPACITY);
  if (batch_size == 0) return rc::note_status(RC_OK);
  int32_t *ws = nullptr;
  int st = rc::zf_scratch_ptr(&ws);
  if (st != RC_OK) return st;
  // the array keeps whatever the last call (or an aborted kernel) left in it
  RC_CUDA_TRY(cudaMemsetAsync(ws, 0, sizeof(int32_t) * RC_ZF_WS_INTS * (size_t)batch_size,
                              static_cast<cudaStream_t>(stream)));
  return rc::launch_zoom_factor(obs, rend, src_pose, K, pixel_means, zoom_factor, bbox, ws,
                                nullptr, nullptr, batch_size, channels, height, width, mode,
                                stream);
}

}  // extern "C"
