// status.cu -- status codes, error bookkeeping and small device queries.
#include "common.cuh"

namespace rc {
static thread_local cudaError_t g_last_cuda = cudaSuccess;
static thread_local int g_last_status = RC_OK;
void set_cuda_error(cudaError_t e) { g_last_cuda = e; }
int note_status(int status) {
  g_last_status = status;
  return status;
}

// SM count of `device` (< 0: the calling thread's current device), cached PER DEVICE: the
// reference drives several GPUs from one process (batch_updater_py_multi.py:176), so a
// process-wide cache filled by whichever device called first would size grids for the wrong part.
int device_sm_count(int device) {
  constexpr int kMaxDev = 64;
  static int cache[kMaxDev];  // 0 = unknown; benign race (every writer stores the same value)
  if (device < 0 && cudaGetDevice(&device) != cudaSuccess) return 148;
  if (device >= 0 && device < kMaxDev && cache[device] > 0) return cache[device];
  int v = 0;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || v <= 0)
    return 148;
  if (device >= 0 && device < kMaxDev) cache[device] = v;
  return v;
}
}  // namespace rc

extern "C" {

int rc_version(void) { return 1; }

const char *rc_status_string(int status) {
  switch (status) {
    case RC_OK: return "ok";
    case RC_ERR_INVALID_ARG: return "invalid argument";
    case RC_ERR_CUDA: return "CUDA runtime error";
    case RC_ERR_UNALIGNED: return "pointer or width not aligned for 128-bit access";
    case RC_ERR_CAPACITY: return "workspace or table capacity exceeded";
    default: return "unknown status";
  }
}

const char *rc_last_cuda_error(void) { return cudaGetErrorString(rc::g_last_cuda); }

int rc_last_status(void) { return rc::g_last_status; }

int rc_device_sm_count(int device_id, int *sm_count) {
  RC_REQUIRE(sm_count != nullptr, RC_ERR_INVALID_ARG);
  int v = 0;
  RC_CUDA_TRY(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device_id));
  *sm_count = v;
  return rc::note_status(RC_OK);
}

}  // extern "C"
