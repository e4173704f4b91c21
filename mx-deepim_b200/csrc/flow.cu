// flow.cu -- ground-truth optical flow with visibility z-test.
//
// Replaces /root/reference/lib/flow_c/gpu_flow_kernel.cu:32-69 (flow_kernel)
// and its host wrapper :82-148 (_flow).
//
// B200 design: the pass is a pure HBM stream (8 B/px read, 12 B/px written,
// one L2-resident gather).  One thread owns four horizontally adjacent pixels:
// one 128-bit no-allocate load of depth_src, three 128-bit stores, and up to
// four scalar gathers of depth_tgt through the read-only path (the target
// pixel is spatially coherent with the source pixel, so the gather hits lines
// that neighbouring threads stream anyway).  A block never straddles samples,
// so the 12 KT coefficients are block-uniform and live in registers.
//
// Bit-exactness: every floating-point operation is written with an explicit
// rounding intrinsic in the association nvcc gives the reference source
// (FMUL/FFMA/FADD sequence read from its sm_100a SASS), so the
// result does not depend on this file's optimisation flags.
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "common.cuh"
#include "flow_device.cuh"

namespace rc {

// grid = (ceil(H*W/4 / 256), B); block = 256.
__global__ void __launch_bounds__(256)
flow_kernel_v4(const float *__restrict__ depth_src, const float *__restrict__ depth_tgt,
               const float *__restrict__ KT, const float *__restrict__ Kinv,
               float *__restrict__ flow, float *__restrict__ valid, int H, int W) {
  const int b = blockIdx.y;
  const int quads = (H * W) >> 2;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= quads) return;
  FlowCoef c;
  load_coef(c, KT, Kinv, b);
  const size_t plane = (size_t)H * W;
  const int p = q << 2;
  const int h = p / W;
  const int w = p - h * W;
  const float4 d = ld_stream_f4(reinterpret_cast<const float4 *>(depth_src + b * plane) + q);
  const float *tgt = depth_tgt + b * plane;
  const float fh = (float)h;
  float4 o0, o1, ov;
  flow_pixel(d.x, (float)(w + 0), fh, c, tgt, H, W, o0.x, o1.x, ov.x);
  flow_pixel(d.y, (float)(w + 1), fh, c, tgt, H, W, o0.y, o1.y, ov.y);
  flow_pixel(d.z, (float)(w + 2), fh, c, tgt, H, W, o0.z, o1.z, ov.z);
  flow_pixel(d.w, (float)(w + 3), fh, c, tgt, H, W, o0.w, o1.w, ov.w);
  float *f = flow + (size_t)b * 2 * plane;
  st_f4(reinterpret_cast<float4 *>(f) + q, o0);
  st_f4(reinterpret_cast<float4 *>(f + plane) + q, o1);
  st_f4(reinterpret_cast<float4 *>(valid + b * plane) + q, ov);
  // possibly launched programmatically behind a grid it never reads (rc_compare_step_dev):
  // waiting after the work makes this grid's completion imply the predecessor's, which is what
  // the normal launch behind it (the resampler) relies on
  pdl_wait();
}

// Scalar variant for widths that are not a multiple of four (host entry only).
__global__ void __launch_bounds__(256)
flow_kernel_v1(const float *__restrict__ depth_src, const float *__restrict__ depth_tgt,
               const float *__restrict__ KT, const float *__restrict__ Kinv,
               float *__restrict__ flow, float *__restrict__ valid, int H, int W) {
  const int b = blockIdx.y;
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= H * W) return;
  FlowCoef c;
  load_coef(c, KT, Kinv, b);
  const size_t plane = (size_t)H * W;
  const int h = p / W;
  const int w = p - h * W;
  float f0, f1, v;
  flow_pixel(__ldg(depth_src + b * plane + p), (float)w, (float)h, c, depth_tgt + b * plane,
             H, W, f0, f1, v);
  flow[(size_t)b * 2 * plane + p] = f0;
  flow[(size_t)b * 2 * plane + plane + p] = f1;
  valid[b * plane + p] = v;
}

static int launch_flow(const float *depth_src, const float *depth_tgt, const float *KT,
                       const float *Kinv, float *flow, float *valid, int B, int H, int W,
                       cudaStream_t stream, bool vec) {
  if (B == 0) return RC_OK;
  if (vec) {
    const int quads = (H * W) / 4;
    dim3 grid((quads + 255) / 256, B);
    if (tl_pdl) {
      RC_CUDA_TRY(launch_maybe_pdl(flow_kernel_v4, grid, dim3(256), 0, stream, true, depth_src,
                                   depth_tgt, KT, Kinv, flow, valid, H, W));
      return RC_OK;
    }
    flow_kernel_v4<<<grid, 256, 0, stream>>>(depth_src, depth_tgt, KT, Kinv, flow, valid, H, W);
  } else {
    dim3 grid((H * W + 255) / 256, B);
    flow_kernel_v1<<<grid, 256, 0, stream>>>(depth_src, depth_tgt, KT, Kinv, flow, valid, H, W);
  }
  return check_launch();
}

// Grow-only staging arena for the host-pointer entry point (one per device).  The reference's
// _flow is re-entrant per device (it mallocs and frees per call, gpu_flow_kernel.cu:86-147);
// the arena is shared, so a per-device mutex is held for the whole host-pointer call: concurrent
// callers on one device take turns (they would share the PCIe link anyway), callers on
// different devices do not interact.
constexpr int kHostWorkers = 8;
struct Arena {
  void *ptr = nullptr;
  size_t bytes = 0;
  std::mutex mu;
  // pipelined host path: page-locked staging (one slot per worker) and one stream per worker
  void *pin = nullptr;
  size_t pin_bytes = 0;
  cudaStream_t streams[kHostWorkers] = {};
  bool have_streams = false;
};
static Arena g_arena[64];

static int arena_reserve(int dev, size_t bytes, void **out);

static int pin_reserve(Arena &a, size_t bytes) {
  if (a.pin_bytes < bytes) {
    if (a.pin) RC_CUDA_TRY(cudaFreeHost(a.pin));
    a.pin = nullptr;
    a.pin_bytes = 0;
    RC_CUDA_TRY(cudaHostAlloc(&a.pin, bytes, cudaHostAllocDefault));
    a.pin_bytes = bytes;
  }
  if (!a.have_streams) {
    for (int i = 0; i < kHostWorkers; ++i)
      RC_CUDA_TRY(cudaStreamCreateWithFlags(&a.streams[i], cudaStreamNonBlocking));
    a.have_streams = true;
  }
  return RC_OK;
}

// Host-pointer path, pipelined by pairs.  The caller's buffers are pageable numpy memory: a
// plain cudaMemcpy of them is staged by the driver on ONE thread at a few GB/s (15 ms per
// 16-pair call, against 3.7 ms of PCIe time).  Here `n` worker threads each own a page-locked
// slot, a device slot and a stream, and walk the pairs k, k + n, ...: memcpy in -> H2D ->
// kernel (one pair) -> D2H -> memcpy out, so the host copies of different pairs run in
// parallel and overlap the transfers of the others.  Results are the same kernel's.
static int flow_host_pipelined(Arena &a, float *flow, float *valid, const float *depth_src,
                               const float *depth_tgt, const float *KT, const float *Kinv, int B,
                               int H, int W, int device_id, int n_workers) {
  const size_t plane = (size_t)H * W * sizeof(float);
  const size_t plane_pad = (plane + 255) & ~(size_t)255;
  const size_t slot = 5 * plane_pad;  // src | tgt | flow (2 planes, contiguous) | valid
  const size_t kt_bytes = (size_t)B * 12 * sizeof(float);
  const size_t kt_pad = (kt_bytes + 255) & ~(size_t)255;
  // flow's two planes must be contiguous for the kernel: place them at 2*plane_pad with the
  // second plane right behind the first (plane is a multiple of 16 when W % 4 == 0)
  void *base = nullptr;
  int st = arena_reserve(device_id, slot * n_workers + kt_pad + 256, &base);
  if (st != RC_OK) return st;
  st = pin_reserve(a, slot * n_workers);
  if (st != RC_OK) return st;
  char *dbase = static_cast<char *>(base);
  float *d_kt = reinterpret_cast<float *>(dbase + slot * n_workers);
  float *d_ki = reinterpret_cast<float *>(dbase + slot * n_workers + kt_pad);
  RC_CUDA_TRY(cudaMemcpyAsync(d_kt, KT, kt_bytes, cudaMemcpyHostToDevice, a.streams[0]));
  RC_CUDA_TRY(cudaMemcpyAsync(d_ki, Kinv, 9 * sizeof(float), cudaMemcpyHostToDevice, a.streams[0]));
  RC_CUDA_TRY(cudaStreamSynchronize(a.streams[0]));
  std::atomic<int> status(RC_OK);
  std::atomic<int> cuda_err((int)cudaSuccess);
  const bool vec = (W % 4 == 0);
  const size_t hw = (size_t)H * W;
  auto work = [&](int k) {
    auto fail = [&](cudaError_t e) {
      int expect = RC_OK;
      if (status.compare_exchange_strong(expect, RC_ERR_CUDA)) cuda_err.store((int)e);
    };
    cudaError_t e = cudaSetDevice(device_id);
    if (e != cudaSuccess) return fail(e);
    char *hp = static_cast<char *>(a.pin) + slot * k;
    char *dp = dbase + slot * k;
    cudaStream_t s = a.streams[k];
    float *d_src = reinterpret_cast<float *>(dp), *d_tgt = reinterpret_cast<float *>(dp + plane_pad);
    float *d_flow = reinterpret_cast<float *>(dp + 2 * plane_pad);
    float *d_valid = reinterpret_cast<float *>(dp + 4 * plane_pad);
    for (int b = k; b < B && status.load() == RC_OK; b += n_workers) {
      std::memcpy(hp, depth_src + b * hw, plane);
      std::memcpy(hp + plane_pad, depth_tgt + b * hw, plane);
      if ((e = cudaMemcpyAsync(d_src, hp, plane, cudaMemcpyHostToDevice, s)) != cudaSuccess) return fail(e);
      if ((e = cudaMemcpyAsync(d_tgt, hp + plane_pad, plane, cudaMemcpyHostToDevice, s)) != cudaSuccess)
        return fail(e);
      if (launch_flow(d_src, d_tgt, d_kt + (size_t)b * 12, d_ki, d_flow, d_valid, 1, H, W, s, vec) != RC_OK)
        return fail(cudaGetLastError());
      if ((e = cudaMemcpyAsync(hp + 2 * plane_pad, d_flow, 2 * plane, cudaMemcpyDeviceToHost, s)) != cudaSuccess)
        return fail(e);
      if ((e = cudaMemcpyAsync(hp + 4 * plane_pad, d_valid, plane, cudaMemcpyDeviceToHost, s)) != cudaSuccess)
        return fail(e);
      if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail(e);
      std::memcpy(flow + (size_t)b * 2 * hw, hp + 2 * plane_pad, 2 * plane);
      std::memcpy(valid + b * hw, hp + 4 * plane_pad, plane);
    }
  };
  std::vector<std::thread> pool;
  pool.reserve(n_workers - 1);
  for (int k = 1; k < n_workers; ++k) pool.emplace_back(work, k);
  work(0);
  for (auto &t : pool) t.join();
  if (status.load() != RC_OK) {
    set_cuda_error((cudaError_t)cuda_err.load());
    return RC_ERR_CUDA;
  }
  return RC_OK;
}

static int arena_reserve(int dev, size_t bytes, void **out) {
  Arena &a = g_arena[dev];
  if (a.bytes < bytes) {
    if (a.ptr) RC_CUDA_TRY(cudaFree(a.ptr));
    a.ptr = nullptr;
    a.bytes = 0;
    RC_CUDA_TRY(cudaMalloc(&a.ptr, bytes));
    a.bytes = bytes;
  }
  *out = a.ptr;
  return RC_OK;
}

}  // namespace rc

extern "C" {

int rc_flow_dev(const float *depth_src, const float *depth_tgt, const float *KT,
                const float *Kinv, float *flow, float *valid, int batch_size, int height,
                int width, void *stream) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);  // gpu_flow.pyx:36
  RC_REQUIRE(depth_src && depth_tgt && KT && Kinv && flow && valid, RC_ERR_INVALID_ARG);
  RC_REQUIRE((int64_t)batch_size * height * width * 2 < (int64_t)INT32_MAX, RC_ERR_INVALID_ARG);
  RC_REQUIRE(width % 4 == 0 && rc::aligned16(depth_src) && rc::aligned16(flow) &&
                 rc::aligned16(valid),
             RC_ERR_UNALIGNED);
  return rc::note_status(rc::launch_flow(depth_src, depth_tgt, KT, Kinv, flow, valid,
                                         batch_size, height, width,
                                         static_cast<cudaStream_t>(stream), true));
}

int rc_flow_host(float *flow, float *valid, const float *depth_src, const float *depth_tgt,
                 const float *KT, const float *Kinv, int batch_size, int height, int width,
                 int device_id) {
  RC_REQUIRE(batch_size >= 0 && height > 0 && width > 0, RC_ERR_INVALID_ARG);
  RC_REQUIRE(device_id >= 0 && device_id < 64, RC_ERR_INVALID_ARG);
  if (batch_size == 0) return rc::note_status(RC_OK);
  RC_REQUIRE(depth_src && depth_tgt && KT && Kinv && flow && valid, RC_ERR_INVALID_ARG);
  std::lock_guard<std::mutex> guard(rc::g_arena[device_id].mu);
  int cur = -1;
  RC_CUDA_TRY(cudaGetDevice(&cur));  // gpu_flow_kernel.cu:71-80
  if (cur != device_id) RC_CUDA_TRY(cudaSetDevice(device_id));
  {
    // RC_FLOW_HOST_THREADS: workers of the pipelined path (default 8, 0 or 1 = the plain
    // pageable copies below); a one-pair call has nothing to overlap
    const char *env = std::getenv("RC_FLOW_HOST_THREADS");
    int n = env ? std::atoi(env) : rc::kHostWorkers;
    const int hc = (int)std::thread::hardware_concurrency();
    if (hc > 0 && n > hc) n = hc;
    if (n > rc::kHostWorkers) n = rc::kHostWorkers;
    if (n > batch_size) n = batch_size;
    if (n >= 2)
      return rc::note_status(rc::flow_host_pipelined(rc::g_arena[device_id], flow, valid, depth_src,
                                                     depth_tgt, KT, Kinv, batch_size, height,
                                                     width, device_id, n));
  }
  const size_t plane = (size_t)batch_size * height * width * sizeof(float);
  const size_t kt_bytes = (size_t)batch_size * 12 * sizeof(float);
  const size_t kt_pad = (kt_bytes + 255) & ~(size_t)255;
  // layout: src | tgt | flow(2) | valid | KT | Kinv ; planes are 16-byte multiples
  // only when H*W % 4 == 0, so pad each region to 256 bytes.
  const size_t plane_pad = (plane + 255) & ~(size_t)255;
  const size_t total = plane_pad * 5 + kt_pad + 256;
  void *base = nullptr;
  int st = rc::arena_reserve(device_id, total, &base);
  if (st != RC_OK) return st;
  char *p = static_cast<char *>(base);
  float *d_src = reinterpret_cast<float *>(p);
  float *d_tgt = reinterpret_cast<float *>(p + plane_pad);
  float *d_flow = reinterpret_cast<float *>(p + 2 * plane_pad);  // 2 planes contiguous
  float *d_valid = reinterpret_cast<float *>(p + 4 * plane_pad);
  float *d_kt = reinterpret_cast<float *>(p + 5 * plane_pad);
  float *d_ki = reinterpret_cast<float *>(p + 5 * plane_pad + kt_pad);
  cudaStream_t s = cudaStreamPerThread;
  RC_CUDA_TRY(cudaMemcpyAsync(d_src, depth_src, plane, cudaMemcpyHostToDevice, s));
  RC_CUDA_TRY(cudaMemcpyAsync(d_tgt, depth_tgt, plane, cudaMemcpyHostToDevice, s));
  RC_CUDA_TRY(cudaMemcpyAsync(d_kt, KT, kt_bytes, cudaMemcpyHostToDevice, s));
  RC_CUDA_TRY(cudaMemcpyAsync(d_ki, Kinv, 9 * sizeof(float), cudaMemcpyHostToDevice, s));
  st = rc::launch_flow(d_src, d_tgt, d_kt, d_ki, d_flow, d_valid, batch_size, height, width, s,
                       width % 4 == 0);
  if (st != RC_OK) return rc::note_status(st);
  RC_CUDA_TRY(cudaMemcpyAsync(flow, d_flow, 2 * plane, cudaMemcpyDeviceToHost, s));
  RC_CUDA_TRY(cudaMemcpyAsync(valid, d_valid, plane, cudaMemcpyDeviceToHost, s));
  RC_CUDA_TRY(cudaStreamSynchronize(s));
  return rc::note_status(RC_OK);
}

int rc_flow_host_release(int device_id) {
  RC_REQUIRE(device_id >= 0 && device_id < 64, RC_ERR_INVALID_ARG);
  rc::Arena &a = rc::g_arena[device_id];
  std::lock_guard<std::mutex> guard(a.mu);
  if (!a.ptr && !a.pin && !a.have_streams) return rc::note_status(RC_OK);
  int cur = -1;
  RC_CUDA_TRY(cudaGetDevice(&cur));
  if (cur != device_id) RC_CUDA_TRY(cudaSetDevice(device_id));
  if (a.have_streams) {
    for (int i = 0; i < rc::kHostWorkers; ++i) (void)cudaStreamDestroy(a.streams[i]);
    a.have_streams = false;
  }
  if (a.ptr) (void)cudaFree(a.ptr);
  if (a.pin) (void)cudaFreeHost(a.pin);
  a.ptr = a.pin = nullptr;
  a.bytes = a.pin_bytes = 0;
  if (cur != device_id) RC_CUDA_TRY(cudaSetDevice(cur));
  return rc::note_status(RC_OK);
}

void _flow(float *flow, float *valid, float *depth_src, float *depth_tgt, float *KT,
           float *Kinv, int batch_size, int height, int width, int device_id) {
  (void)rc_flow_host(flow, valid, depth_src, depth_tgt, KT, Kinv, batch_size, height, width,
                     device_id);
}

}  // extern "C"
