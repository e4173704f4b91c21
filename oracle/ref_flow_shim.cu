// oracle/ref_flow_shim.cu -- TEST / BENCH INFRASTRUCTURE ONLY (never on the product path).
//
// Compiles the reference's own CUDA source WHERE IT LIES (nothing is copied into the repo):
// the #include below pulls /root/reference/lib/flow_c/gpu_flow_kernel.cu into this translation
// unit (oracle/build.py passes -I /root/reference/lib/flow_c), so the library holds the
// UNMODIFIED `flow_kernel` (:32-69) and `_flow` (:82-148), plus ONE extra entry point that
// launches that same kernel on DEVICE pointers with the reference's own launch configuration
// (:129-130: CAFFE_GET_BLOCKS(count) blocks of 512 threads) -- the kernel-only baseline
// "G0" of BASELINE.md section 3, timed with CUDA events next to flow_kernel_v4 without the
// reference wrapper's six cudaMalloc/cudaFree and pageable PCIe copies.
#include "gpu_flow_kernel.cu"

extern "C" int ref_flow_kernel_dev(const float *depth_src, const float *depth_tgt, float *KT,
                                   float *Kinv, float *flow, float *valid, int batch_size,
                                   int height, int width, void *stream) {
  const int count = batch_size * height * width;
  flow_kernel<<<CAFFE_GET_BLOCKS(count), CAFFE_CUDA_NUM_THREADS, 0,
                static_cast<cudaStream_t>(stream)>>>(count, depth_src, depth_tgt, height, width,
                                                     KT, Kinv, flow, valid);
  return (int)cudaPeekAtLastError();
}
