// flow_cxx.cu -- the reference's `_flow` under its ORIGINAL C++ linkage.
//
// /root/reference/lib/flow_c/gpu_flow.hpp:1-3 declares `void _flow(float*, float*, float*,
// float*, float*, float*, int, int, int, int)` without extern "C", and the Cython stub
// (lib/flow_c/gpu_flow.pyx:12-16, `cdef extern from "gpu_flow.hpp"`) is compiled as C++, so
// the symbol it links against is the mangled `_Z5_flowPfS_S_S_S_S_iiii`.  flow.cu exports
// the C symbol `_flow` (for ctypes / cgo-style bindings); this translation unit -- which must
// NOT see include/rendercompare.h's extern "C" declaration of the same name -- adds the
// mangled one, so the reference's extension links against librendercompare.so with no header
// edit.  Both forward to rc_flow_host.
extern "C" int rc_flow_host(float *flow, float *valid, const float *depth_src,
                            const float *depth_tgt, const float *KT, const float *Kinv,
                            int batch_size, int height, int width, int device_id);

__attribute__((visibility("default"))) void _flow(float *flow, float *valid, float *depth_src,
                                                  float *depth_tgt, float *KT, float *Kinv,
                                                  int batch_size, int height, int width,
                                                  int device_id) {
  (void)rc_flow_host(flow, valid, depth_src, depth_tgt, KT, Kinv, batch_size, height, width,
                     device_id);
}
