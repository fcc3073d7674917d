// Shared by the translation units that implement the C ABI (vc_api*.cu): handle layout, error
// reporting, launch counter.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <string>

#include <cuda.h>
#include <cuda_runtime.h>

#include "../../include/vc_b200.h"

namespace vc {

int fail(int code, const char* fmt, ...);
void add_launches(int n);
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                   const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

}  // namespace vc

#define VC_CUDA(expr)                                                                             \
  do {                                                                                            \
    cudaError_t _e = (expr);                                                                      \
    if (_e != cudaSuccess)                                                                        \
      return vc::fail(VC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, \
                      __LINE__);                                                                  \
  } while (0)

struct vc_handle {
  int device = 0;
  int num_sms = 0;
  vc::EncodeTiledFn encode = nullptr;
  // cached descriptor for the last (ptr, n_images, kmax, d)
  const void* map_ptr = nullptr;
  int map_n = 0, map_k = 0, map_d = 0;
  CUtensorMap map;
  // device staging owned by the host-buffer entry point
  void* stage = nullptr;
  size_t stage_bytes = 0;
  // streams / events / pinned scratch of the host-buffer entry point (created on first use)
  cudaStream_t host_streams[3] = {nullptr, nullptr, nullptr};  // compute, upload, download
  cudaEvent_t host_events[16] = {};                           // [0..7] chunk uploaded, [8..15] group matched
  void* pin = nullptr;
  size_t pin_bytes = 0;
  // row-block counters of vc_vit_linear_ln (the forward keeps its own in the caller's workspace)
  int* ln_counters = nullptr;
  size_t ln_counters_len = 0;
};
