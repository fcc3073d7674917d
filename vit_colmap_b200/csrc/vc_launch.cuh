// Launch-side helpers shared by the kernel translation units.
#pragma once
#include <cstdlib>

#include <cuda_runtime.h>

namespace vc {

// Opt-in to more than 48 KB of dynamic shared memory.  The attribute belongs to the (kernel, device)
// pair, not to the process, so the "already done" state is kept per device: a second vc_handle on
// another GPU of the same process gets its own opt-in (tests/test_gpu_matching.py two-device test).
constexpr int kMaxDevices = 64;
struct SmemOptIn {
  int bytes[kMaxDevices] = {};
};
template <typename Kernel>
inline cudaError_t ensure_dynamic_smem(Kernel kernel, int bytes, SmemOptIn& state) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev >= 0 && dev < kMaxDevices && state.bytes[dev] >= bytes) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
  if (e == cudaSuccess && dev >= 0 && dev < kMaxDevices) state.bytes[dev] = bytes;
  return e;
}

// Launch with programmatic stream serialization (see griddep_wait in vc_ptx.cuh): the kernel's
// prologue may overlap the tail of the previous kernel of `stream`.  Only for kernels that call
// griddep_wait() before their first access to memory another kernel writes or reads.
// VC_PDL=0 launches them as ordinary stream-ordered kernels (A/B timing).
inline bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("VC_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

}  // namespace vc
