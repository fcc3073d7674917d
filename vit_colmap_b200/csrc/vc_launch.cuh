// Launch-side helpers shared by the kernel translation units.
#pragma once
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

}  // namespace vc
