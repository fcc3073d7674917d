// C ABI: the allocation-free variant of the score-map entry point (see include/vc_b200.h,
// vc_score_map_static).  It has no counterpart in the reference; it exists for callers that may not
// call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// The score map needs 5 floats of scratch per patch-grid cell (the four moment maps and the second
// response of the "combined" method).  vc_score_map takes that scratch from the caller; here it is a
// file-scope __device__ array, so nothing is allocated by the call.  Every GPU that loads this
// library holds the array (kStaticScratchBytes) until the process exits.
//
// The array's name is only a host-side stand-in in host code: the device address is looked up with
// cudaGetSymbolAddress, once per device, with that device current, and that pointer is what the
// kernels receive.
#include <mutex>

#include "extract_kernels.cuh"
#include "vc_internal.cuh"
#include "vc_launch.cuh"

using vc::fail;

namespace {

constexpr size_t kStaticScratchFloats = VC_SCORE_STATIC_MAX_FLOATS;
constexpr size_t kStaticScratchBytes = kStaticScratchFloats * sizeof(float);

static __device__ __align__(256) float g_score_scratch[kStaticScratchFloats];

// Device address of g_score_scratch on the current device (looked up once per device).
int static_scratch(int device, float** out) {
  static std::mutex mu;
  static float* addr[vc::kMaxDevices] = {};
  if (device < 0 || device >= vc::kMaxDevices)
    return fail(VC_ERR_INVALID, "device %d outside the supported range [0, %d)", device, vc::kMaxDevices);
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    int cur = -1;
    VC_CUDA(cudaGetDevice(&cur));
    if (cur != device) return fail(VC_ERR_CUDA, "device %d is not current (current is %d)", device, cur);
    float* p = nullptr;
    VC_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_score_scratch));
    if (!p || (reinterpret_cast<uintptr_t>(p) & 255) != 0)
      return fail(VC_ERR_CUDA, "static scratch address %p is null or not 256-byte aligned", (void*)p);
    addr[device] = p;
  }
  *out = addr[device];
  return VC_OK;
}

void taps(int k, float sigma, float* out) {  // vit_extractor.py:396-402, as in vc_score_map
  float sum = 0.f;
  for (int i = 0; i < k; ++i) {
    const float x = (float)(i - k / 2);
    out[i] = expf(-(x * x) / (2.0f * sigma * sigma));
    sum += out[i];
  }
  for (int i = 0; i < k; ++i) out[i] /= sum;
}

}  // namespace

extern "C" {

int vc_score_map_static(vc_handle* h, const void* tokens, int dtype, int n_images, int grid_h, int grid_w,
                        int channels, int64_t image_stride, int method, float* score, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (n_images <= 0 || grid_h <= 0 || grid_w <= 0)
    return fail(VC_ERR_INVALID, "bad sizes B=%d H=%d W=%d", n_images, grid_h, grid_w);
  if ((long long)grid_h * grid_w > vc::kMaxGridCells)
    return fail(VC_ERR_INVALID, "patch grid %dx%d exceeds %d cells", grid_h, grid_w, vc::kMaxGridCells);
  if (!tokens || !score) return fail(VC_ERR_INVALID, "null tokens / output");
  if (dtype != VC_DTYPE_F32 && dtype != VC_DTYPE_BF16) return fail(VC_ERR_INVALID, "tokens must be f32 or bf16");
  if (channels <= 0 || channels % 4 != 0 || channels > vc::kMaxChannels)
    return fail(VC_ERR_INVALID, "C=%d must be a multiple of 4 and <= %d", channels, vc::kMaxChannels);
  if ((reinterpret_cast<uintptr_t>(tokens) & 15) != 0) return fail(VC_ERR_INVALID, "tokens must be 16-byte aligned");
  if (method < 0 || method > 2) return fail(VC_ERR_INVALID, "unknown detection method %d", method);
  const int el = dtype == VC_DTYPE_F32 ? 4 : 2;
  if (image_stride < (int64_t)grid_h * grid_w * channels || (image_stride * el) % 16 != 0)
    return fail(VC_ERR_INVALID, "bad image_stride");
  const size_t need = (size_t)n_images * 5 * grid_h * grid_w;
  if (need > kStaticScratchFloats)
    return fail(VC_ERR_WORKSPACE,
                "vc_score_map_static: %d images of %dx%d cells need %zu B of scratch, the static array holds %zu B "
                "(n_images * grid_h * grid_w <= %zu); use vc_score_map with a caller workspace",
                n_images, grid_h, grid_w, need * sizeof(float), kStaticScratchBytes, kStaticScratchFloats / 5);
  VC_CUDA(cudaSetDevice(h->device));
  float* scratch = nullptr;
  if (int rc = static_scratch(h->device, &scratch); rc != VC_OK) return rc;
  vc::ScoreParams p{};
  p.tokens = tokens;
  p.is_bf16 = dtype == VC_DTYPE_BF16;
  p.image_stride = image_stride;
  p.B = n_images;
  p.H = grid_h;
  p.W = grid_w;
  p.C = channels;
  p.method = method;
  // Both regions are written in full before they are read (every cell's four moments by
  // moments_kernel, the second response by response_kernel), so nothing has to start at zero and the
  // previous call's contents are never seen.
  p.moments = scratch;
  p.scratch = scratch + (size_t)n_images * 4 * grid_h * grid_w;
  p.score = score;
  taps(3, 1.0f, p.g3);
  taps(7, 1.0f, p.g7);
  taps(11, 1.6f, p.g11);
  VC_CUDA(vc::launch_score_map(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(2);
  return VC_OK;
}

}  // extern "C"
