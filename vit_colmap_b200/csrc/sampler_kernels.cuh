// Internal interface between the C ABI (vc_api_sampler.cu) and the training-time similarity /
// hard-negative kernels (sampler_kernels.cu).  fp32 token-major feature maps [B, H*W, C].
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

namespace vc {

struct WarpSimParams {
  const float* tokens1;  // source map (image 1), warped into image 2's frame
  const float* tokens2;  // target map (image 2)
  int64_t image_stride;  // elements between images
  int B, H, W, C;
  const float* h_inv;    // device [B, 9]: inverse homographies (target pixel -> source pixel)
  float patch_size, img_h, img_w;
  float* similarity;     // out [B, H, W]
};
cudaError_t launch_warp_similarity(const WarpSimParams& p, cudaStream_t stream);

struct SampleBilinearParams {
  const float* tokens;
  int64_t image_stride;
  int B, H, W, C;
  const float* xy;  // [B, n_pts, 2] (x, y) in feature-map units; NaN -> zero row
  int n_pts;
  float* out;       // [B, n_pts, C]
};
cudaError_t launch_sample_bilinear(const SampleBilinearParams& p, cudaStream_t stream);

struct TopkParams {
  const float* values;  // [n_rows, n]
  int n_rows, n, k;
  float threshold;
  int32_t* order;    // out [n_rows, k]: indices by value descending (lower index first on ties), -1 beyond n; or null
  int32_t* above;    // out [n_rows, n]: ascending indices with value >= threshold (first n_above valid); or null
  int32_t* n_above;  // out [n_rows]; or null
};
int topk_smem_bytes(int n);
cudaError_t launch_topk_rows(const TopkParams& p, cudaStream_t stream);

struct HardNegParams {
  const float* tokens;
  int64_t image_stride;
  int B, H, W, C;
  const float* pos_feat;  // [B, K, C]
  const float* pos_xy;    // [B, K, 2] (x, y) feature-map units
  int K, n_hard;          // n_hard <= 32
  float radius;
  int32_t* idx;           // out [B, K, n_hard] cell indices by similarity descending, -1 padded
  int32_t* counts;        // out [B, K]
};
cudaError_t launch_hard_negatives(const HardNegParams& p, cudaStream_t stream);

cudaError_t launch_nchw_to_tokens(const float* in, float* out, int B, int C, int HW, cudaStream_t stream);

}  // namespace vc
