// Internal interface between the C ABI (vc_api_extract.cu) and the extraction post-processing
// kernels (score_kernels.cu, select_kernels.cu, sample_kernels.cu).  These are the bandwidth-bound
// stages of the path: coalesced vector loads, shared-memory staging, warp-shuffle reductions.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace vc {

constexpr int kMaxChannels = 1024;   // ViT-S/B/L: 384 / 768 / 1024
constexpr int kMaxGridCells = 16384;  // 85 x 114 = 9690 at 1600x1200

struct ScoreParams {
  const void* tokens;   // [B, H*W, C] token-major, f32 or bf16
  int is_bf16;
  int64_t image_stride; // elements between images
  int B, H, W, C;
  int method;           // 0 harris, 1 dog, 2 combined
  float* moments;       // workspace [B, 4, H, W]: Ixx, Iyy, Ixy, channel mean
  float* scratch;       // workspace [B, H, W]: second response for "combined"
  float* score;         // out [B, H, W]
  float g3[3], g7[7], g11[11];  // 1-D Gaussian taps (sigma 1, 1, 1.6), normalised
};
cudaError_t launch_score_map(const ScoreParams& p, cudaStream_t stream);

struct SelectParams {
  const float* score;  // [B, H, W]
  int B, H, W;
  int num_keypoints;   // K
  int bin_size;        // 16 in the reference
  float nms_radius;    // 1.5 in the reference
  int cap;             // rows per image in the outputs
  int32_t* coords;     // out [B, cap, 2] (y, x)
  float* scores;       // out [B, cap]
  int32_t* counts;     // out [B]
};
int select_smem_bytes(int H, int W, int K, int bin_size);
cudaError_t launch_select_keypoints(const SelectParams& p, cudaStream_t stream);

struct SampleParams {
  const void* tokens;  // [B, H*W, C]
  int is_bf16;
  int64_t image_stride;
  int B, H, W, C;
  const void* coords;  // [B, cap, 2] (y, x): int32 grid cells or float32 grid units
  int coords_f32;
  const int32_t* counts;  // [B]
  int cap;
  const float* projection;  // [C, d] or nullptr (then d == C)
  const float* proj_bias;   // [d] added after the projection, or nullptr
  int d;
  int quant_mode;      // 0: trunc(clip(512 v, 0, 255)) (vit); 1: round, COLMAP; 2: (v+1)*127.5
  int rootsift;        // 1: L1-normalise + sqrt(clamp 1e-8) before the L2 step (hybrid_extractor.py:286-288)
  float norm_eps;      // F.normalize eps: 1e-12 (vit_extractor.py:243), 1e-8 (vit_feature_model.py:226)
  float sx1, sx2, sy1, sy2;  // keypoint scale factors: x = (col + .5) * sx1 * sx2
  float* keypoints;    // out [B, cap, 2] (x, y) or nullptr
  __nv_bfloat16* desc_bf16;  // out [B, cap, d] or nullptr
  uint8_t* desc_u8;          // out [B, cap, d] or nullptr
  float* desc_f32;           // out [B, cap, d] or nullptr
};
cudaError_t launch_sample_pack(const SampleParams& p, cudaStream_t stream);

struct TrainableSelectParams {
  const float* kp_map;   // [B, H*W, 4] channel-last: score logit, dx, dy, orientation
  int B, H, W;
  int num_keypoints, nms_radius, apply_tanh;
  float score_threshold;
  float scale_x, scale_y;  // w_orig / w_new, h_orig / h_new
  float max_x, max_y;      // w_orig - 1, h_orig - 1
  int cap;
  float* score_scratch;               // [B, H*W]
  unsigned long long* key_scratch;    // [B, key_capacity]
  size_t key_capacity;                // power of two >= H*W
  int32_t* coords;       // out [B, cap, 2] (y, x)
  float* keypoints;      // out [B, cap, 6] x, y, 1, orientation, score, 0
  int32_t* counts;       // out [B]
};
cudaError_t launch_trainable_select(const TrainableSelectParams& p, cudaStream_t stream);

}  // namespace vc
