// C ABI, extraction post-processing half (see include/vc_b200.h): argument checks and launches
// for the score map, keypoint selection and descriptor sampling / packing kernels.
#include <cmath>

#include "extract_kernels.cuh"
#include "vc_internal.cuh"

using vc::fail;

namespace {

// 1-D taps of vit_extractor.py:396-402: exp(-x^2 / (2 sigma^2)) normalised to sum 1, in fp32
void gaussian_taps(int k, float sigma, float* out) {
  float sum = 0.f;
  for (int i = 0; i < k; ++i) {
    const float x = (float)(i - k / 2);
    out[i] = expf(-(x * x) / (2.0f * sigma * sigma));
    sum += out[i];
  }
  for (int i = 0; i < k; ++i) out[i] /= sum;
}

int check_grid(vc_handle* h, int B, int H, int W) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (B <= 0 || H <= 0 || W <= 0) return fail(VC_ERR_INVALID, "bad sizes B=%d H=%d W=%d", B, H, W);
  if ((long long)H * W > vc::kMaxGridCells)
    return fail(VC_ERR_INVALID, "patch grid %dx%d exceeds %d cells", H, W, vc::kMaxGridCells);
  return VC_OK;
}

int check_tokens(const void* tokens, int dtype, int C) {
  if (!tokens) return fail(VC_ERR_INVALID, "null tokens");
  if (dtype != VC_DTYPE_F32 && dtype != VC_DTYPE_BF16)
    return fail(VC_ERR_INVALID, "tokens must be f32 or bf16");
  if (C <= 0 || C % 4 != 0 || C > vc::kMaxChannels)
    return fail(VC_ERR_INVALID, "C=%d must be a multiple of 4 and <= %d", C, vc::kMaxChannels);
  if ((reinterpret_cast<uintptr_t>(tokens) & 15) != 0)
    return fail(VC_ERR_INVALID, "tokens must be 16-byte aligned");
  return VC_OK;
}

}  // namespace

extern "C" {

size_t vc_score_workspace_bytes(int n_images, int grid_h, int grid_w) {
  if (n_images <= 0 || grid_h <= 0 || grid_w <= 0) return 256;
  return vc::align_up((size_t)n_images * 5 * grid_h * grid_w * sizeof(float), 256);
}

int vc_score_map(vc_handle* h, const void* tokens, int dtype, int n_images, int grid_h, int grid_w,
                 int channels, int64_t image_stride, int method, float* score, void* workspace,
                 size_t workspace_bytes, void* stream) {
  int rc = check_grid(h, n_images, grid_h, grid_w);
  if (rc != VC_OK) return rc;
  rc = check_tokens(tokens, dtype, channels);
  if (rc != VC_OK) return rc;
  if (!score || !workspace) return fail(VC_ERR_INVALID, "null output / workspace");
  if (method < 0 || method > 2) return fail(VC_ERR_INVALID, "unknown detection method %d", method);
  if (workspace_bytes < vc_score_workspace_bytes(n_images, grid_h, grid_w))
    return fail(VC_ERR_WORKSPACE, "workspace %zu B < required %zu B", workspace_bytes,
                vc_score_workspace_bytes(n_images, grid_h, grid_w));
  const int el = dtype == VC_DTYPE_F32 ? 4 : 2;
  if (image_stride < (int64_t)grid_h * grid_w * channels || (image_stride * el) % 16 != 0)
    return fail(VC_ERR_INVALID, "bad image_stride");
  VC_CUDA(cudaSetDevice(h->device));
  vc::ScoreParams p{};
  p.tokens = tokens;
  p.is_bf16 = dtype == VC_DTYPE_BF16;
  p.image_stride = image_stride;
  p.B = n_images;
  p.H = grid_h;
  p.W = grid_w;
  p.C = channels;
  p.method = method;
  p.moments = static_cast<float*>(workspace);
  p.scratch = p.moments + (size_t)n_images * 4 * grid_h * grid_w;
  p.score = score;
  gaussian_taps(3, 1.0f, p.g3);
  gaussian_taps(7, 1.0f, p.g7);
  gaussian_taps(11, 1.6f, p.g11);
  VC_CUDA(vc::launch_score_map(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(2);
  return VC_OK;
}

int vc_select_keypoints(vc_handle* h, const float* score, int n_images, int grid_h, int grid_w,
                        int num_keypoints, int bin_size, float nms_radius, int32_t* coords,
                        float* scores, int32_t* counts, int capacity, void* stream) {
  int rc = check_grid(h, n_images, grid_h, grid_w);
  if (rc != VC_OK) return rc;
  if (!score || !coords || !scores || !counts) return fail(VC_ERR_INVALID, "null pointer argument");
  if (num_keypoints <= 0 || bin_size <= 0 || !(nms_radius >= 0.f) || nms_radius > 8.f)
    return fail(VC_ERR_INVALID, "bad selection parameters");
  const int most = std::min(num_keypoints, grid_h * grid_w);
  if (capacity < most)
    return fail(VC_ERR_INVALID, "capacity %d < min(num_keypoints, grid cells) = %d", capacity, most);
  if (vc::select_smem_bytes(grid_h, grid_w, num_keypoints, bin_size) > 227 * 1024)
    return fail(VC_ERR_INVALID, "selection state does not fit in shared memory");
  VC_CUDA(cudaSetDevice(h->device));
  vc::SelectParams p{};
  p.score = score;
  p.B = n_images;
  p.H = grid_h;
  p.W = grid_w;
  p.num_keypoints = num_keypoints;
  p.bin_size = bin_size;
  p.nms_radius = nms_radius;
  p.cap = capacity;
  p.coords = coords;
  p.scores = scores;
  p.counts = counts;
  VC_CUDA(vc::launch_select_keypoints(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_sample_pack(vc_handle* h, const void* tokens, int dtype, int n_images, int grid_h, int grid_w,
                   int channels, int64_t image_stride, const void* coords, int coords_are_float,
                   const int32_t* counts, int capacity, const float* projection,
                   const float* projection_bias, int d, int quant_mode, const float* kp_scale, float* keypoints, void* desc_bf16, uint8_t* desc_u8,
                   float* desc_f32, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (n_images <= 0 || grid_h <= 0 || grid_w <= 0) return fail(VC_ERR_INVALID, "bad sizes");
  int rc = check_tokens(tokens, dtype, channels);
  if (rc != VC_OK) return rc;
  if (!coords || !counts) return fail(VC_ERR_INVALID, "null coords / counts");
  if (capacity <= 0) return fail(VC_ERR_INVALID, "bad capacity");
  if (projection == nullptr && d != channels)
    return fail(VC_ERR_INVALID, "d=%d must equal C=%d when no projection is given", d, channels);
  if (d <= 0 || d % 4 != 0 || d > vc::kMaxChannels) return fail(VC_ERR_INVALID, "bad d=%d", d);
  const int rootsift = (quant_mode & VC_NORM_ROOTSIFT) != 0;
  quant_mode &= ~VC_NORM_ROOTSIFT;
  if (quant_mode < 0 || quant_mode > 2) return fail(VC_ERR_INVALID, "bad quant_mode");
  if (keypoints && !kp_scale) return fail(VC_ERR_INVALID, "keypoints requested without kp_scale");
  const int el = dtype == VC_DTYPE_F32 ? 4 : 2;
  if (image_stride < (int64_t)grid_h * grid_w * channels || (image_stride * el) % 16 != 0)
    return fail(VC_ERR_INVALID, "bad image_stride");
  if (projection && (size_t)(channels * 16 + 16 * d) * 4 > 227 * 1024)
    return fail(VC_ERR_INVALID, "projection tile does not fit in shared memory");
  VC_CUDA(cudaSetDevice(h->device));
  vc::SampleParams p{};
  p.tokens = tokens;
  p.is_bf16 = dtype == VC_DTYPE_BF16;
  p.image_stride = image_stride;
  p.B = n_images;
  p.H = grid_h;
  p.W = grid_w;
  p.C = channels;
  p.coords = coords;
  p.coords_f32 = coords_are_float;
  p.counts = counts;
  p.cap = capacity;
  p.projection = projection;
  p.proj_bias = projection ? projection_bias : nullptr;
  p.d = d;
  p.quant_mode = quant_mode;
  p.rootsift = rootsift;
  p.norm_eps = quant_mode == VC_QUANT_TRAINABLE ? 1e-8f : 1e-12f;
  if (kp_scale) {
    p.sx1 = kp_scale[0];
    p.sx2 = kp_scale[1];
    p.sy1 = kp_scale[2];
    p.sy2 = kp_scale[3];
  }
  p.keypoints = keypoints;
  p.desc_bf16 = static_cast<__nv_bfloat16*>(desc_bf16);
  p.desc_u8 = desc_u8;
  p.desc_f32 = desc_f32;
  VC_CUDA(vc::launch_sample_pack(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

size_t vc_trainable_workspace_bytes(int n_images, int map_h, int map_w) {
  if (n_images <= 0 || map_h <= 0 || map_w <= 0) return 256;
  size_t cap = 2;
  while (cap < (size_t)map_h * map_w) cap <<= 1;
  return vc::align_up((size_t)n_images * map_h * map_w * sizeof(float), 256) + (size_t)n_images * cap * 8;
}

int vc_trainable_select(vc_handle* h, const float* kp_map, int n_images, int map_h, int map_w,
                        int num_keypoints, float score_threshold, int nms_radius, int apply_tanh,
                        const float* kp_scale, int32_t* coords, float* keypoints, int32_t* counts,
                        int capacity, void* workspace, size_t workspace_bytes, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!kp_map || !kp_scale || !coords || !keypoints || !counts || !workspace)
    return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_images <= 0 || map_h <= 0 || map_w <= 0 || num_keypoints <= 0 || nms_radius < 0 || nms_radius > 16)
    return fail(VC_ERR_INVALID, "bad selection parameters");
  if ((long long)map_h * map_w > (1ll << 30)) return fail(VC_ERR_INVALID, "score map too large");
  if (capacity < std::min(num_keypoints, map_h * map_w))
    return fail(VC_ERR_INVALID, "capacity %d < min(num_keypoints, cells)", capacity);
  if (workspace_bytes < vc_trainable_workspace_bytes(n_images, map_h, map_w))
    return fail(VC_ERR_WORKSPACE, "workspace %zu B < required %zu B", workspace_bytes,
                vc_trainable_workspace_bytes(n_images, map_h, map_w));
  if ((reinterpret_cast<uintptr_t>(kp_map) & 15) != 0) return fail(VC_ERR_INVALID, "kp_map must be 16-byte aligned");
  VC_CUDA(cudaSetDevice(h->device));
  vc::TrainableSelectParams p{};
  p.kp_map = kp_map;
  p.B = n_images;
  p.H = map_h;
  p.W = map_w;
  p.num_keypoints = num_keypoints;
  p.nms_radius = nms_radius;
  p.apply_tanh = apply_tanh;
  p.score_threshold = score_threshold;
  p.scale_x = kp_scale[0];
  p.scale_y = kp_scale[1];
  p.max_x = kp_scale[2];
  p.max_y = kp_scale[3];
  p.cap = capacity;
  size_t cap = 2;
  while (cap < (size_t)map_h * map_w) cap <<= 1;
  p.key_capacity = cap;
  p.score_scratch = static_cast<float*>(workspace);
  p.key_scratch = reinterpret_cast<unsigned long long*>(
      static_cast<uint8_t*>(workspace) + vc::align_up((size_t)n_images * map_h * map_w * sizeof(float), 256));
  p.coords = coords;
  p.keypoints = keypoints;
  p.counts = counts;
  VC_CUDA(vc::launch_trainable_select(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

}  // extern "C"
