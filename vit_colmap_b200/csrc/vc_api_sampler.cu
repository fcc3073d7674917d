// C ABI, training-time similarity / hard-negative mining (see include/vc_b200.h):
// argument checks and launches for sampler_kernels.cu.  fp32 token-major feature maps only.
#include "sampler_kernels.cuh"
#include "vc_internal.cuh"

using vc::fail;

namespace {

constexpr int kMaxChannels = 1024;
constexpr int kMaxCells = 16384;

int check_map(vc_handle* h, const void* tokens, int B, int H, int W, int C) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!tokens) return fail(VC_ERR_INVALID, "null feature map");
  if (B <= 0 || H <= 0 || W <= 0) return fail(VC_ERR_INVALID, "bad sizes B=%d H=%d W=%d", B, H, W);
  if (C <= 0 || C % 4 != 0 || C > kMaxChannels)
    return fail(VC_ERR_INVALID, "C=%d must be a multiple of 4 and <= %d", C, kMaxChannels);
  if ((reinterpret_cast<uintptr_t>(tokens) & 15) != 0) return fail(VC_ERR_INVALID, "feature map must be 16-byte aligned");
  return VC_OK;
}

}  // namespace

extern "C" {

int vc_nchw_to_tokens(vc_handle* h, const float* nchw, int n_images, int channels, int grid_h, int grid_w,
                      float* tokens, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!nchw || !tokens) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_images <= 0 || channels <= 0 || grid_h <= 0 || grid_w <= 0 || n_images > 65535)
    return fail(VC_ERR_INVALID, "bad sizes");
  VC_CUDA(cudaSetDevice(h->device));
  VC_CUDA(vc::launch_nchw_to_tokens(nchw, tokens, n_images, channels, grid_h * grid_w,
                                    reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_warp_similarity(vc_handle* h, const float* tokens1, const float* tokens2, int n_pairs, int grid_h,
                       int grid_w, int channels, const float* h_inv, float patch_size, float img_h, float img_w,
                       float* similarity, void* stream) {
  int rc = check_map(h, tokens1, n_pairs, grid_h, grid_w, channels);
  if (rc != VC_OK) return rc;
  rc = check_map(h, tokens2, n_pairs, grid_h, grid_w, channels);
  if (rc != VC_OK) return rc;
  if (!h_inv || !similarity) return fail(VC_ERR_INVALID, "null pointer argument");
  if (!(patch_size > 0.f) || !(img_h > 0.f) || !(img_w > 0.f)) return fail(VC_ERR_INVALID, "bad image geometry");
  VC_CUDA(cudaSetDevice(h->device));
  vc::WarpSimParams p{};
  p.tokens1 = tokens1;
  p.tokens2 = tokens2;
  p.image_stride = (int64_t)grid_h * grid_w * channels;
  p.B = n_pairs;
  p.H = grid_h;
  p.W = grid_w;
  p.C = channels;
  p.h_inv = h_inv;
  p.patch_size = patch_size;
  p.img_h = img_h;
  p.img_w = img_w;
  p.similarity = similarity;
  VC_CUDA(vc::launch_warp_similarity(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_topk_rows(vc_handle* h, const float* values, int n_rows, int n, int k, float threshold, int32_t* order,
                 int32_t* above, int32_t* n_above, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!values) return fail(VC_ERR_INVALID, "null values");
  if (n_rows <= 0 || n <= 0 || k < 0) return fail(VC_ERR_INVALID, "bad sizes rows=%d n=%d k=%d", n_rows, n, k);
  if (n > kMaxCells) return fail(VC_ERR_INVALID, "n=%d exceeds %d entries per row", n, kMaxCells);
  if (k > 0 && !order) return fail(VC_ERR_INVALID, "k > 0 without an order buffer");
  VC_CUDA(cudaSetDevice(h->device));
  vc::TopkParams p{};
  p.values = values;
  p.n_rows = n_rows;
  p.n = n;
  p.k = k;
  p.threshold = threshold;
  p.order = k > 0 ? order : nullptr;
  p.above = above;
  p.n_above = n_above;
  VC_CUDA(vc::launch_topk_rows(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_hard_negatives(vc_handle* h, const float* tokens, const float* pos_feat, const float* pos_xy, int n_images,
                      int grid_h, int grid_w, int channels, int n_pos, int n_hard, float radius, int32_t* idx,
                      int32_t* counts, void* stream) {
  int rc = check_map(h, tokens, n_images, grid_h, grid_w, channels);
  if (rc != VC_OK) return rc;
  if (!pos_feat || !pos_xy || !idx || !counts) return fail(VC_ERR_INVALID, "null pointer argument");
  if ((reinterpret_cast<uintptr_t>(pos_feat) & 15) != 0 || (reinterpret_cast<uintptr_t>(pos_xy) & 7) != 0)
    return fail(VC_ERR_INVALID, "pos_feat must be 16-byte and pos_xy 8-byte aligned");
  if (n_pos <= 0 || n_hard <= 0 || n_hard > 32) return fail(VC_ERR_INVALID, "bad n_pos=%d / n_hard=%d (1..32)", n_pos, n_hard);
  VC_CUDA(cudaSetDevice(h->device));
  vc::HardNegParams p{};
  p.tokens = tokens;
  p.image_stride = (int64_t)grid_h * grid_w * channels;
  p.B = n_images;
  p.H = grid_h;
  p.W = grid_w;
  p.C = channels;
  p.pos_feat = pos_feat;
  p.pos_xy = pos_xy;
  p.K = n_pos;
  p.n_hard = n_hard;
  p.radius = radius;
  p.idx = idx;
  p.counts = counts;
  VC_CUDA(vc::launch_hard_negatives(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_sample_bilinear(vc_handle* h, const float* tokens, int n_images, int grid_h, int grid_w, int channels,
                       const float* xy, int n_points, float* out, void* stream) {
  int rc = check_map(h, tokens, n_images, grid_h, grid_w, channels);
  if (rc != VC_OK) return rc;
  if (!xy || !out) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_points <= 0) return fail(VC_ERR_INVALID, "bad n_points=%d", n_points);
  if ((reinterpret_cast<uintptr_t>(xy) & 7) != 0 || (reinterpret_cast<uintptr_t>(out) & 15) != 0)
    return fail(VC_ERR_INVALID, "xy must be 8-byte and out 16-byte aligned");
  VC_CUDA(cudaSetDevice(h->device));
  vc::SampleBilinearParams p{};
  p.tokens = tokens;
  p.image_stride = (int64_t)grid_h * grid_w * channels;
  p.B = n_images;
  p.H = grid_h;
  p.W = grid_w;
  p.C = channels;
  p.xy = xy;
  p.n_pts = n_points;
  p.out = out;
  VC_CUDA(vc::launch_sample_bilinear(p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

}  // extern "C"
