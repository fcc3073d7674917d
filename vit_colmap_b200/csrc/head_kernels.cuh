// Internal interface of the ViTFeatureModel convolutional head kernels (conv_gemm.cu,
// head_pointwise.cu).  Reference: vit_colmap/model/vit_feature_model.py:12-29 (UpsampleBlock),
// :90-121 (upsampler, trunk, heads), :173-232 (forward).  All activations are channels-last bf16
// [B, H, W, C]; the NCHW tensors of the reference are never materialised.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace vc {

constexpr int kConvTileH = 8, kConvTileW = 16;  // one GEMM M tile = 8 x 16 output pixels of one image
constexpr int kConvMaxTaps = 9, kConvMaxPhases = 4;

enum ConvEpilogue {
  kConvBiasBf16 = 0,      // out_bf16 = acc + shift                    (transposed convolution)
  kConvBnGeluBf16 = 1,    // out_bf16 = gelu_erf(acc * scale + shift)  (conv + eval BatchNorm + GELU)
  kConvBiasF32 = 2        // out_f32  = acc + shift                    (1x1 descriptor projection)
};

// Implicit-GEMM convolution: for each phase p and output pixel (y, x) of the INPUT grid,
//   out[b, y*sy + oy[p], x*sx + ox[p], n] = epi( sum_t sum_c in[b, y+dy[p][t], x+dx[p][t], c] *
//                                                      w[p*Cout + n, t*Cin + c] )
// with zero padding outside the input.  Conv3x3: 1 phase, 9 taps; ConvTranspose k4 s2 p1: 4 sub-pixel
// phases of 2x2 taps; 1x1: 1 phase, 1 tap.
struct ConvParams {
  int B, H, W;            // input grid
  int Cin, Cout;          // Cin % 64 == 0, Cout % 32 == 0
  int phases, taps;
  int8_t dy[kConvMaxPhases][kConvMaxTaps], dx[kConvMaxPhases][kConvMaxTaps];
  int sy, sx;             // output stride multiplier (1 conv, 2 transposed conv)
  int8_t oy[kConvMaxPhases], ox[kConvMaxPhases];
  int Ho, Wo;             // output grid
  int epilogue;
  const float* scale;     // [Cout] (kConvBnGeluBf16)
  const float* shift;     // [Cout]
  void* out;              // [B, Ho, Wo, ldo]
  int ldo;
  int split_n;            // columns >= split_n go to out2 (column n - split_n); 0 = no split
  void* out2;
  int ldo2;
};
int conv_smem_bytes();
// map_a: 4-D map over in bf16 [B, H, W, Cin] with box {64, 16, 8, 1}; map_w: 2-D map over
// w bf16 [phases*Cout, taps*Cin] with box {64, 256}; both 128-byte swizzle.
cudaError_t launch_conv(const CUtensorMap& map_a, const CUtensorMap& map_w, const ConvParams& p,
                        int num_sms, cudaStream_t stream);

// Two-SM (cta_group::2) form, conv_gemm2.cu: a cluster of two CTAs per pair of patches.  Same results bit for bit.
// map_w must be built with a {64, 128} box (each CTA stages half of a block's weight rows).
cudaError_t launch_conv2(const CUtensorMap& map_a, const CUtensorMap& map_w, const ConvParams& p,
                         int num_sms, cudaStream_t stream);

// Bilinear resize, align_corners=False (F.interpolate, vit_feature_model.py:205-212), bf16
// channels-last [B, Hi, Wi, C] -> [B, Ho, Wo, C]; C % 8 == 0.
cudaError_t launch_resize_bilinear(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int Hi, int Wi,
                                   int Ho, int Wo, int C, cudaStream_t stream);
// Keypoint head 1x1 convolution 64 -> 4 (+bias), tanh(.)*pi on channel 3
// (vit_feature_model.py:112,222-223): hidden bf16 [pixels, 64] -> kp f32 [pixels, 4].
cudaError_t launch_kp_head(const __nv_bfloat16* hidden, const float* w, const float* b, float* kp,
                           long long pixels, cudaStream_t stream);
// F.normalize(p=2, dim=channel, eps=1e-8) in place on f32 [pixels, d] (vit_feature_model.py:226).
cudaError_t launch_normalize_rows(float* desc, long long pixels, int d, cudaStream_t stream);

}  // namespace vc
