// Internal interface of the DINOv2 ViT forward kernels (vit_gemm.cu, vit_attention.cu,
// vit_pointwise.cu).  The backbone is external to the reference (torch.hub facebookresearch/dinov2,
// called at vit_colmap/features/vit_extractor.py:95-98,135-146); SURVEY.md Appendix A restates it.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace vc {

enum GemmEpilogue {
  kEpiBiasBf16 = 0,      // out_bf16 = acc + bias                       (QKV)
  kEpiBiasGeluBf16 = 1,  // out_bf16 = gelu_erf(acc + bias)             (fc1)
  kEpiResidualF32 = 2,   // out_f32 += gamma * (acc + bias)             (proj / fc2 + LayerScale + residual)
  kEpiPatchEmbed = 3,    // out_f32[b*T + 1 + t] = acc + bias + pos[1+t] (patch embedding)
  // kEpiResidualF32 followed, in the same kernel, by the LayerNorm of the rows it has completed: the CTA
  // that stores the LAST column tile of a 128-row block (per-block counter) normalises those rows straight
  // out of L2 and writes the bf16 operand of the next GEMM.  N must be the full row (N == ldo <= 1024).
  kEpiResidualLnF32 = 4,
  // kEpiResidualF32 whose eight epilogue warps each bump the counter of a tile's 128-row block once their part of
  // the tile is visible device-wide (one tile late, when the fence costs nothing), and lets the next kernel of the stream start early (programmatic dependent launch): that
  // kernel is launch_layernorm_follow, which normalises row blocks WHILE this GEMM is still running.
  kEpiResidualPublishF32 = 5,
  // LayerNorm folded into the GEMMs on either side of it (VC_VIT_LN=fold).  The residual GEMM (kEpiResidualStatsF32)
  // also stores a bf16 copy of the updated rows and, per row and 128-column slice, the sums of x and x^2.  The next
  // GEMM takes that bf16 copy as its A operand with weights W' = W * ln_w (folded on the host) and finishes the
  // normalisation per row in its epilogue: out = rstd * (acc - mean * colsum[n]) + bias'[n]
  // (colsum[n] = sum_k W'[n,k], bias' = bias + W ln_b; kEpiFoldBiasBf16 / kEpiFoldBiasGeluBf16).
  kEpiFoldBiasBf16 = 6,
  kEpiFoldBiasGeluBf16 = 7,
  kEpiResidualStatsF32 = 8
};

struct GemmParams {
  int M, N, K;            // D[M,N] = A[M,K] * W[N,K]^T, A and W bf16 with K contiguous
  int epilogue;
  const float* bias;      // [N]
  const float* gamma;     // [N] LayerScale (kEpiResidualF32)
  const float* pos;       // [T, N] position embedding incl. cls row (kEpiPatchEmbed)
  int tokens_per_image;   // HW (kEpiPatchEmbed): row m -> image m / HW, token 1 + m % HW
  void* out;              // bf16 [M, ldo] or f32 [rows, ldo]
  int ldo;
  // kEpiResidualLnF32 only
  const float* ln_w;      // [N] LayerNorm weight / bias applied to the updated rows
  const float* ln_b;
  __nv_bfloat16* ln_out;  // bf16 [M, N]; must not alias the A operand
  int* tile_counters;     // [ceil(M / 128)] zero before the launch; the kernel leaves them zero
  // kEpiResidualStatsF32 (N % 128 == 0)
  __nv_bfloat16* xb_out;  // bf16 [M, N]: copy of the updated rows
  float* part_out;        // f32 [M, N / 128, 2]: per row and 128-column slice, sum x and sum x^2
  // kEpiFoldBias*Bf16
  const float* ln_part;   // f32 [M, ln_slices, 2] written by the producing GEMM (row statistics over K columns)
  int ln_slices;          // K / 128
  const float* colsum;    // f32 [N]
};
int gemm_smem_bytes();
// Two-SM (cta_group::2) form, vit_gemm2.cu: a cluster of two CTAs per 256 x 256 tile.  Same results bit for bit.
// map_w must be built with a {64, 128} box (each CTA stages half of the 256 weight rows).
bool gemm2_supports(int epilogue);
cudaError_t launch_gemm2(const CUtensorMap& map_a, const CUtensorMap& map_w, const GemmParams& p,
                         int num_sms, cudaStream_t stream);
// LayerNorm (f32 [rows, C] -> bf16) of the rows a kEpiResidualPublishF32 GEMM with `tiles_n` column tiles is
// producing: launched right after that GEMM on the same stream, runs next to it (one small CTA per SM), takes
// each 32-row unit as soon as the unit's 128-row block has been counted 8 * `tiles_n` times, reads it out of L2.
// Same arithmetic as launch_layernorm_bf16 (bit-identical).  counters / done: [ceil(rows / 128)] ints each, zero
// before the GEMM's launch; left zero.
cudaError_t launch_layernorm_follow(const float* x, const float* w, const float* b, __nv_bfloat16* out, int rows,
                                    int C, int tiles_n, int* counters, int* done, int num_sms, cudaStream_t stream);
cudaError_t launch_gemm(const CUtensorMap& map_a, const CUtensorMap& map_w, const GemmParams& p,
                        int num_sms, cudaStream_t stream);

struct AttnParams {
  int B, T, heads;        // head_dim is 64
  float scale;            // 1/sqrt(64)
  __nv_bfloat16* out;     // [B*T, heads*64]
};
int attn_smem_bytes();
// map_qkv: 3-D map over qkv bf16 [B, T, 3*heads*64] with box {64, 128, 1}
cudaError_t launch_attention(const CUtensorMap& map_qkv, const AttnParams& p, cudaStream_t stream);

// LayerNorm over the last dim (eps 1e-6): f32 [rows, C] -> bf16 [rows, C] (GEMM input) or, for the
// final norm, f32 [B, T-1, C] with the cls row dropped.
cudaError_t launch_layernorm_bf16(const float* x, const float* w, const float* b, __nv_bfloat16* out,
                                  int rows, int C, cudaStream_t stream);
cudaError_t launch_layernorm_final(const float* x, const float* w, const float* b, float* out, int B,
                                   int T, int C, cudaStream_t stream);
// u8 HWC RGB images -> ImageNet-normalised bf16 patch rows [B*gh*gw, kpad] (3*14*14 = 588 values in
// conv-weight order c, dy, dx, zero padded to kpad) and the cls rows of x (cls + pos[0]).
cudaError_t launch_patchify(const uint8_t* images, int B, int img_h, int img_w, int bgr,
                            __nv_bfloat16* patches, int kpad, const float* cls_token, const float* pos,
                            float* x, int C, cudaStream_t stream);

}  // namespace vc
