// C ABI, ViT forward half (see include/vc_b200.h): tensor-map creation, workspace carving and the
// per-layer launch sequence.  No CPU path.
#include "vc_internal.cuh"
#include "vit_kernels.cuh"

#include <cstdlib>
#include <cstring>

using vc::fail;

namespace {

// 2-D bf16 row-major [rows, cols] map with a {64, box_rows} box, 128-byte swizzle
int make_map_2d(vc_handle* h, CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols,
                uint32_t box_rows) {
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || cols % 8 != 0)
    return fail(VC_ERR_INVALID, "GEMM operand must be 16-byte aligned with a multiple-of-8 row length");
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {cols * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = h->encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(VC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return VC_OK;
}

// 3-D map over qkv bf16 [B, T, 3C]: box {64, 128, 1}
int make_map_qkv(vc_handle* h, CUtensorMap* map, const void* ptr, int B, int T, int C3) {
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return fail(VC_ERR_INVALID, "qkv must be 16-byte aligned");
  cuuint64_t gdim[3] = {(cuuint64_t)C3, (cuuint64_t)T, (cuuint64_t)B};
  cuuint64_t gstride[2] = {(cuuint64_t)C3 * 2, (cuuint64_t)C3 * 2 * (cuuint64_t)T};
  cuuint32_t box[3] = {64, 128, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = h->encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstride,
                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(VC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return VC_OK;
}

int run_linear(vc_handle* h, const void* a, const void* w, int m, int n, int k, vc::GemmParams p,
               cudaStream_t s) {
  // VC_VIT_2CTA=0 keeps every GEMM on the one-SM kernel (default: the two-SM kernel, vit_gemm2.cu, where it exists)
  const char* env_2cta = getenv("VC_VIT_2CTA");
  const bool two_sm_on = !(env_2cta && env_2cta[0] == '0');
  // K < 768 (ViT-S): six k-steps per tile do not pay for the cluster's hand-over latency (measured 1 % slower)
  const bool two_sm = two_sm_on && vc::gemm2_supports(p.epilogue) && k >= 768;
  CUtensorMap ma, mw;
  int rc = make_map_2d(h, &ma, a, (uint64_t)m, (uint64_t)k, 128);
  if (rc != VC_OK) return rc;
  rc = make_map_2d(h, &mw, w, (uint64_t)n, (uint64_t)k, two_sm ? 128 : 256);
  if (rc != VC_OK) return rc;
  p.M = m;
  p.N = n;
  p.K = k;
  if (two_sm)
    VC_CUDA(vc::launch_gemm2(ma, mw, p, h->num_sms, s));
  else
    VC_CUDA(vc::launch_gemm(ma, mw, p, h->num_sms, s));
  vc::add_launches(1);
  return VC_OK;
}

struct VitWs {
  float* x;                // [B*T, C] residual stream
  __nv_bfloat16* y;        // [B*T, C] LayerNorm-1 output / attention output
  __nv_bfloat16* y2;       // [B*T, C] LayerNorm-2 output (written by the proj GEMM while it still reads y)
  __nv_bfloat16* wide;     // [B*T, 4C] qkv (3C) / MLP hidden (4C) / patch rows (kpad)
  int* counters;           // [2 * ceil(B*T / 128)] row-block counters (tiles stored | units normalised)
  float* part;             // [2][B*T, C / 128, 2] per-slice row sums of the folded LayerNorms (ln1 | ln2)
};

size_t ws_bytes(int B, int T, int C, VitWs* out, void* base) {
  const size_t rows = (size_t)B * T;
  const size_t bx = vc::align_up(rows * C * 4, 1024), by = vc::align_up(rows * C * 2, 1024);
  const size_t bw = vc::align_up(rows * (size_t)(4 * C > 640 ? 4 * C : 640) * 2, 1024);
  const size_t bc = vc::align_up(2 * ((rows + 127) / 128) * sizeof(int), 1024);
  const size_t bp = vc::align_up(rows * (size_t)((C + 127) / 128) * 2 * sizeof(float), 1024);
  if (out) {
    uint8_t* p = static_cast<uint8_t*>(base);
    out->x = reinterpret_cast<float*>(p);
    out->y = reinterpret_cast<__nv_bfloat16*>(p + bx);
    out->y2 = reinterpret_cast<__nv_bfloat16*>(p + bx + by);
    out->wide = reinterpret_cast<__nv_bfloat16*>(p + bx + 2 * by);
    out->counters = reinterpret_cast<int*>(p + bx + 2 * by + bw);
    out->part = reinterpret_cast<float*>(p + bx + 2 * by + bw + bc);
  }
  return bx + 2 * by + bw + bc + 2 * bp;
}

}  // namespace

extern "C" {

size_t vc_vit_workspace_bytes(int n_images, int grid_h, int grid_w, int embed_dim) {
  if (n_images <= 0 || grid_h <= 0 || grid_w <= 0 || embed_dim <= 0) return 1024;
  return ws_bytes(n_images, grid_h * grid_w + 1, embed_dim, nullptr, nullptr) + 1024;
}

int vc_vit_linear(vc_handle* h, const void* a, const void* w, const float* bias, const float* gamma,
                  int m, int n, int k, int epilogue, void* out, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!a || !w || !bias || !out) return fail(VC_ERR_INVALID, "null pointer argument");
  if (m <= 0 || n <= 0 || k <= 0 || k % 64 != 0 || n % 32 != 0)
    return fail(VC_ERR_INVALID, "bad GEMM shape m=%d n=%d k=%d (k %% 64, n %% 32)", m, n, k);
  if (epilogue < 0 || epilogue > 2) return fail(VC_ERR_INVALID, "bad epilogue %d", epilogue);
  if (epilogue == 2 && !gamma) return fail(VC_ERR_INVALID, "residual epilogue needs gamma");
  VC_CUDA(cudaSetDevice(h->device));
  vc::GemmParams p{};
  p.epilogue = epilogue;
  p.bias = bias;
  p.gamma = gamma;
  p.out = out;
  p.ldo = n;
  return run_linear(h, a, w, m, n, k, p, reinterpret_cast<cudaStream_t>(stream));
}

int vc_vit_linear_ln(vc_handle* h, const void* a, const void* w, const float* bias, const float* gamma,
                     const float* ln_w, const float* ln_b, int m, int n, int k, float* x, void* ln_out_bf16,
                     void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!a || !w || !bias || !gamma || !ln_w || !ln_b || !x || !ln_out_bf16)
    return fail(VC_ERR_INVALID, "null pointer argument");
  if (m <= 0 || n <= 0 || k <= 0 || k % 64 != 0 || n % 32 != 0 || n > 1024)
    return fail(VC_ERR_INVALID, "bad GEMM shape m=%d n=%d k=%d (k %% 64, n %% 32, n <= 1024)", m, n, k);
  if (ln_out_bf16 == a) return fail(VC_ERR_INVALID, "the LayerNorm output must not alias the A operand");
  VC_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const size_t blocks = ((size_t)m + 127) / 128;
  if (h->ln_counters_len < blocks) {
    if (h->ln_counters) cudaFree(h->ln_counters);
    h->ln_counters = nullptr;
    h->ln_counters_len = 0;
    VC_CUDA(cudaMalloc(&h->ln_counters, blocks * sizeof(int)));
    h->ln_counters_len = blocks;
  }
  VC_CUDA(cudaMemsetAsync(h->ln_counters, 0, blocks * sizeof(int), s));
  vc::GemmParams p{};
  p.epilogue = vc::kEpiResidualLnF32;
  p.bias = bias;
  p.gamma = gamma;
  p.out = x;
  p.ldo = n;
  p.ln_w = ln_w;
  p.ln_b = ln_b;
  p.ln_out = static_cast<__nv_bfloat16*>(ln_out_bf16);
  p.tile_counters = h->ln_counters;
  return run_linear(h, a, w, m, n, k, p, s);
}

int vc_vit_attention(vc_handle* h, const void* qkv, int n_images, int tokens, int heads, void* out,
                     void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!qkv || !out) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_images <= 0 || tokens <= 0 || heads <= 0) return fail(VC_ERR_INVALID, "bad attention shape");
  VC_CUDA(cudaSetDevice(h->device));
  CUtensorMap map;
  int rc = make_map_qkv(h, &map, qkv, n_images, tokens, 3 * heads * 64);
  if (rc != VC_OK) return rc;
  vc::AttnParams p{};
  p.B = n_images;
  p.T = tokens;
  p.heads = heads;
  p.scale = 0.125f;
  p.out = static_cast<__nv_bfloat16*>(out);
  VC_CUDA(vc::launch_attention(map, p, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_vit_layernorm(vc_handle* h, const float* x, const float* w, const float* b, int rows, int channels,
                     void* out_bf16, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!x || !w || !b || !out_bf16) return fail(VC_ERR_INVALID, "null pointer argument");
  if (rows <= 0 || channels <= 0 || channels % 4 != 0 || channels > 1024)
    return fail(VC_ERR_INVALID, "bad LayerNorm shape");
  VC_CUDA(cudaSetDevice(h->device));
  VC_CUDA(vc::launch_layernorm_bf16(x, w, b, static_cast<__nv_bfloat16*>(out_bf16), rows, channels,
                                    reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_vit_forward(vc_handle* h, const vc_vit_weights* w, const uint8_t* images, int bgr, int n_images,
                   int grid_h, int grid_w, float* tokens, void* workspace, size_t workspace_bytes,
                   void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!w || !images || !tokens || !workspace) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_images <= 0 || grid_h <= 0 || grid_w <= 0) return fail(VC_ERR_INVALID, "bad sizes");
  const int C = w->embed_dim, heads = w->num_heads, kpad = w->patch_kpad;
  if (C <= 0 || C % 64 != 0 || C > 1024 || heads * 64 != C || w->depth <= 0 || !w->blocks)
    return fail(VC_ERR_INVALID, "bad model (embed_dim=%d heads=%d depth=%d)", C, heads, w->depth);
  if (kpad < 588 || kpad % 64 != 0) return fail(VC_ERR_INVALID, "patch_kpad=%d must be >= 588 and %% 64", kpad);
  if (workspace_bytes < vc_vit_workspace_bytes(n_images, grid_h, grid_w, C))
    return fail(VC_ERR_WORKSPACE, "workspace %zu B < required %zu B", workspace_bytes,
                vc_vit_workspace_bytes(n_images, grid_h, grid_w, C));
  VC_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int HW = grid_h * grid_w, T = HW + 1, rows = n_images * T;
  VitWs ws;
  uint8_t* base = reinterpret_cast<uint8_t*>(vc::align_up(reinterpret_cast<uintptr_t>(workspace), 1024));
  ws_bytes(n_images, T, C, &ws, base);

  // patch embedding: u8 image -> normalised bf16 patch rows -> GEMM (+bias +pos) into x[:, 1:, :]
  VC_CUDA(vc::launch_patchify(images, n_images, grid_h * 14, grid_w * 14, bgr, ws.wide, kpad, w->cls_token,
                              w->pos_embed, ws.x, C, s));
  vc::add_launches(1);
  {
    vc::GemmParams p{};
    p.epilogue = vc::kEpiPatchEmbed;
    p.bias = w->patch_b;
    p.pos = w->pos_embed;
    p.tokens_per_image = HW;
    p.out = ws.x;
    p.ldo = C;
    int rc = run_linear(h, ws.wide, w->patch_w, n_images * HW, C, kpad, p, s);
    if (rc != VC_OK) return rc;
  }
  CUtensorMap map_qkv;
  int rc = make_map_qkv(h, &map_qkv, ws.wide, n_images, T, 3 * C);
  if (rc != VC_OK) return rc;
  // LayerNorm placement.  Default: one launch per LayerNorm.  VC_VIT_FUSED_LN=1 runs LN2 of a block in the
  // epilogue of its proj GEMM (into y2: the GEMM still reads its A operand from y) and LN1 of block l + 1 in
  // the epilogue of block l's fc2 GEMM (into y) -- bit-identical tokens, but measured SLOWER (13.95 vs
  // 11.96 ms per 32 images, see vit_gemm.cu), so it is not the default.
  // VC_VIT_LN=follow: the residual GEMM publishes its row blocks and a small LayerNorm kernel launched right
  // behind it (programmatic dependent launch) normalises them out of L2 while the GEMM runs -- bit-identical,
  // measured SLOWER (12.9 vs 11.9 ms per 32 images: next to the GEMM's 384 x 144 registers only 4 warps of
  // LayerNorm fit on an SM, a third of the stand-alone kernel's speed), so not the default.
  const char* fused_env = getenv("VC_VIT_FUSED_LN");
  const bool fused_ln = fused_env && fused_env[0] == '1';
  const char* ln_env = getenv("VC_VIT_LN");
  const bool follow_ln = !fused_ln && ln_env && strcmp(ln_env, "follow") == 0;
  // VC_VIT_LN=fold: no LayerNorm launches between the first and the last one -- the residual GEMMs store a bf16
  // copy of the stream and per-slice row sums, the next GEMM runs on that copy with ln_w folded into its weights
  // and applies mean / rstd in its epilogue (vit_kernels.cuh).  Needs the folded weights and C % 128 == 0.
  bool fold_ln = !fused_ln && ln_env && strcmp(ln_env, "fold") == 0 && C % 128 == 0;
  for (int l = 0; fold_ln && l < w->depth; ++l) {
    const vc_vit_block& b = w->blocks[l];
    fold_ln = b.qkv_wf && b.qkv_cs && b.qkv_bf && b.fc1_wf && b.fc1_cs && b.fc1_bf;
  }
  const int slices = C / 128;
  float* part1 = ws.part;                                         // statistics of x before a block's LN1
  float* part2 = ws.part + (size_t)rows * ((C + 127) / 128) * 2;  // ... before its LN2
  const int n_blocks = (rows + 127) / 128;
  int* done = ws.counters + n_blocks;
  const int tiles_c = (C + 255) / 256;
  VC_CUDA(cudaMemsetAsync(ws.counters, 0, (size_t)2 * n_blocks * sizeof(int), s));
  VC_CUDA(vc::launch_layernorm_bf16(ws.x, w->blocks[0].ln1_w, w->blocks[0].ln1_b, ws.y, rows, C, s));
  vc::add_launches(1);
  for (int l = 0; l < w->depth; ++l) {
    const vc_vit_block& b = w->blocks[l];
    const bool last = l + 1 == w->depth;
    vc::GemmParams p{};
    const bool fold_qkv = fold_ln && l > 0;  // block 0's LN1 follows the patch embedding: a plain launch
    p.epilogue = fold_qkv ? vc::kEpiFoldBiasBf16 : vc::kEpiBiasBf16;
    p.bias = fold_qkv ? b.qkv_bf : b.qkv_b;
    p.ln_part = part1;
    p.ln_slices = slices;
    p.colsum = b.qkv_cs;
    p.out = ws.wide;
    p.ldo = 3 * C;
    rc = run_linear(h, ws.y, fold_qkv ? b.qkv_wf : b.qkv_w, rows, 3 * C, C, p, s);
    if (rc != VC_OK) return rc;
    vc::AttnParams ap{};
    ap.B = n_images;
    ap.T = T;
    ap.heads = heads;
    ap.scale = 0.125f;
    ap.out = ws.y;
    VC_CUDA(vc::launch_attention(map_qkv, ap, s));
    vc::add_launches(1);
    p = vc::GemmParams{};
    p.epilogue = fused_ln ? vc::kEpiResidualLnF32 : (follow_ln ? vc::kEpiResidualPublishF32 : vc::kEpiResidualF32);
    if (fold_ln) p.epilogue = vc::kEpiResidualStatsF32;
    p.xb_out = ws.y2;
    p.part_out = part2;
    p.bias = b.proj_b;
    p.gamma = b.ls1;
    p.out = ws.x;
    p.ldo = C;
    p.ln_w = b.ln2_w;
    p.ln_b = b.ln2_b;
    p.ln_out = ws.y2;
    p.tile_counters = ws.counters;
    rc = run_linear(h, ws.y, b.proj_w, rows, C, C, p, s);
    if (rc != VC_OK) return rc;
    if (follow_ln) {
      VC_CUDA(vc::launch_layernorm_follow(ws.x, b.ln2_w, b.ln2_b, ws.y2, rows, C, tiles_c, ws.counters, done,
                                          h->num_sms, s));
      vc::add_launches(1);
    } else if (!fused_ln && !fold_ln) {
      VC_CUDA(vc::launch_layernorm_bf16(ws.x, b.ln2_w, b.ln2_b, ws.y2, rows, C, s));
      vc::add_launches(1);
    }
    p = vc::GemmParams{};
    p.epilogue = fold_ln ? vc::kEpiFoldBiasGeluBf16 : vc::kEpiBiasGeluBf16;
    p.bias = fold_ln ? b.fc1_bf : b.fc1_b;
    p.ln_part = part2;
    p.ln_slices = slices;
    p.colsum = b.fc1_cs;
    p.out = ws.wide;
    p.ldo = 4 * C;
    rc = run_linear(h, ws.y2, fold_ln ? b.fc1_wf : b.fc1_w, rows, 4 * C, C, p, s);
    if (rc != VC_OK) return rc;
    p = vc::GemmParams{};
    p.epilogue = (fused_ln && !last) ? vc::kEpiResidualLnF32
                                     : ((follow_ln && !last) ? vc::kEpiResidualPublishF32 : vc::kEpiResidualF32);
    if (fold_ln && !last) p.epilogue = vc::kEpiResidualStatsF32;
    p.xb_out = ws.y;
    p.part_out = part1;
    p.bias = b.fc2_b;
    p.gamma = b.ls2;
    p.out = ws.x;
    p.ldo = C;
    if (!last) {
      p.ln_w = w->blocks[l + 1].ln1_w;
      p.ln_b = w->blocks[l + 1].ln1_b;
      p.ln_out = ws.y;
      p.tile_counters = ws.counters;
      }
    rc = run_linear(h, ws.wide, b.fc2_w, rows, C, 4 * C, p, s);
    if (rc != VC_OK) return rc;
    if (follow_ln && !last) {
      VC_CUDA(vc::launch_layernorm_follow(ws.x, w->blocks[l + 1].ln1_w, w->blocks[l + 1].ln1_b, ws.y, rows, C, tiles_c,
                                          ws.counters, done, h->num_sms, s));
      vc::add_launches(1);
    } else if (!fused_ln && !fold_ln && !last) {
      VC_CUDA(vc::launch_layernorm_bf16(ws.x, w->blocks[l + 1].ln1_w, w->blocks[l + 1].ln1_b, ws.y, rows, C, s));
      vc::add_launches(1);
    }
  }
  VC_CUDA(vc::launch_layernorm_final(ws.x, w->norm_w, w->norm_b, tokens, n_images, T, C, s));
  vc::add_launches(1);
  return VC_OK;
}

}  // extern "C"
