// C ABI, ViTFeatureModel convolutional head (see include/vc_b200.h): tensor-map creation,
// workspace carving and the launch sequence of vit_feature_model.py:234-295.  No CPU path.
#include "vc_internal.cuh"
#include <cstdlib>

#include "head_kernels.cuh"

using vc::fail;

namespace {

// 4-D map over bf16 activations [B, H, W, C] (C may be a channel sub-range of rows `ld` wide)
int make_map_act(vc_handle* h, CUtensorMap* map, const void* ptr, int B, int H, int W, int C, int ld) {
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || ld % 8 != 0 || C % 64 != 0)
    return fail(VC_ERR_INVALID, "conv input must be 16-byte aligned, channels %% 64 == 0");
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)B};
  cuuint64_t gstride[3] = {(cuuint64_t)ld * 2, (cuuint64_t)ld * 2 * W, (cuuint64_t)ld * 2 * W * H};
  cuuint32_t box[4] = {64, (cuuint32_t)vc::kConvTileW, (cuuint32_t)vc::kConvTileH, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = h->encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(ptr), gdim, gstride,
                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(VC_ERR_CUDA, "cuTensorMapEncodeTiled (activations) failed with CUresult %d", (int)r);
  return VC_OK;
}

int make_map_w(vc_handle* h, CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint32_t box_rows) {
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || cols % 64 != 0)
    return fail(VC_ERR_INVALID, "conv weights must be 16-byte aligned with K %% 64 == 0");
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {cols * 2};
  cuuint32_t box[2] = {64, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = h->encode(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                         box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(VC_ERR_CUDA, "cuTensorMapEncodeTiled (weights) failed with CUresult %d", (int)r);
  return VC_OK;
}

enum ConvKind { kConv3x3 = 0, kDeconv4x4s2 = 1, kConv1x1 = 2 };

void fill_taps(vc::ConvParams& p, int kind) {
  p.sy = p.sx = 1;
  p.phases = 1;
  if (kind == kConv3x3) {
    p.taps = 9;
    for (int t = 0; t < 9; ++t) {
      p.dy[0][t] = (int8_t)(t / 3 - 1);
      p.dx[0][t] = (int8_t)(t % 3 - 1);
    }
  } else if (kind == kConv1x1) {
    p.taps = 1;
  } else {
    // oy = 2*iy - 1 + ky: phase 0 reads (y-1: k3, y: k1), phase 1 reads (y: k2, y+1: k0)
    p.phases = 4;
    p.taps = 4;
    p.sy = p.sx = 2;
    const int off[2][2] = {{-1, 0}, {0, 1}};
    for (int ph = 0; ph < 4; ++ph) {
      const int py = ph >> 1, px = ph & 1;
      p.oy[ph] = (int8_t)py;
      p.ox[ph] = (int8_t)px;
      for (int t = 0; t < 4; ++t) {
        p.dy[ph][t] = (int8_t)off[py][t >> 1];
        p.dx[ph][t] = (int8_t)off[px][t & 1];
      }
    }
  }
}

struct ConvOut {
  void* out;
  int ldo;
  int split_n = 0;
  void* out2 = nullptr;
  int ldo2 = 0;
};

int run_conv(vc_handle* h, const void* in, int in_ld, int B, int H, int W, int Cin, const void* w, int Cout,
             int kind, int epilogue, const float* scale, const float* shift, const ConvOut& o, cudaStream_t s) {
  vc::ConvParams p{};
  fill_taps(p, kind);
  p.B = B;
  p.H = H;
  p.W = W;
  p.Cin = Cin;
  p.Cout = Cout;
  p.Ho = H * p.sy;
  p.Wo = W * p.sx;
  p.epilogue = epilogue;
  p.scale = scale;
  p.shift = shift;
  p.out = o.out;
  p.ldo = o.ldo;
  p.split_n = o.split_n;
  p.out2 = o.out2;
  p.ldo2 = o.ldo2;
  CUtensorMap ma, mw;
  int rc = make_map_act(h, &ma, in, B, H, W, Cin, in_ld);
  if (rc != VC_OK) return rc;
  // VC_CONV_2CTA=0 keeps the convolutions on the one-SM kernel (default: the two-SM kernel, conv_gemm2.cu)
  const char* env_2cta = getenv("VC_CONV_2CTA");
  const bool two_sm = !(env_2cta && env_2cta[0] == '0');
  rc = make_map_w(h, &mw, w, (uint64_t)p.phases * Cout, (uint64_t)p.taps * Cin, two_sm ? 128 : 256);
  if (rc != VC_OK) return rc;
  if (two_sm)
    VC_CUDA(vc::launch_conv2(ma, mw, p, h->num_sms, s));
  else
    VC_CUDA(vc::launch_conv(ma, mw, p, h->num_sms, s));
  vc::add_launches(1);
  return VC_OK;
}

struct HeadWs {
  __nv_bfloat16 *a, *b;   // ping-pong activation buffers
  __nv_bfloat16 *trunk, *kp_hidden, *desc_hidden;
};

size_t head_ws(int B, int gh, int gw, int C, int oh, int ow, HeadWs* out, void* base) {
  const size_t up2 = (size_t)B * 16 * gh * gw * 512 * 2;          // [B, 4gh, 4gw, 512]
  const size_t up1 = (size_t)B * 4 * gh * gw * 512 * 2;
  const size_t tok = (size_t)B * gh * gw * C * 2;
  const size_t rs = (size_t)B * oh * ow * 512 * 2;
  // a: tokens -> conv1 out -> conv2 out ; b: deconv1 out -> deconv2 out -> resized (after conv2 read b)
  size_t sa = tok > up1 ? tok : up1;
  sa = sa > up2 ? sa : up2;
  size_t sb = up2 > rs ? up2 : rs;
  sa = vc::align_up(sa, 1024);
  sb = vc::align_up(sb, 1024);
  const size_t st = vc::align_up((size_t)B * oh * ow * 256 * 2, 1024);
  const size_t sk = vc::align_up((size_t)B * oh * ow * 64 * 2, 1024);
  const size_t sd = vc::align_up((size_t)B * oh * ow * 128 * 2, 1024);
  if (out) {
    uint8_t* p = static_cast<uint8_t*>(base);
    out->a = reinterpret_cast<__nv_bfloat16*>(p);
    out->b = reinterpret_cast<__nv_bfloat16*>(p + sa);
    out->trunk = reinterpret_cast<__nv_bfloat16*>(p + sa + sb);
    out->kp_hidden = reinterpret_cast<__nv_bfloat16*>(p + sa + sb + st);
    out->desc_hidden = reinterpret_cast<__nv_bfloat16*>(p + sa + sb + st + sk);
  }
  return sa + sb + st + sk + sd;
}

}  // namespace

extern "C" {

size_t vc_head_workspace_bytes(int n_images, int grid_h, int grid_w, int backbone_dim, int out_h, int out_w) {
  if (n_images <= 0 || grid_h <= 0 || grid_w <= 0 || backbone_dim <= 0 || out_h <= 0 || out_w <= 0) return 1024;
  return head_ws(n_images, grid_h, grid_w, backbone_dim, out_h, out_w, nullptr, nullptr) + 1024;
}

int vc_conv2d(vc_handle* h, const void* in, int n, int height, int width, int c_in, const void* w, int c_out,
              int kind, int epilogue, const float* scale, const float* shift, void* out, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!in || !w || !shift || !out) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n <= 0 || height <= 0 || width <= 0 || c_in <= 0 || c_in % 64 != 0 || c_out <= 0 || c_out % 32 != 0)
    return fail(VC_ERR_INVALID, "bad conv shape n=%d h=%d w=%d c_in=%d c_out=%d (c_in %% 64, c_out %% 32)", n,
                height, width, c_in, c_out);
  if (kind < 0 || kind > 2 || epilogue < 0 || epilogue > 2) return fail(VC_ERR_INVALID, "bad kind/epilogue");
  if (epilogue == 1 && !scale) return fail(VC_ERR_INVALID, "BatchNorm epilogue needs scale");
  VC_CUDA(cudaSetDevice(h->device));
  ConvOut o{out, c_out};
  return run_conv(h, in, c_in, n, height, width, c_in, w, c_out, kind, epilogue, scale, shift, o,
                  reinterpret_cast<cudaStream_t>(stream));
}

int vc_resize_bilinear(vc_handle* h, const void* in, int n, int in_h, int in_w, int out_h, int out_w,
                       int channels, void* out, void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!in || !out) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n <= 0 || in_h <= 0 || in_w <= 0 || out_h <= 0 || out_w <= 0 || channels <= 0 || channels % 8 != 0)
    return fail(VC_ERR_INVALID, "bad resize shape");
  VC_CUDA(cudaSetDevice(h->device));
  VC_CUDA(vc::launch_resize_bilinear(static_cast<const __nv_bfloat16*>(in), static_cast<__nv_bfloat16*>(out), n,
                                     in_h, in_w, out_h, out_w, channels, reinterpret_cast<cudaStream_t>(stream)));
  vc::add_launches(1);
  return VC_OK;
}

int vc_conv_head_forward(vc_handle* h, const vc_head_weights* w, const void* tokens, int dtype, int n_images,
                         int grid_h, int grid_w, int out_h, int out_w, float* kp_map, float* desc_map,
                         void* features, void* desc_hidden, void* workspace, size_t workspace_bytes,
                         void* stream) {
  if (!h) return fail(VC_ERR_INVALID, "null handle");
  if (!w || !tokens || !kp_map || !workspace) return fail(VC_ERR_INVALID, "null pointer argument");
  if (n_images <= 0 || grid_h <= 0 || grid_w <= 0 || out_h <= 0 || out_w <= 0)
    return fail(VC_ERR_INVALID, "bad sizes");
  const int C = w->backbone_dim, d = w->descriptor_dim;
  if (C <= 0 || C % 64 != 0) return fail(VC_ERR_INVALID, "backbone_dim=%d must be a multiple of 64", C);
  if (d <= 0 || d % 32 != 0) return fail(VC_ERR_INVALID, "descriptor_dim=%d must be a multiple of 32", d);
  if (dtype != VC_DTYPE_F32 && dtype != VC_DTYPE_BF16) return fail(VC_ERR_INVALID, "tokens must be f32 or bf16");
  const size_t need = vc_head_workspace_bytes(n_images, grid_h, grid_w, C, out_h, out_w);
  if (workspace_bytes < need) return fail(VC_ERR_WORKSPACE, "workspace %zu B < required %zu B", workspace_bytes, need);
  VC_CUDA(cudaSetDevice(h->device));
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  HeadWs ws;
  uint8_t* base = reinterpret_cast<uint8_t*>(vc::align_up(reinterpret_cast<uintptr_t>(workspace), 1024));
  head_ws(n_images, grid_h, grid_w, C, out_h, out_w, &ws, base);
  const int B = n_images;

  const void* x = tokens;
  if (dtype == VC_DTYPE_F32) {
    int rc = vc_convert_f32_to_bf16(static_cast<const float*>(tokens), ws.a, (size_t)B * grid_h * grid_w * C, stream);
    if (rc != VC_OK) return rc;
    x = ws.a;
  }
  // upsampler: (deconv -> conv3x3 + BN + GELU) x 2
  int H = grid_h, W = grid_w, rc;
  rc = run_conv(h, x, C, B, H, W, C, w->deconv_w[0], 512, kDeconv4x4s2, vc::kConvBiasBf16, nullptr,
                w->deconv_b[0], ConvOut{ws.b, 512}, s);
  if (rc != VC_OK) return rc;
  H *= 2;
  W *= 2;
  rc = run_conv(h, ws.b, 512, B, H, W, 512, w->conv_w[0], 512, kConv3x3, vc::kConvBnGeluBf16, w->conv_scale[0],
                w->conv_shift[0], ConvOut{ws.a, 512}, s);
  if (rc != VC_OK) return rc;
  rc = run_conv(h, ws.a, 512, B, H, W, 512, w->deconv_w[1], 512, kDeconv4x4s2, vc::kConvBiasBf16, nullptr,
                w->deconv_b[1], ConvOut{ws.b, 512}, s);
  if (rc != VC_OK) return rc;
  H *= 2;
  W *= 2;
  rc = run_conv(h, ws.b, 512, B, H, W, 512, w->conv_w[1], 512, kConv3x3, vc::kConvBnGeluBf16, w->conv_scale[1],
                w->conv_shift[1], ConvOut{ws.a, 512}, s);
  if (rc != VC_OK) return rc;
  // resize to the target grid (skipped when it already matches, vit_feature_model.py:205)
  const __nv_bfloat16* up = ws.a;
  if (H != out_h || W != out_w) {
    VC_CUDA(vc::launch_resize_bilinear(ws.a, ws.b, B, H, W, out_h, out_w, 512, s));
    vc::add_launches(1);
    up = ws.b;
  }
  __nv_bfloat16* trunk = features ? static_cast<__nv_bfloat16*>(features) : ws.trunk;
  rc = run_conv(h, up, 512, B, out_h, out_w, 512, w->trunk_w, 256, kConv3x3, vc::kConvBnGeluBf16, w->trunk_scale,
                w->trunk_shift, ConvOut{trunk, 256}, s);
  if (rc != VC_OK) return rc;
  // both head 3x3 convolutions as one GEMM (N = 64 + 128), outputs split into two tensors
  __nv_bfloat16* dh = desc_hidden ? static_cast<__nv_bfloat16*>(desc_hidden) : ws.desc_hidden;
  ConvOut ho{ws.kp_hidden, 64, 64, dh, 128};
  rc = run_conv(h, trunk, 256, B, out_h, out_w, 256, w->heads_w, 192, kConv3x3, vc::kConvBnGeluBf16,
                w->heads_scale, w->heads_shift, ho, s);
  if (rc != VC_OK) return rc;
  const long long pixels = (long long)B * out_h * out_w;
  VC_CUDA(vc::launch_kp_head(ws.kp_hidden, w->kp_w, w->kp_b, kp_map, pixels, s));
  vc::add_launches(1);
  if (desc_map) {
    rc = run_conv(h, dh, 128, B, out_h, out_w, 128, w->desc_w, d, kConv1x1, vc::kConvBiasF32, nullptr, w->desc_b,
                  ConvOut{desc_map, d}, s);
    if (rc != VC_OK) return rc;
    VC_CUDA(vc::launch_normalize_rows(desc_map, pixels, d, s));
    vc::add_launches(1);
  }
  return VC_OK;
}

}  // extern "C"
