// Bandwidth-bound pieces of the ViT forward: image -> patch rows (fused ToTensor + ImageNet
// normalisation + im2col of the 14x14/s14 patch convolution, vit_extractor.py:117-132 and
// SURVEY.md Appendix A), LayerNorm (eps 1e-6) feeding the GEMMs in bf16, and the final norm that
// also drops the cls token (x_norm_patchtokens).
#include "vc_launch.cuh"
#include "vc_ptx.cuh"
#include "vit_kernels.cuh"

namespace vc {

namespace {

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kVec = 8;  // float4 per lane, C <= 1024

// One warp per row.  Statistics as torch.nn.LayerNorm: biased variance around the mean.
template <bool kFinal>
__global__ void __launch_bounds__(256)
layernorm_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                 void* __restrict__ out, int rows, int T, int C) {
  const int lane = threadIdx.x & 31;
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  griddep_wait();  // x is the previous kernel's output
  if (row >= rows) return;
  size_t out_row = row;
  if (kFinal) {
    const int img = row / T, t = row - img * T;
    if (t == 0) return;  // cls token
    out_row = (size_t)img * (T - 1) + t - 1;
  }
  const int nvec = C >> 2;
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * C);
  float4 v[kVec];
  float s = 0.f;
#pragma unroll
  for (int k = 0; k < kVec; ++k) {
    const int i = lane + 32 * k;
    v[k] = i < nvec ? xr[i] : make_float4(0, 0, 0, 0);
    s += v[k].x + v[k].y + v[k].z + v[k].w;
  }
  const float mean = warp_sum(s) / (float)C;
  float ss = 0.f;
#pragma unroll
  for (int k = 0; k < kVec; ++k) {
    const int i = lane + 32 * k;
    if (i < nvec) {
      const float a = v[k].x - mean, c = v[k].y - mean, d = v[k].z - mean, e = v[k].w - mean;
      ss += a * a + c * c + d * d + e * e;
    }
  }
  const float rstd = rsqrtf(warp_sum(ss) / (float)C + 1e-6f);
#pragma unroll
  for (int k = 0; k < kVec; ++k) {
    const int i = lane + 32 * k;
    if (i < nvec) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(w) + i);
      const float4 be = __ldg(reinterpret_cast<const float4*>(b) + i);
      float4 o;
      o.x = (v[k].x - mean) * rstd * g.x + be.x;
      o.y = (v[k].y - mean) * rstd * g.y + be.y;
      o.z = (v[k].z - mean) * rstd * g.z + be.z;
      o.w = (v[k].w - mean) * rstd * g.w + be.w;
      if (kFinal) {
        reinterpret_cast<float4*>(static_cast<float*>(out) + out_row * C)[i] = o;
      } else {
        __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
        uint2 pk;
        pk.x = *reinterpret_cast<unsigned*>(&lo);
        pk.y = *reinterpret_cast<unsigned*>(&hi);
        reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(out) + out_row * C)[i] = pk;
      }
    }
  }
}

// One thread produces 8 consecutive bf16 of a patch row (one 16-byte store).  The per-channel
// ToTensor + Normalize value of every byte, (u / 255 - mean) / std in the reference's fp32 order
// (vit_extractor.py:117-132), comes from a 3 x 256 table built once per CTA, so the result is bit-identical
// to evaluating the two divisions per element.
__global__ void __launch_bounds__(256)
patchify_kernel(const uint8_t* __restrict__ images, int B, int img_h, int img_w, int bgr,
                __nv_bfloat16* __restrict__ patches, int kpad, const float* __restrict__ cls_token,
                const float* __restrict__ pos, float* __restrict__ x, int C) {
  __shared__ float lut[3][256];
  for (int e = threadIdx.x; e < 768; e += blockDim.x) {
    const int c = e >> 8, u = e & 255;
    const float mean = c == 0 ? 0.485f : (c == 1 ? 0.456f : 0.406f);
    const float sd = c == 0 ? 0.229f : (c == 1 ? 0.224f : 0.225f);
    lut[c][u] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)u, 255.0f), mean), sd);
  }
  __syncthreads();
  griddep_wait();  // the table above is built from constants; images / patches / x belong to other kernels
  const int gw = img_w / 14, gh = img_h / 14;
  const int vec_per_patch = kpad >> 3;
  const size_t total = (size_t)B * gh * gw * vec_per_patch;
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < total) {
    const int k0 = (int)(i % vec_per_patch) << 3;
    const size_t patch = i / vec_per_patch;
    const int px = (int)(patch % gw);
    const size_t t = patch / gw;
    const int py = (int)(t % gh), img = (int)(t / gh);
    const uint8_t* base = images + (((size_t)img * img_h + py * 14) * img_w + px * 14) * 3;
    int c = k0 / 196, r = k0 - c * 196, dy = r / 14, dx = r - dy * 14;
    float v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      float val = 0.f;
      if (k0 + e < 588) val = lut[c][base[((size_t)dy * img_w + dx) * 3 + (bgr ? 2 - c : c)]];
      v[e] = val;
      if (++dx == 14) {
        dx = 0;
        if (++dy == 14) {
          dy = 0;
          ++c;
        }
      }
    }
    uint4 o;
    __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]);
    __nv_bfloat162 h2 = __floats2bfloat162_rn(v[4], v[5]), h3 = __floats2bfloat162_rn(v[6], v[7]);
    o.x = *reinterpret_cast<unsigned*>(&h0);
    o.y = *reinterpret_cast<unsigned*>(&h1);
    o.z = *reinterpret_cast<unsigned*>(&h2);
    o.w = *reinterpret_cast<unsigned*>(&h3);
    *reinterpret_cast<uint4*>(patches + patch * kpad + k0) = o;
  }
  // cls rows of the residual stream: cls_token + pos[0]
  const int T = gh * gw + 1;
  for (size_t n = i; n < (size_t)B * C; n += (size_t)gridDim.x * blockDim.x) {
    const int img = (int)(n / C), ch = (int)(n % C);
    x[(size_t)img * T * C + ch] = cls_token[ch] + pos[ch];
  }
}

}  // namespace

cudaError_t launch_layernorm_bf16(const float* x, const float* w, const float* b, __nv_bfloat16* out,
                                  int rows, int C, cudaStream_t stream) {
  if (rows <= 0) return cudaSuccess;
  const int blocks = (int)(((long long)rows * 32 + 255) / 256);
  return launch_pdl(layernorm_kernel<false>, dim3(blocks), dim3(256), 0, stream, x, w, b, out, rows, 1, C);
}

cudaError_t launch_layernorm_final(const float* x, const float* w, const float* b, float* out, int B,
                                   int T, int C, cudaStream_t stream) {
  const long long rows = (long long)B * T;
  if (rows <= 0) return cudaSuccess;
  const int blocks = (int)((rows * 32 + 255) / 256);
  return launch_pdl(layernorm_kernel<true>, dim3(blocks), dim3(256), 0, stream, x, w, b, out, (int)rows, T, C);
}

cudaError_t launch_patchify(const uint8_t* images, int B, int img_h, int img_w, int bgr,
                            __nv_bfloat16* patches, int kpad, const float* cls_token, const float* pos,
                            float* x, int C, cudaStream_t stream) {
  const size_t total = (size_t)B * (img_h / 14) * (img_w / 14) * (kpad / 8);  // kpad is a multiple of 64
  if (total == 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((total + 255) / 256);
  return launch_pdl(patchify_kernel, dim3(blocks), dim3(256), 0, stream, images, B, img_h, img_w, bgr, patches, kpad,
                    cls_token, pos, x, C);
}

}  // namespace vc
