// Bandwidth-bound pieces of the ViTFeatureModel head (vit_colmap/model/vit_feature_model.py):
// bilinear resize of the upsampled map to (14H//4, 14W//4) :199-212, the keypoint head's 1x1
// convolution 64 -> 4 with tanh(.)*pi on the orientation channel :112,222-223, and
// F.normalize(dim=1, eps=1e-8) of the descriptor map :226.  Channels-last, 16-byte accesses.
#include "head_kernels.cuh"
#include "vc_launch.cuh"
#include "vc_ptx.cuh"

namespace vc {

namespace {

__device__ __forceinline__ void unpack8(const uint4& r, float (&f)[8]) {
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
  }
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// ATen's area_pixel_compute_source_index (align_corners=False, non-cubic): the source coordinate is
// clamped at 0; the upper neighbour is clamped at the last index.
__device__ __forceinline__ void src_index(int dst, float scale, int n_in, int& i0, int& i1, float& l1) {
  const float s = fmaxf(__fsub_rn(__fmul_rn(scale, __fadd_rn((float)dst, 0.5f)), 0.5f), 0.f);
  i0 = min((int)s, n_in - 1);
  i1 = min(i0 + 1, n_in - 1);
  l1 = __fsub_rn(s, (float)i0);
}

// thread = 8 channels of one output pixel
__global__ void __launch_bounds__(256) resize_bilinear_kernel(const __nv_bfloat16* __restrict__ in,
                                                              __nv_bfloat16* __restrict__ out, int B, int Hi,
                                                              int Wi, int Ho, int Wo, int C, float sy, float sx) {
  const int cv = C >> 3;
  const long long total = (long long)B * Ho * Wo * cv;
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const int v = (int)(i % cv);
    long long pix = i / cv;
    const int x = (int)(pix % Wo);
    pix /= Wo;
    const int y = (int)(pix % Ho);
    const int b = (int)(pix / Ho);
    int y0, y1, x0, x1;
    float ly1, lx1;
    src_index(y, sy, Hi, y0, y1, ly1);
    src_index(x, sx, Wi, x0, x1, lx1);
    const float ly0 = __fsub_rn(1.f, ly1), lx0 = __fsub_rn(1.f, lx1);
    const uint4* base = reinterpret_cast<const uint4*>(in) + (size_t)b * Hi * Wi * cv + v;
    float a[8], c[8], e[8], g[8];
    unpack8(__ldg(base + ((size_t)y0 * Wi + x0) * cv), a);
    unpack8(__ldg(base + ((size_t)y0 * Wi + x1) * cv), c);
    unpack8(__ldg(base + ((size_t)y1 * Wi + x0) * cv), e);
    unpack8(__ldg(base + ((size_t)y1 * Wi + x1) * cv), g);
    float o[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const float top = __fadd_rn(__fmul_rn(a[k], lx0), __fmul_rn(c[k], lx1));
      const float bot = __fadd_rn(__fmul_rn(e[k], lx0), __fmul_rn(g[k], lx1));
      o[k] = __fadd_rn(__fmul_rn(top, ly0), __fmul_rn(bot, ly1));
    }
    uint4 w;
    w.x = pack2(o[0], o[1]);
    w.y = pack2(o[2], o[3]);
    w.z = pack2(o[4], o[5]);
    w.w = pack2(o[6], o[7]);
    reinterpret_cast<uint4*>(out)[i] = w;
  }
}

// thread = one pixel: 64 bf16 of hidden (8 x 16 B), 4 x 64 fp32 weights from shared memory
__global__ void __launch_bounds__(256) kp_head_kernel(const __nv_bfloat16* __restrict__ hidden,
                                                      const float* __restrict__ w, const float* __restrict__ b,
                                                      float* __restrict__ kp, long long pixels) {
  __shared__ float sw[4 * 64];
  __shared__ float sb[4];
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  for (int i = threadIdx.x; i < 256; i += blockDim.x) sw[i] = w[i];
  if (threadIdx.x < 4) sb[threadIdx.x] = b[threadIdx.x];
  __syncthreads();
  for (long long px = (long long)blockIdx.x * blockDim.x + threadIdx.x; px < pixels;
       px += (long long)gridDim.x * blockDim.x) {
    const uint4* row = reinterpret_cast<const uint4*>(hidden + px * 64);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int v = 0; v < 8; ++v) {
      float f[8];
      unpack8(__ldg(row + v), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
#pragma unroll
        for (int o = 0; o < 4; ++o) acc[o] = fmaf(f[k], sw[o * 64 + v * 8 + k], acc[o]);
      }
    }
    float4 r;
    r.x = acc[0] + sb[0];
    r.y = acc[1] + sb[1];
    r.z = acc[2] + sb[2];
    r.w = tanhf(acc[3] + sb[3]) * 3.14159265358979323846f;
    reinterpret_cast<float4*>(kp)[px] = r;
  }
}

// warp = one pixel row of d floats
__global__ void __launch_bounds__(256) normalize_rows_kernel(float* __restrict__ desc, long long pixels, int d) {
  const int lane = threadIdx.x & 31;
  const int dv = d >> 2;
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  for (long long px = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; px < pixels;
       px += ((long long)gridDim.x * blockDim.x) >> 5) {
    float4* row = reinterpret_cast<float4*>(desc + px * d);
    float ss = 0.f;
    for (int v = lane; v < dv; v += 32) {
      const float4 x = row[v];
      ss += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
    const float den = fmaxf(sqrtf(ss), 1e-8f);
    for (int v = lane; v < dv; v += 32) {
      float4 x = row[v];
      x.x = __fdiv_rn(x.x, den);
      x.y = __fdiv_rn(x.y, den);
      x.z = __fdiv_rn(x.z, den);
      x.w = __fdiv_rn(x.w, den);
      row[v] = x;
    }
  }
}

inline int grid_for(long long threads, int block, int cap) {
  long long g = (threads + block - 1) / block;
  return (int)(g < 1 ? 1 : (g > cap ? cap : g));
}

}  // namespace

cudaError_t launch_resize_bilinear(const __nv_bfloat16* in, __nv_bfloat16* out, int B, int Hi, int Wi,
                                   int Ho, int Wo, int C, cudaStream_t stream) {
  if (B <= 0 || Ho <= 0 || Wo <= 0) return cudaSuccess;
  if (C % 8 != 0) return cudaErrorInvalidValue;
  const long long total = (long long)B * Ho * Wo * (C >> 3);
  // F.interpolate(size=...) uses in/out as the scale (fp32)
  const float sy = (float)Hi / (float)Ho, sx = (float)Wi / (float)Wo;
  return launch_pdl(resize_bilinear_kernel, dim3(grid_for(total, 256, 148 * 64)), dim3(256), 0, stream, in, out, B, Hi, Wi,
                    Ho, Wo, C, sy, sx);
}

cudaError_t launch_kp_head(const __nv_bfloat16* hidden, const float* w, const float* b, float* kp,
                           long long pixels, cudaStream_t stream) {
  if (pixels <= 0) return cudaSuccess;
  return launch_pdl(kp_head_kernel, dim3(grid_for(pixels, 256, 148 * 16)), dim3(256), 0, stream, hidden, w, b, kp, pixels);
}

cudaError_t launch_normalize_rows(float* desc, long long pixels, int d, cudaStream_t stream) {
  if (pixels <= 0) return cudaSuccess;
  if (d % 4 != 0) return cudaErrorInvalidValue;
  return launch_pdl(normalize_rows_kernel, dim3(grid_for(pixels * 32, 256, 148 * 32)), dim3(256), 0, stream, desc, pixels, d);
}

}  // namespace vc
