// Distinctiveness score on the patch grid (reference: vit_colmap/features/vit_extractor.py
// _harris_response :281-348, _dog_response :350-394, _compute_distinctiveness :254-279).
//
// HBM-bound stage.  Kernel 1 (moments_kernel) streams the token-major feature map [B, H*W, C] once:
// a warp walks a strip of cells along x, keeps the current token row in registers (128-bit
// coalesced loads, lanes split the channels), reuses it as the next cell's left neighbour, and
// reduces the channel sums gx^2, gy^2, gx*gy, f with warp shuffles -> 16 B per cell.  The "below"
// row is the only re-read and is served by L2.  Algorithmic bytes = H*W*C*sizeof(elem).
// Kernel 2 (response_kernel) is one CTA per image on the [H, W] moment maps: 3x3 / 7x7 / 11x11
// zero-padded Gaussian stencils, corner/edge response, block-wide min / max, normalisation.
#include "extract_kernels.cuh"
#include "vc_launch.cuh"
#include "vc_ptx.cuh"

namespace vc {

namespace {

constexpr int kStrip = 8;

template <typename T>
__device__ __forceinline__ float4 ld4(const T* row, int v);
template <>
__device__ __forceinline__ float4 ld4<float>(const float* row, int v) {
  return __ldg(reinterpret_cast<const float4*>(row) + v);
}
template <>
__device__ __forceinline__ float4 ld4<__nv_bfloat16>(const __nv_bfloat16* row, int v) {
  const uint2 r = __ldg(reinterpret_cast<const uint2*>(row) + v);
  float4 o;
  o.x = __uint_as_float(r.x << 16);
  o.y = __uint_as_float(r.x & 0xFFFF0000u);
  o.z = __uint_as_float(r.y << 16);
  o.w = __uint_as_float(r.y & 0xFFFF0000u);
  return o;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

template <typename T, int KV>  // KV = ceil(C / 128) float4 slots per lane
__global__ void __launch_bounds__(256)
moments_kernel(const T* __restrict__ tokens, int64_t image_stride, int B, int H, int W, int C,
               float* __restrict__ moments) {
  const int lane = threadIdx.x & 31;
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  const int strips = (W + kStrip - 1) / kStrip;
  const int total = B * H * strips;
  if (warp_global >= total) return;
  const int b = warp_global / (H * strips);
  const int r = warp_global - b * H * strips;
  const int y = r / strips;
  const int x0 = (r - y * strips) * kStrip;
  const int x1 = min(W, x0 + kStrip);
  const int nvec = C >> 2;
  const T* img = tokens + (int64_t)b * image_stride;
  const size_t plane = (size_t)H * W;
  float* mom = moments + (size_t)b * 4 * plane;
  const float inv_c = 1.0f / (float)C;

  float4 cur[KV], nxt[KV];
#pragma unroll
  for (int k = 0; k < KV; ++k) {
    const int v = lane + 32 * k;
    cur[k] = v < nvec ? ld4(img + (size_t)(y * W + x0) * C, v) : make_float4(0, 0, 0, 0);
  }
  for (int x = x0; x < x1; ++x) {
    const bool has_r = x + 1 < W, has_d = y + 1 < H;
    const T* row_r = img + (size_t)(y * W + x + 1) * C;
    const T* row_d = img + (size_t)((y + 1) * W + x) * C;
    float sxx = 0.f, syy = 0.f, sxy = 0.f, sav = 0.f;
#pragma unroll
    for (int k = 0; k < KV; ++k) {
      const int v = lane + 32 * k;
      const bool ok = v < nvec;
      const float4 c = cur[k];
      float4 rr = c, dd = c;  // zero gradient at the last column / row (F.pad with 0)
      if (ok && has_r) rr = ld4(row_r, v);
      if (ok && has_d) dd = ld4(row_d, v);
      nxt[k] = rr;
      const float gx0 = rr.x - c.x, gx1 = rr.y - c.y, gx2 = rr.z - c.z, gx3 = rr.w - c.w;
      const float gy0 = dd.x - c.x, gy1 = dd.y - c.y, gy2 = dd.z - c.z, gy3 = dd.w - c.w;
      sxx += gx0 * gx0 + gx1 * gx1 + gx2 * gx2 + gx3 * gx3;
      syy += gy0 * gy0 + gy1 * gy1 + gy2 * gy2 + gy3 * gy3;
      sxy += gx0 * gy0 + gx1 * gy1 + gx2 * gy2 + gx3 * gy3;
      sav += c.x + c.y + c.z + c.w;
    }
    sxx = warp_sum(sxx);
    syy = warp_sum(syy);
    sxy = warp_sum(sxy);
    sav = warp_sum(sav);
    if (lane == 0) {
      const size_t o = (size_t)y * W + x;
      mom[o] = sxx * inv_c;
      mom[plane + o] = syy * inv_c;
      mom[2 * plane + o] = sxy * inv_c;
      mom[3 * plane + o] = sav * inv_c;
    }
#pragma unroll
    for (int k = 0; k < KV; ++k) cur[k] = nxt[k];
  }
}

__device__ __forceinline__ float block_reduce(float v, bool is_min, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float t = __shfl_xor_sync(0xffffffffu, v, o);
    v = is_min ? fminf(v, t) : fmaxf(v, t);
  }
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  const int nw = blockDim.x >> 5;
  float r = red[0];
  for (int i = 1; i < nw; ++i) r = is_min ? fminf(r, red[i]) : fmaxf(r, red[i]);
  return r;
}

// zero-padded K x K stencil with separable-product weights w[i]*w[j] (vit_extractor.py:396-402)
template <int K>
__device__ __forceinline__ float stencil(const float* map, int H, int W, int y, int x, const float* w) {
  constexpr int R = K / 2;
  float acc = 0.f;
#pragma unroll
  for (int dy = -R; dy <= R; ++dy) {
    const int yy = y + dy;
    if (yy < 0 || yy >= H) continue;
#pragma unroll
    for (int dx = -R; dx <= R; ++dx) {
      const int xx = x + dx;
      if (xx < 0 || xx >= W) continue;
      acc += map[yy * W + xx] * (w[dy + R] * w[dx + R]);
    }
  }
  return acc;
}

// min-max normalisation of vit_extractor.py:344-346: s -= min; if (max > 0) s /= max
__device__ void minmax_inplace(float* s, int n, float* red) {
  float mn = INFINITY, mx = -INFINITY;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    mn = fminf(mn, s[i]);
    mx = fmaxf(mx, s[i]);
  }
  mn = block_reduce(mn, true, red);
  mx = block_reduce(mx, false, red);
  const float top = mx - mn;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    float v = s[i] - mn;
    if (top > 0.f) v = __fdiv_rn(v, top);
    s[i] = v;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(1024, 1) response_kernel(const ScoreParams p) {
  extern __shared__ float s_avg[];  // [H*W] channel mean (DoG only)
  __shared__ float red[32];
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  const int b = blockIdx.x;
  const int n = p.H * p.W;
  const size_t plane = (size_t)n;
  const float* mom = p.moments + (size_t)b * 4 * plane;
  float* out = p.score + (size_t)b * plane;
  float* tmp = p.scratch + (size_t)b * plane;

  if (p.method == 0 || p.method == 2) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int y = i / p.W, x = i - y * p.W;
      const float ixx = stencil<3>(mom, p.H, p.W, y, x, p.g3);
      const float iyy = stencil<3>(mom + plane, p.H, p.W, y, x, p.g3);
      const float ixy = stencil<3>(mom + 2 * plane, p.H, p.W, y, x, p.g3);
      const float det = __fsub_rn(__fmul_rn(ixx, iyy), __fmul_rn(ixy, ixy));
      const float tr = __fadd_rn(ixx, iyy);
      const float corner = __fsub_rn(det, __fmul_rn(0.04f, __fmul_rn(tr, tr)));
      const float edge = sqrtf(tr);
      out[i] = __fadd_rn(__fmul_rn(0.7f, corner), __fmul_rn(0.3f, edge));
    }
    __syncthreads();
    minmax_inplace(out, n, red);
  }
  if (p.method == 1 || p.method == 2) {
    float* dst = p.method == 1 ? out : tmp;
    for (int i = threadIdx.x; i < n; i += blockDim.x) s_avg[i] = mom[3 * plane + i];
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const int y = i / p.W, x = i - y * p.W;
      const float a = stencil<7>(s_avg, p.H, p.W, y, x, p.g7);
      const float c = stencil<11>(s_avg, p.H, p.W, y, x, p.g11);
      dst[i] = fabsf(a - c);
    }
    __syncthreads();
    minmax_inplace(dst, n, red);
  }
  if (p.method == 2) {
    // vit_extractor.py:271-277: both maps are re-normalised with a +1e-8 denominator
    float hmn = INFINITY, hmx = -INFINITY, dmn = INFINITY, dmx = -INFINITY;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      hmn = fminf(hmn, out[i]);
      hmx = fmaxf(hmx, out[i]);
      dmn = fminf(dmn, tmp[i]);
      dmx = fmaxf(dmx, tmp[i]);
    }
    hmn = block_reduce(hmn, true, red);
    hmx = block_reduce(hmx, false, red);
    dmn = block_reduce(dmn, true, red);
    dmx = block_reduce(dmx, false, red);
    const float hden = __fadd_rn(__fsub_rn(hmx, hmn), 1e-8f);
    const float dden = __fadd_rn(__fsub_rn(dmx, dmn), 1e-8f);
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
      const float h = __fdiv_rn(__fsub_rn(out[i], hmn), hden);
      const float d = __fdiv_rn(__fsub_rn(tmp[i], dmn), dden);
      out[i] = __fadd_rn(__fmul_rn(0.5f, h), __fmul_rn(0.5f, d));
    }
  }
}

}  // namespace

cudaError_t launch_score_map(const ScoreParams& p, cudaStream_t stream) {
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(response_kernel, kMaxGridCells * (int)sizeof(float), opt_in); e != cudaSuccess)
    return e;
  const int strips = (p.W + kStrip - 1) / kStrip;
  const long long warps = (long long)p.B * p.H * strips;
  const int blocks = (int)((warps * 32 + 255) / 256);
  const int kv = (p.C / 4 + 31) / 32;
  cudaError_t le = cudaSuccess;
#define VC_LAUNCH_MOMENTS(KV)                                                                          \
  do {                                                                                                 \
    if (p.is_bf16)                                                                                     \
      le = launch_pdl(moments_kernel<__nv_bfloat16, KV>, dim3(blocks), dim3(256), 0, stream,           \
                      static_cast<const __nv_bfloat16*>(p.tokens), p.image_stride, p.B, p.H, p.W, p.C, p.moments); \
    else                                                                                               \
      le = launch_pdl(moments_kernel<float, KV>, dim3(blocks), dim3(256), 0, stream,                   \
                      static_cast<const float*>(p.tokens), p.image_stride, p.B, p.H, p.W, p.C, p.moments); \
  } while (0)
  if (kv <= 1) VC_LAUNCH_MOMENTS(1);
  else if (kv <= 2) VC_LAUNCH_MOMENTS(2);
  else if (kv <= 3) VC_LAUNCH_MOMENTS(3);
  else if (kv <= 4) VC_LAUNCH_MOMENTS(4);
  else if (kv <= 6) VC_LAUNCH_MOMENTS(6);
  else VC_LAUNCH_MOMENTS(8);
#undef VC_LAUNCH_MOMENTS
  if (le != cudaSuccess) return le;
  const int smem = (p.method == 0) ? 0 : p.H * p.W * (int)sizeof(float);
  return launch_pdl(response_kernel, dim3(p.B), dim3(1024), smem, stream, p);
}

}  // namespace vc
