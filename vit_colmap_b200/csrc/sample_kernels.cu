// Descriptor sampling, optional PCA projection, L2 normalisation and packing into the matcher /
// database layouts.  Reference: vit_colmap/features/vit_extractor.py
// _extract_descriptors_interp :545-586 (F.grid_sample bilinear, border, align_corners=True),
// keypoint coordinates :229-236, projection apply :651, F.normalize + uint8 quantisation :243-250.
//
// HBM-bound: per keypoint one (up to four) token rows of C elements are read with 128-bit
// coalesced loads and d bf16 + d uint8 (+ 8 B of keypoint) are written.
//   sample_pack_direct : d == C, warp per keypoint, the row never leaves registers.
//   sample_pack_project: 16 keypoints per CTA are staged in shared memory (channel-major), the
//                        [C, d] projection streams through L2 once per CTA, fp32 SIMT FMAs.
#include "extract_kernels.cuh"
#include "vc_launch.cuh"
#include "vc_ptx.cuh"

namespace vc {

namespace {


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

// grid_sample source index of one axis: normalise as the reference does (2 v / (n-1) - 1), then
// PyTorch's align_corners=True un-normalisation and border clipping, all in rounded fp32 steps.
__device__ __forceinline__ float source_index(float v, int n) {
  const float nm1 = (float)(n - 1);
  const float c = __fsub_rn(__fdiv_rn(__fmul_rn(2.0f, v), nm1), 1.0f);
  const float u = __fmul_rn(__fdiv_rn(__fadd_rn(c, 1.0f), 2.0f), nm1);
  return fminf(nm1, fmaxf(u, 0.0f));
}

struct Taps {
  int idx[4];   // token index of nw, ne, sw, se (or -1 when outside / zero weight)
  float w[4];
};

__device__ __forceinline__ Taps make_taps(float fy, float fx, int H, int W) {
  const float iy = source_index(fy, H), ix = source_index(fx, W);
  const float y0 = floorf(iy), x0 = floorf(ix);
  const float y1 = y0 + 1.f, x1 = x0 + 1.f;
  // weights as ATen's CPU kernel forms them: w = x - floor(x), e = 1 - w, n = y - floor(y), s = 1 - n
  const float ww = __fsub_rn(ix, x0), ee = __fsub_rn(1.0f, ww);
  const float nn = __fsub_rn(iy, y0), ss = __fsub_rn(1.0f, nn);
  Taps t;
  t.w[0] = __fmul_rn(ee, ss);
  t.w[1] = __fmul_rn(ww, ss);
  t.w[2] = __fmul_rn(ee, nn);
  t.w[3] = __fmul_rn(ww, nn);
  const int yi[2] = {(int)y0, (int)y1}, xi[2] = {(int)x0, (int)x1};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int yy = yi[k >> 1], xx = xi[k & 1];
    const bool in = yy >= 0 && yy < H && xx >= 0 && xx < W && t.w[k] != 0.f;
    t.idx[k] = in ? yy * W + xx : -1;
  }
  return t;
}

__device__ __forceinline__ void read_coord(const SampleParams& p, size_t kp, float& fy, float& fx) {
  if (p.coords_f32) {
    const float2 c = reinterpret_cast<const float2*>(p.coords)[kp];
    fy = c.x;
    fx = c.y;
  } else {
    const int2 c = reinterpret_cast<const int2*>(p.coords)[kp];
    fy = (float)c.x;
    fx = (float)c.y;
  }
}

__device__ __forceinline__ uint8_t quantise(float v, int mode) {
  float q;
  if (mode == 0) {         // vit_extractor.py:250  trunc(clip(512 v, 0, 255))
    q = fminf(fmaxf(__fmul_rn(v, 512.0f), 0.f), 255.f);
  } else if (mode == 1) {  // COLMAP/feature/utils.cc:56-71  min(255, round(512 v))
    q = fminf(fmaxf(roundf(__fmul_rn(v, 512.0f)), 0.f), 255.f);
  } else {                 // trainable_vit_extractor.py:265-267  trunc(clip((v + 1) 127.5, 0, 255))
    q = fminf(fmaxf(__fmul_rn(__fadd_rn(v, 1.0f), 127.5f), 0.f), 255.f);
  }
  return (uint8_t)(int)q;
}

// RootSIFT pre-step of hybrid_extractor.py:286-287: F.normalize(p=1) (eps 1e-12) then
// sqrt(clamp(min=1e-8)); the L2 step that follows is the common one.
__device__ __forceinline__ float rootsift(float v, float l1_den) {
  return sqrtf(fmaxf(__fdiv_rn(v, l1_den), 1e-8f));
}

__device__ __forceinline__ void write_keypoint(const SampleParams& p, size_t kp, float fy, float fx) {
  if (p.keypoints == nullptr) return;
  float2 o;
  o.x = __fmul_rn(__fmul_rn(__fadd_rn(fx, 0.5f), p.sx1), p.sx2);
  o.y = __fmul_rn(__fmul_rn(__fadd_rn(fy, 0.5f), p.sy1), p.sy2);
  reinterpret_cast<float2*>(p.keypoints)[kp] = o;
}

__device__ __forceinline__ void store4(const SampleParams& p, size_t row, int v, float4 o) {
  const size_t e = row * p.d + 4 * (size_t)v;
  if (p.desc_f32) *reinterpret_cast<float4*>(p.desc_f32 + e) = o;
  if (p.desc_bf16) {
    __nv_bfloat162 a = __floats2bfloat162_rn(o.x, o.y), c = __floats2bfloat162_rn(o.z, o.w);
    uint2 pk;
    pk.x = *reinterpret_cast<unsigned*>(&a);
    pk.y = *reinterpret_cast<unsigned*>(&c);
    *reinterpret_cast<uint2*>(p.desc_bf16 + e) = pk;
  }
  if (p.desc_u8) {
    uchar4 q;
    q.x = quantise(o.x, p.quant_mode);
    q.y = quantise(o.y, p.quant_mode);
    q.z = quantise(o.z, p.quant_mode);
    q.w = quantise(o.w, p.quant_mode);
    *reinterpret_cast<uchar4*>(p.desc_u8 + e) = q;
  }
}

// Rows at or beyond counts[b] are zero-filled here, so the caller can hand over uninitialised
// output buffers (saves a full extra write pass over [B, cap, d] x 3 outputs).
__device__ __forceinline__ void zero_row(const SampleParams& p, size_t kp, int lane) {
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int v = lane; v < (p.d >> 2); v += 32) {
    const size_t e = kp * p.d + 4 * (size_t)v;
    if (p.desc_f32) *reinterpret_cast<float4*>(p.desc_f32 + e) = z;
    if (p.desc_bf16) *reinterpret_cast<uint2*>(p.desc_bf16 + e) = make_uint2(0u, 0u);
    if (p.desc_u8) *reinterpret_cast<uchar4*>(p.desc_u8 + e) = make_uchar4(0, 0, 0, 0);
  }
  if (lane == 0 && p.keypoints) reinterpret_cast<float2*>(p.keypoints)[kp] = make_float2(0.f, 0.f);
}

// KV = float4 slots per lane = ceil(C / 128): sized to the channel count so that no dead accumulator
// registers limit the occupancy of this bandwidth-bound kernel.
template <typename T, int KV>
__global__ void __launch_bounds__(256) sample_pack_direct(const SampleParams p) {
  const int lane = threadIdx.x & 31;
  const long long wg = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int b = (int)(wg / p.cap), k = (int)(wg - (long long)b * p.cap);
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  if (b >= p.B) return;
  const size_t kp = (size_t)b * p.cap + k;
  if (k >= min(p.counts[b], p.cap)) {
    zero_row(p, kp, lane);
    return;
  }
  float fy, fx;
  read_coord(p, kp, fy, fx);
  const Taps t = make_taps(fy, fx, p.H, p.W);
  const T* img = static_cast<const T*>(p.tokens) + (int64_t)b * p.image_stride;
  const int nvec = p.C >> 2;
  float4 acc[KV];
#pragma unroll
  for (int j = 0; j < KV; ++j) acc[j] = make_float4(0, 0, 0, 0);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    if (t.idx[s] < 0) continue;  // warp-uniform
    const T* row = img + (size_t)t.idx[s] * p.C;
    const float w = t.w[s];
#pragma unroll
    for (int j = 0; j < KV; ++j) {
      const int v = lane + 32 * j;
      if (v < nvec) {
        const float4 x = ld4(row, v);
        acc[j].x = __fadd_rn(acc[j].x, __fmul_rn(x.x, w));
        acc[j].y = __fadd_rn(acc[j].y, __fmul_rn(x.y, w));
        acc[j].z = __fadd_rn(acc[j].z, __fmul_rn(x.z, w));
        acc[j].w = __fadd_rn(acc[j].w, __fmul_rn(x.w, w));
      }
    }
  }
  if (p.rootsift) {
    float l1 = 0.f;
#pragma unroll
    for (int j = 0; j < KV; ++j)
      if (lane + 32 * j < nvec) l1 += fabsf(acc[j].x) + fabsf(acc[j].y) + fabsf(acc[j].z) + fabsf(acc[j].w);
    const float den1 = fmaxf(warp_sum(l1), 1e-12f);
#pragma unroll
    for (int j = 0; j < KV; ++j) {
      if (lane + 32 * j < nvec) {
        acc[j].x = rootsift(acc[j].x, den1);
        acc[j].y = rootsift(acc[j].y, den1);
        acc[j].z = rootsift(acc[j].z, den1);
        acc[j].w = rootsift(acc[j].w, den1);
      }
    }
  }
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < KV; ++j)
    ss += acc[j].x * acc[j].x + acc[j].y * acc[j].y + acc[j].z * acc[j].z + acc[j].w * acc[j].w;
  ss = warp_sum(ss);
  const float den = fmaxf(sqrtf(ss), p.norm_eps);  // F.normalize eps
#pragma unroll
  for (int j = 0; j < KV; ++j) {
    const int v = lane + 32 * j;
    if (v < nvec) {
      float4 o;
      o.x = __fdiv_rn(acc[j].x, den);
      o.y = __fdiv_rn(acc[j].y, den);
      o.z = __fdiv_rn(acc[j].z, den);
      o.w = __fdiv_rn(acc[j].w, den);
      store4(p, kp, v, o);
    }
  }
  if (lane == 0) write_keypoint(p, kp, fy, fx);
}

constexpr int kTile = 16;  // keypoints per CTA in the projecting kernel

template <typename T>
__global__ void __launch_bounds__(256) sample_pack_project(const SampleParams p, int tiles_per_image) {
  extern __shared__ __align__(16) float sm[];
  float* vt = sm;                       // [C][kTile] sampled rows, channel-major
  float* outp = sm + (size_t)p.C * kTile;  // [kTile][d] projected rows
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  const int b = blockIdx.x / tiles_per_image;
  const int k0 = (blockIdx.x - b * tiles_per_image) * kTile;
  const int cnt = min(p.counts[b], p.cap);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nk = max(0, min(kTile, cnt - k0));
  for (int kk = nk + warp; kk < kTile && k0 + kk < p.cap; kk += 8) zero_row(p, (size_t)b * p.cap + k0 + kk, lane);
  if (nk == 0) return;
  const T* img = static_cast<const T*>(p.tokens) + (int64_t)b * p.image_stride;
  const int nvec = p.C >> 2;

  // phase 1: gather (each warp samples 2 keypoints)
  for (int kk = warp; kk < kTile; kk += 8) {
    float fy = 0.f, fx = 0.f;
    Taps t;
    const bool live = kk < nk;
    if (live) {
      read_coord(p, (size_t)b * p.cap + k0 + kk, fy, fx);
      t = make_taps(fy, fx, p.H, p.W);
      if (lane == 0) write_keypoint(p, (size_t)b * p.cap + k0 + kk, fy, fx);
    }
    for (int v = lane; v < nvec; v += 32) {
      float4 a = make_float4(0, 0, 0, 0);
      if (live) {
#pragma unroll
        for (int s = 0; s < 4; ++s) {
          if (t.idx[s] < 0) continue;
          const float4 x = ld4(img + (size_t)t.idx[s] * p.C, v);
          const float w = t.w[s];
          a.x = __fadd_rn(a.x, __fmul_rn(x.x, w));
          a.y = __fadd_rn(a.y, __fmul_rn(x.y, w));
          a.z = __fadd_rn(a.z, __fmul_rn(x.z, w));
          a.w = __fadd_rn(a.w, __fmul_rn(x.w, w));
        }
      }
      vt[(4 * v + 0) * kTile + kk] = a.x;
      vt[(4 * v + 1) * kTile + kk] = a.y;
      vt[(4 * v + 2) * kTile + kk] = a.z;
      vt[(4 * v + 3) * kTile + kk] = a.w;
    }
  }
  __syncthreads();

  // phase 2: out[k][j] = sum_c v[k][c] * P[c][j]; thread = (8-keypoint group, column)
  const int grp = threadIdx.x >> 7, jj = threadIdx.x & 127;
  for (int j0 = 0; j0 < p.d; j0 += 128) {
    const int j = j0 + jj;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    if (j < p.d) {
      const float* P = p.projection + j;
#pragma unroll 4
      for (int c = 0; c < p.C; ++c) {
        const float w = __ldg(P + (size_t)c * p.d);
        const float4 a = *reinterpret_cast<const float4*>(vt + c * kTile + grp * 8);
        const float4 e = *reinterpret_cast<const float4*>(vt + c * kTile + grp * 8 + 4);
        acc[0] = fmaf(a.x, w, acc[0]);
        acc[1] = fmaf(a.y, w, acc[1]);
        acc[2] = fmaf(a.z, w, acc[2]);
        acc[3] = fmaf(a.w, w, acc[3]);
        acc[4] = fmaf(e.x, w, acc[4]);
        acc[5] = fmaf(e.y, w, acc[5]);
        acc[6] = fmaf(e.z, w, acc[6]);
        acc[7] = fmaf(e.w, w, acc[7]);
      }
      const float bj = p.proj_bias ? __ldg(p.proj_bias + j) : 0.f;
#pragma unroll
      for (int i = 0; i < 8; ++i) outp[(size_t)(grp * 8 + i) * p.d + j] = acc[i] + bj;
    }
  }
  __syncthreads();

  // phase 3: normalise + pack (each warp finishes 2 keypoints)
  const int dvec = p.d >> 2;
  for (int kk = warp; kk < nk; kk += 8) {
    float* row = outp + (size_t)kk * p.d;
    if (p.rootsift) {  // each lane rewrites only the elements it reads back below
      float l1 = 0.f;
      for (int v = lane; v < dvec; v += 32) {
        const float4 x = *reinterpret_cast<const float4*>(row + 4 * v);
        l1 += fabsf(x.x) + fabsf(x.y) + fabsf(x.z) + fabsf(x.w);
      }
      const float den1 = fmaxf(warp_sum(l1), 1e-12f);
      for (int v = lane; v < dvec; v += 32) {
        float4 x = *reinterpret_cast<const float4*>(row + 4 * v);
        x.x = rootsift(x.x, den1);
        x.y = rootsift(x.y, den1);
        x.z = rootsift(x.z, den1);
        x.w = rootsift(x.w, den1);
        *reinterpret_cast<float4*>(row + 4 * v) = x;
      }
    }
    float ss = 0.f;
    for (int v = lane; v < dvec; v += 32) {
      const float4 x = *reinterpret_cast<const float4*>(row + 4 * v);
      ss += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
    }
    ss = warp_sum(ss);
    const float den = fmaxf(sqrtf(ss), p.norm_eps);
    for (int v = lane; v < dvec; v += 32) {
      const float4 x = *reinterpret_cast<const float4*>(row + 4 * v);
      float4 o;
      o.x = __fdiv_rn(x.x, den);
      o.y = __fdiv_rn(x.y, den);
      o.z = __fdiv_rn(x.z, den);
      o.w = __fdiv_rn(x.w, den);
      store4(p, (size_t)b * p.cap + k0 + kk, v, o);
    }
  }
}

}  // namespace

cudaError_t launch_sample_pack(const SampleParams& p, cudaStream_t stream) {
  if (p.projection == nullptr) {
    const long long warps = (long long)p.B * p.cap;
    const int blocks = (int)((warps * 32 + 255) / 256);
    if (blocks == 0) return cudaSuccess;
    const int kv = (p.C / 4 + 31) / 32;
    cudaError_t le = cudaSuccess;
#define VC_LAUNCH_DIRECT(KV)                                                    \
  do {                                                                          \
    if (p.is_bf16)                                                              \
      le = launch_pdl(sample_pack_direct<__nv_bfloat16, KV>, dim3(blocks), dim3(256), 0, stream, p); \
    else                                                                        \
      le = launch_pdl(sample_pack_direct<float, KV>, dim3(blocks), dim3(256), 0, stream, p); \
  } while (0)
    if (kv <= 1) VC_LAUNCH_DIRECT(1);
    else if (kv <= 2) VC_LAUNCH_DIRECT(2);
    else if (kv <= 3) VC_LAUNCH_DIRECT(3);
    else if (kv <= 4) VC_LAUNCH_DIRECT(4);
    else if (kv <= 6) VC_LAUNCH_DIRECT(6);
    else VC_LAUNCH_DIRECT(8);
#undef VC_LAUNCH_DIRECT
    return le;
  }
  const int smem = (p.C * kTile + kTile * p.d) * (int)sizeof(float);
  static SmemOptIn opt_in[2];
  if (cudaError_t e = p.is_bf16 ? ensure_dynamic_smem(sample_pack_project<__nv_bfloat16>, smem, opt_in[1])
                                : ensure_dynamic_smem(sample_pack_project<float>, smem, opt_in[0]);
      e != cudaSuccess)
    return e;
  const int tiles = (p.cap + kTile - 1) / kTile;
  if (tiles == 0 || p.B == 0) return cudaSuccess;
  if (p.is_bf16)
    return launch_pdl(sample_pack_project<__nv_bfloat16>, dim3(p.B * tiles), dim3(256), smem, stream, p, tiles);
  return launch_pdl(sample_pack_project<float>, dim3(p.B * tiles), dim3(256), smem, stream, p, tiles);
}

}  // namespace vc
