// Per-tile epilogue of the ViT tcgen05 GEMMs, shared by the one-SM kernel (vit_gemm.cu) and the two-SM
// kernel (vit_gemm2.cu): one epilogue warp drains its 32 rows x 128 columns of a 128 x 256 fp32 accumulator
// tile.  tcgen05.ld hands thread t row t of the warp's 32 x 128-byte chunk; the chunk is transposed through a
// swizzled shared-memory buffer so that every global access of a warp covers 4 whole 128-byte lines.
#pragma once
#include "vc_epilogue.cuh"
#include "vc_ptx.cuh"
#include "vit_kernels.cuh"

namespace vc {

constexpr int kGemmBM = 128, kGemmBN = 256;
constexpr bool gemm_is_residual(int epi) {
  return epi == kEpiResidualF32 || epi == kEpiResidualLnF32 || epi == kEpiResidualPublishF32 ||
         epi == kEpiResidualStatsF32;
}
constexpr bool gemm_is_bf16_out(int epi) {
  return epi == kEpiBiasBf16 || epi == kEpiBiasGeluBf16 || epi == kEpiFoldBiasBf16 || epi == kEpiFoldBiasGeluBf16;
}
constexpr bool gemm_is_fold(int epi) { return epi == kEpiFoldBiasBf16 || epi == kEpiFoldBiasGeluBf16; }
constexpr bool gemm_has_gelu(int epi) { return epi == kEpiBiasGeluBf16 || epi == kEpiFoldBiasGeluBf16; }

// Folded LayerNorm (kEpiFoldBias*Bf16): mean and 1 / std of row m from the producing GEMM's per-slice sums, added
// in slice order.  Callers ask for it BEFORE they wait for the accumulator, so the L2 latency is hidden.
__device__ __forceinline__ void gemm_row_stats(const GemmParams& p, int m, float& mean, float& rstd) {
  mean = 0.f;
  rstd = 1.f;
  if (m >= p.M) return;
  const float2* part = reinterpret_cast<const float2*>(p.ln_part) + (size_t)m * p.ln_slices;
  float s1 = 0.f, s2 = 0.f;
  for (int j = 0; j < p.ln_slices; ++j) {
    const float2 v = __ldcg(part + j);
    s1 += v.x;
    s2 += v.y;
  }
  const float inv_c = 1.0f / (float)p.K;
  mean = s1 * inv_c;
  rstd = rsqrtf(fmaxf(s2 * inv_c - mean * mean, 0.f) + 1e-6f);
}

// t_addr: TMEM address of the warp's first column (lane quarter q, column half h of the tile's accumulator).
// tm / tn: 128-row block and 256-column block of D.  stg: this warp's 32 x 128 B transposition buffer.
// release(): called once, as soon as the accumulator has been read (hands the TMEM buffer back).
template <int EPI, typename Release>
__device__ __forceinline__ void gemm_epilogue_tile(const GemmParams& p, uint32_t t_addr, int tm, int tn, int q, int h,
                                                   int lane, uint8_t* stg, Release release, float ln_mean = 0.f,
                                                   float ln_rstd = 1.f) {
  constexpr int BM = kGemmBM, BN = kGemmBN;
  uint8_t* stg_row = stg + lane * 128;          // row layout: thread = row
  const int tc = lane & 7, tr = lane >> 3;       // transposed layout: 8 lanes per row, rows tr + 4 i
  const int row0 = tm * BM + q * 32;
  bool released = false;
  if (gemm_is_bf16_out(EPI)) {

    // bf16 outputs: 64 columns (= 128 B per row) per round
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      const int n0 = tn * BN + h * 128 + c * 64;
      if (n0 >= p.N) break;  // warp-uniform
      const bool last = (c == 1) || (n0 + 64 >= p.N);
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int nh = n0 + half * 32;
        uint32_t vr[32];
        if (nh < p.N) {  // warp-uniform (N % 32 == 0)
          tmem_ld_32x32(t_addr + c * 64 + half * 32, vr);
          tmem_wait_ld();
        }
        if (half == 1 && last) {  // accumulator fully read: hand the TMEM buffer back early
          release();
          released = true;
        }
        if (nh >= p.N) break;
        float v[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(vr[e]);
        const float4* b4 = reinterpret_cast<const float4*>(p.bias + nh);
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const float4 b = __ldg(b4 + g);
          if (gemm_is_fold(EPI)) {
            const float4 cs = __ldg(reinterpret_cast<const float4*>(p.colsum + nh) + g);
            v[4 * g + 0] = fmaf(ln_rstd, fmaf(-ln_mean, cs.x, v[4 * g + 0]), b.x);
            v[4 * g + 1] = fmaf(ln_rstd, fmaf(-ln_mean, cs.y, v[4 * g + 1]), b.y);
            v[4 * g + 2] = fmaf(ln_rstd, fmaf(-ln_mean, cs.z, v[4 * g + 2]), b.z);
            v[4 * g + 3] = fmaf(ln_rstd, fmaf(-ln_mean, cs.w, v[4 * g + 3]), b.w);
          } else {
            v[4 * g + 0] += b.x;
            v[4 * g + 1] += b.y;
            v[4 * g + 2] += b.z;
            v[4 * g + 3] += b.w;
          }
        }
        if (gemm_has_gelu(EPI)) {
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            const float2 g = gelu_erf2(make_float2(v[e], v[e + 1]));
            v[e] = g.x;
            v[e + 1] = g.y;
          }
        }
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          uint4 w;
          w.x = pack_bf16(v[8 * g + 0], v[8 * g + 1]);
          w.y = pack_bf16(v[8 * g + 2], v[8 * g + 3]);
          w.z = pack_bf16(v[8 * g + 4], v[8 * g + 5]);
          w.w = pack_bf16(v[8 * g + 6], v[8 * g + 7]);
          *reinterpret_cast<uint4*>(stg_row + (((half * 4 + g) ^ (lane & 7)) << 4)) = w;
        }
      }
      __syncwarp();
      const int col = n0 + tc * 8;
      if (col < p.N) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int r = tr + 4 * i, m = row0 + r;
          if (m < p.M) {
            const uint4 w = *reinterpret_cast<const uint4*>(stg + r * 128 + ((tc ^ (r & 7)) << 4));
            *reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(p.out) + (size_t)m * p.ldo + col) = w;
          }
        }
      }
      __syncwarp();
    }
  } else {
    // fp32 outputs (residual update / patch embedding): 32 columns (= 128 B per row) per round;
    // bias, LayerScale and position embedding are applied in the transposed layout, where a
    // lane's 4 columns are fixed
    float st1[8], st2[8];  // kEpiResidualStatsF32: sums of x and x^2 of this lane's columns, rows tr + 4 i
    if (EPI == kEpiResidualStatsF32) {
#pragma unroll
      for (int i = 0; i < 8; ++i) st1[i] = st2[i] = 0.f;
    }
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      const int n0 = tn * BN + h * 128 + c * 32;
      if (n0 >= p.N) break;  // warp-uniform
      const bool last = (c == 3) || (n0 + 32 >= p.N);
      const int col = n0 + tc * 4;
      // residual rows are requested first, so their HBM latency overlaps the TMEM read and the
      // transposition (loads and stores are kept in separate loops: no load waits on a store)
      float4 x[8];
      if (gemm_is_residual(EPI)) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int m = row0 + tr + 4 * i;
          if (m < p.M) x[i] = *reinterpret_cast<const float4*>(static_cast<const float*>(p.out) + (size_t)m * p.ldo + col);
        }
      }
      uint32_t vr[32];
      tmem_ld_32x32(t_addr + c * 32, vr);
      tmem_wait_ld();
      if (last) {
        release();
        released = true;
      }
#pragma unroll
      for (int g = 0; g < 8; ++g)
        *reinterpret_cast<uint4*>(stg_row + ((g ^ (lane & 7)) << 4)) =
            make_uint4(vr[4 * g], vr[4 * g + 1], vr[4 * g + 2], vr[4 * g + 3]);
      __syncwarp();
      const float4 b = __ldg(reinterpret_cast<const float4*>(p.bias + col));
      float4 ga = make_float4(1.f, 1.f, 1.f, 1.f);
      if (gemm_is_residual(EPI)) ga = __ldg(reinterpret_cast<const float4*>(p.gamma + col));
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = tr + 4 * i, m = row0 + r;
        if (m < p.M) {
          const float4 a = *reinterpret_cast<const float4*>(stg + r * 128 + ((tc ^ (r & 7)) << 4));
          if (gemm_is_residual(EPI)) {
            float4 y = x[i];
            y.x = fmaf(ga.x, a.x + b.x, y.x);
            y.y = fmaf(ga.y, a.y + b.y, y.y);
            y.z = fmaf(ga.z, a.z + b.z, y.z);
            y.w = fmaf(ga.w, a.w + b.w, y.w);
            *reinterpret_cast<float4*>(static_cast<float*>(p.out) + (size_t)m * p.ldo + col) = y;
            if (EPI == kEpiResidualStatsF32) {
              st1[i] += (y.x + y.y) + (y.z + y.w);
              st2[i] = fmaf(y.x, y.x, fmaf(y.y, y.y, fmaf(y.z, y.z, fmaf(y.w, y.w, st2[i]))));
              *reinterpret_cast<uint2*>(p.xb_out + (size_t)m * p.N + col) = make_uint2(pack_bf16(y.x, y.y), pack_bf16(y.z, y.w));
            }
          } else {
            const int img = m / p.tokens_per_image, t = m - img * p.tokens_per_image;
            const size_t out_row = (size_t)img * (p.tokens_per_image + 1) + 1 + t;
            const float4 pe = __ldg(reinterpret_cast<const float4*>(p.pos + (size_t)(1 + t) * p.N + col));
            *reinterpret_cast<float4*>(static_cast<float*>(p.out) + out_row * p.ldo + col) =
                make_float4(a.x + b.x + pe.x, a.y + b.y + pe.y, a.z + b.z + pe.z, a.w + b.w + pe.w);
          }
        }
      }
      __syncwarp();
    }
    if (EPI == kEpiResidualStatsF32) {
      // the eight lanes of a row (tc = 0 .. 7) add up in a fixed order; lane tc == 0 stores the slice's pair
      const int slice = tn * (BN / 128) + h;
      if (slice * 128 < p.N) {  // warp-uniform
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          float a = st1[i], b = st2[i];
#pragma unroll
          for (int o = 1; o < 8; o <<= 1) {
            a += __shfl_xor_sync(0xffffffffu, a, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
          }
          const int m = row0 + tr + 4 * i;
          if (tc == 0 && m < p.M)
            reinterpret_cast<float2*>(p.part_out)[(size_t)m * (p.N >> 7) + slice] = make_float2(a, b);
        }
      }
    }
  }
  if (!released) release();  // this warp's columns lie entirely beyond N
}

}  // namespace vc
