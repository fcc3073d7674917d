// Per-tile epilogue of the implicit-GEMM convolution kernels, shared by the one-SM kernel (conv_gemm.cu) and the
// two-SM kernel (conv_gemm2.cu): one epilogue warp drains its 32 pixels x 128 channels of a 128-pixel x 256-channel
// fp32 accumulator tile (BatchNorm scale / shift, exact-erf GELU, bias, bf16 or fp32 stores, sub-pixel scatter of the
// transposed convolution).  See conv_gemm.cu for the layout notes.
#pragma once
#include "head_kernels.cuh"
#include "vc_epilogue.cuh"
#include "vc_ptx.cuh"

namespace vc {

constexpr int kConvBM = 128, kConvBN = 256;

struct ConvTileCoord {
  int b, y0, x0, ph, n0;
  bool valid;  // false: a pair's second patch beyond the last one (two-SM kernel): nothing is stored
};

// patch: index of the 8 x 16 pixel patch over (image, patch row, patch column); tn: 256-channel block
__device__ __forceinline__ ConvTileCoord conv_tile_coord(const ConvParams& p, int patch, int ph, int tn, int tiles_x,
                                                         int tiles_per_image) {
  ConvTileCoord t;
  t.ph = ph;
  t.b = patch / tiles_per_image;
  const int r = patch - t.b * tiles_per_image;
  const int ty = r / tiles_x;
  t.y0 = ty * kConvTileH;
  t.x0 = (r - ty * tiles_x) * kConvTileW;
  t.n0 = tn * kConvBN;
  t.valid = t.b < p.B;
  return t;
}

// Output pixels of the 8 rows a lane stores (rows tr + 4 i of the warp's 32); computed before the wait for the
// accumulator.
struct ConvLanePixels {
  size_t out_pix[8];
  bool ok[8];
};
__device__ __forceinline__ ConvLanePixels conv_lane_pixels(const ConvParams& p, const ConvTileCoord& t, int q, int lane) {
  ConvLanePixels px;
  const int tr = lane >> 3;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = q * 32 + tr + 4 * i;
    const int y = t.y0 + r / kConvTileW, x = t.x0 + r % kConvTileW;
    px.ok[i] = t.valid && y < p.H && x < p.W;
    px.out_pix[i] = ((size_t)t.b * p.Ho + (size_t)(y * p.sy + p.oy[t.ph])) * p.Wo + (size_t)(x * p.sx + p.ox[t.ph]);
  }
  return px;
}

// t_addr: TMEM address of the warp's first column; stg: the warp's 32 x 128 B transposition buffer;
// release(): called once, as soon as the accumulator has been read.
template <int EPI, typename Release>
__device__ __forceinline__ void conv_epilogue_tile(const ConvParams& p, const ConvTileCoord& t, const ConvLanePixels& px,
                                                   uint32_t t_addr, int h, int lane, uint8_t* stg, Release release) {
  uint8_t* stg_row = stg + lane * 128;
  const int tc = lane & 7, tr = lane >> 3;
  const size_t* out_pix = px.out_pix;
  const bool* ok = px.ok;
  bool released = false;
  if (EPI != kConvBiasF32) {
    // bf16 outputs: 64 channels (= 128 B per pixel) per round
#pragma unroll 1
    for (int c = 0; c < 2; ++c) {
      const int n0 = t.n0 + h * 128 + c * 64;
      if (n0 >= p.Cout) break;  // warp-uniform
      const bool last = (c == 1) || (n0 + 64 >= p.Cout);
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        const int nh = n0 + half * 32;
        uint32_t vr[32];
        if (nh < p.Cout) {  // warp-uniform (Cout % 32 == 0)
          tmem_ld_32x32(t_addr + c * 64 + half * 32, vr);
          tmem_wait_ld();
        }
        if (half == 1 && last) {  // accumulator fully read: hand the TMEM buffer back early
          release();
          released = true;
        }
        if (nh >= p.Cout) break;
        float v[32];
#pragma unroll
        for (int e = 0; e < 32; ++e) v[e] = __uint_as_float(vr[e]);
        const float4* b4 = reinterpret_cast<const float4*>(p.shift + nh);
        if (EPI == kConvBnGeluBf16) {
          const float4* s4 = reinterpret_cast<const float4*>(p.scale + nh);
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const float4 sc = __ldg(s4 + g), b = __ldg(b4 + g);
            v[4 * g + 0] = fmaf(v[4 * g + 0], sc.x, b.x);
            v[4 * g + 1] = fmaf(v[4 * g + 1], sc.y, b.y);
            v[4 * g + 2] = fmaf(v[4 * g + 2], sc.z, b.z);
            v[4 * g + 3] = fmaf(v[4 * g + 3], sc.w, b.w);
          }
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            const float2 g = gelu_erf2(make_float2(v[e], v[e + 1]));
            v[e] = g.x;
            v[e + 1] = g.y;
          }
        } else {
#pragma unroll
          for (int g = 0; g < 8; ++g) {
            const float4 b = __ldg(b4 + g);
            v[4 * g + 0] += b.x;
            v[4 * g + 1] += b.y;
            v[4 * g + 2] += b.z;
            v[4 * g + 3] += b.w;
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
      int col = n0 + tc * 8;
      if (col < p.Cout) {
        __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(p.out);
        int ld = p.ldo;
        if (p.split_n > 0 && col >= p.split_n) {
          dst = static_cast<__nv_bfloat16*>(p.out2);
          ld = p.ldo2;
          col -= p.split_n;
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (ok[i]) {
            const int r = tr + 4 * i;
            const uint4 w = *reinterpret_cast<const uint4*>(stg + r * 128 + ((tc ^ (r & 7)) << 4));
            *reinterpret_cast<uint4*>(dst + out_pix[i] * ld + col) = w;
          }
        }
      }
      __syncwarp();
    }
  } else {
    // fp32 outputs (1x1 projection): 32 channels (= 128 B per pixel) per round, bias added in the
    // transposed layout where a lane's 4 channels are fixed
#pragma unroll 1
    for (int c = 0; c < 4; ++c) {
      const int n0 = t.n0 + h * 128 + c * 32;
      if (n0 >= p.Cout) break;  // warp-uniform
      const bool last = (c == 3) || (n0 + 32 >= p.Cout);
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
      int col = n0 + tc * 4;
      const float4 b = __ldg(reinterpret_cast<const float4*>(p.shift + col));
      float* dst = static_cast<float*>(p.out);
      int ld = p.ldo;
      if (p.split_n > 0 && col >= p.split_n) {
        dst = static_cast<float*>(p.out2);
        ld = p.ldo2;
        col -= p.split_n;
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (ok[i]) {
          const int r = tr + 4 * i;
          const float4 a = *reinterpret_cast<const float4*>(stg + r * 128 + ((tc ^ (r & 7)) << 4));
          *reinterpret_cast<float4*>(dst + out_pix[i] * ld + col) =
              make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w);
        }
      }
      __syncwarp();
    }
  }
  if (!released) release();  // this warp's channels lie entirely beyond Cout
}

}  // namespace vc
