// Implicit-GEMM convolution on tcgen05 for the ViTFeatureModel head
// (vit_colmap/model/vit_feature_model.py:12-29,90-121: ConvTranspose2d k4 s2 p1, Conv3x3 + BatchNorm
// + GELU, 1x1 projections).  Activations are channels-last bf16 [B, H, W, C].
//   - GEMM M tile = an 8 x 16 pixel patch of one image, fetched per filter tap as ONE 4-D TMA box
//     {64 channels, 16, 8, 1} at the tap's (dy, dx) offset: out-of-image rows/columns arrive as
//     zeros, which is the convolution's zero padding, so no im2col buffer and no halo logic exist.
//     The box lands as 128 rows x 128 B with the 128-byte swizzle = the K-major UMMA A operand.
//   - K loop = taps x (Cin / 64); B operand = weight rows [Cout, taps*Cin] (tap-major K), 2-D TMA.
//   - the transposed convolution runs as 4 sub-pixel phases (2x2 taps each) of the same kernel;
//     the epilogue scatters phase (py, px) to out[2y+py, 2x+px].
//   - persistent, one CTA per SM, 4-stage TMA ring, fp32 accumulators double-buffered in TMEM,
//     8 epilogue warps; UMMA N is min(256, Cout - n0) so Cout = 192 / 128 / 64 waste no MMA work.
//   - epilogues: +bias -> bf16; eval-BatchNorm scale/shift + exact-erf GELU -> bf16 (optionally
//     split over two output tensors: merged keypoint/descriptor head conv); +bias -> f32.
// The per-tile epilogue lives in conv_gemm_epilogue.cuh and is shared with the two-SM kernel (conv_gemm2.cu, the
// default); this kernel runs under VC_CONV_2CTA=0.
#include "head_kernels.cuh"

#include "conv_gemm_epilogue.cuh"
#include "vc_epilogue.cuh"
#include "vc_ptx.cuh"
#include "vc_launch.cuh"

namespace vc {

namespace {

constexpr int BM = 128, BN = 256, BK = 64, kStages = 4;
constexpr int kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kThreads = 384;
constexpr int kTmemCols = 512;
constexpr int kStageOutBytes = 32 * 128;  // per epilogue warp: 32 rows x 128 B transposition buffer
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256 + 8 * kStageOutBytes;
static_assert(kConvTileH * kConvTileW == BM && BM == kConvBM && BN == kConvBN, "M tile is one 8x16 pixel patch");

using TileCoord = ConvTileCoord;

// tile -> (256-channel block fastest, then sub-pixel phase, then patch)
__device__ __forceinline__ TileCoord decode_tile(int tile, const ConvParams& p, int tiles_n, int tiles_x,
                                                 int tiles_per_image) {
  const int tn = tile % tiles_n;
  const int r = tile / tiles_n;
  return conv_tile_coord(p, r / p.phases, r % p.phases, tn, tiles_x, tiles_per_image);
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
conv_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                 const __grid_constant__ ConvParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const uint32_t sA = base;
  const uint32_t sB = base + kStages * kABytes;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kStages + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * kStages + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * kStages + 2 + s); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * kStages + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cpb = p.Cin / BK;
  const int nkb = p.taps * cpb;
  const int tiles_n = (p.Cout + BN - 1) / BN;
  const int tiles_x = (p.W + kConvTileW - 1) / kConvTileW;
  const int tiles_per_image = ((p.H + kConvTileH - 1) / kConvTileH) * tiles_x;
  const int n_tiles = p.B * tiles_per_image * p.phases * tiles_n;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_a);
    tma_prefetch_desc(&map_w);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 256);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();  // the set-up above overlaps the previous kernel's tail

  if (warp == 0 && lane == 0) {
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const TileCoord t = decode_tile(tile, p, tiles_n, tiles_x, tiles_per_image);
      const int wrow = t.ph * p.Cout + t.n0;
      int kb = 0;
      for (int tap = 0; tap < p.taps; ++tap) {
        const int ay = t.y0 + p.dy[t.ph][tap], ax = t.x0 + p.dx[t.ph][tap];
        for (int cb = 0; cb < cpb; ++cb, ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          mbar_arrive_expect_tx(full_bar(stage), kStageBytes);
          tma_load_4d(sA + stage * kABytes, &map_a, full_bar(stage), cb * BK, ax, ay, t.b);
          tma_load_2d(sB + stage * kBBytes, &map_w, full_bar(stage), kb * BK, wrow);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    uint32_t stage = 0, phase = 0, it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int n0 = (tile % tiles_n) * BN;
      const int n_cur = min(BN, p.Cout - n0);
      const uint32_t idesc = umma_idesc_bf16(BM, (uint32_t)n_cur);
      const uint32_t as = it & 1, aph = (it >> 1) & 1;
      mbar_wait(tempty_bar(as), aph ^ 1);
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + as * BN;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint64_t da = umma_desc_k_sw128(sA + stage * kABytes);
        const uint64_t db = umma_desc_k_sw128(sB + stage * kBBytes);
#pragma unroll
        for (int k = 0; k < BK / 16; ++k) tc_mma_bf16(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
        tc_commit(empty_bar(stage));
        if (kb == nkb - 1) tc_commit(tfull_bar(as));
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int h = (warp - 4) >> 2;
    // transposition buffer of this warp (32 rows x 128 B, 16-byte chunks XOR-swizzled with row & 7):
    // tcgen05.ld gives thread t pixel t of the warp's 32; stores then run with 8 lanes per pixel so a
    // warp instruction covers whole 128-byte lines (see vit_gemm.cu)
    uint8_t* stg = smem + kStages * kStageBytes + 256 + (warp - 4) * kStageOutBytes;
    uint32_t it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const TileCoord t = decode_tile(tile, p, tiles_n, tiles_x, tiles_per_image);
      const ConvLanePixels px = conv_lane_pixels(p, t, q, lane);
      const uint32_t as = it & 1, aph = (it >> 1) & 1;
      mbar_wait_sleep(tfull_bar(as), aph, 32);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + h * 128;
      conv_epilogue_tile<EPI>(p, t, px, t_addr, h, lane, stg, [&] {
        tc_fence_before();
        mbar_arrive(tempty_bar(as));
      });
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

template <int EPI>
cudaError_t launch_one(const CUtensorMap& map_a, const CUtensorMap& map_w, const ConvParams& p, int grid,
                       cudaStream_t stream) {
  static SmemOptIn opt_in;  // one per EPI instantiation
  if (cudaError_t e = ensure_dynamic_smem(conv_gemm_kernel<EPI>, kSmemBytes, opt_in); e != cudaSuccess) return e;
  return launch_pdl(conv_gemm_kernel<EPI>, dim3(grid), dim3(kThreads), kSmemBytes, stream, map_a, map_w, p);
}

}  // namespace

int conv_smem_bytes() { return kSmemBytes; }

cudaError_t launch_conv(const CUtensorMap& map_a, const CUtensorMap& map_w, const ConvParams& p,
                        int num_sms, cudaStream_t stream) {
  if (p.B <= 0 || p.H <= 0 || p.W <= 0) return cudaSuccess;
  if (p.Cin <= 0 || p.Cin % BK != 0 || p.Cout <= 0 || p.Cout % 32 != 0) return cudaErrorInvalidValue;
  if (p.phases < 1 || p.phases > kConvMaxPhases || p.taps < 1 || p.taps > kConvMaxTaps)
    return cudaErrorInvalidValue;
  if (p.split_n % 64 != 0) return cudaErrorInvalidValue;
  const long long tiles = (long long)p.B * ((p.H + kConvTileH - 1) / kConvTileH) *
                          ((p.W + kConvTileW - 1) / kConvTileW) * p.phases * ((p.Cout + BN - 1) / BN);
  if (tiles > 0x7fffffffLL) return cudaErrorInvalidValue;
  const int grid = tiles < num_sms ? (int)tiles : num_sms;
  switch (p.epilogue) {
    case kConvBiasBf16: return launch_one<kConvBiasBf16>(map_a, map_w, p, grid, stream);
    case kConvBnGeluBf16: return launch_one<kConvBnGeluBf16>(map_a, map_w, p, grid, stream);
    case kConvBiasF32: return launch_one<kConvBiasF32>(map_a, map_w, p, grid, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace vc
