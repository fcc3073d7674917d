// Two-SM (cta_group::2) form of the implicit-GEMM convolution in conv_gemm.cu.  A cluster of two CTAs works on TWO
// consecutive 8 x 16 pixel patches and one block of up to 256 output channels: each CTA fetches its own patch per
// filter tap (4-D TMA box, zero padding by out-of-bounds fill) and HALF of the weight rows, the leader issues one
// tcgen05.mma.cta_group::2 (M = 256, N = channels of the block) per k-step, and each CTA finds the accumulators of
// its own 128 pixels in its TMEM.  A third less operand traffic through L2 per MMA and 6 pipeline stages instead of
// 4 (as in vit_gemm2.cu / match_fused2.cu, whose synchronisation scheme this follows).  The per-tile epilogue is
// conv_gemm.cu's (conv_gemm_epilogue.cuh): outputs are bit-identical.
#include "head_kernels.cuh"

#include "conv_gemm_epilogue.cuh"
#include "vc_launch.cuh"
#include "vc_ptx.cuh"

namespace vc {

namespace {

constexpr int BM = kConvBM, BN = kConvBN, BK = 64, kStages = 6;
constexpr int kABytes = BM * BK * 2, kBBytes = (BN / 2) * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kThreads = 384;
constexpr int kEpiWarps = 8;
constexpr int kTmemCols = 512;
constexpr int kStageOutBytes = 32 * 128;
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256 + kEpiWarps * kStageOutBytes;

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t map_to_rank(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst_smem, const CUtensorMap* m, uint32_t bar_leader, int c0,
                                                int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_leader), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d_2sm(uint32_t dst_smem, const CUtensorMap* m, uint32_t bar_leader, int c0,
                                                int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_leader), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc_2sm(uint32_t dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem), "r"(ncols)
               : "memory");
}
__device__ __forceinline__ void tmem_relinquish_2sm() {
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_2sm(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_commit_2sm(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
      ::"r"(bar), "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tc_mma_bf16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                                uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
conv_gemm2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
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
  const int patch_pairs = (p.B * tiles_per_image + 1) / 2;  // CTA rank r takes patch 2 * pair + r
  const int n_tiles = patch_pairs * p.phases * tiles_n;
  const uint32_t rank = cluster_rank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;
  // tile -> (256-channel block fastest, then sub-pixel phase, then patch pair)
  auto coord = [&](int tile) {
    const int tn = tile % tiles_n;
    const int r = tile / tiles_n;
    return conv_tile_coord(p, 2 * (r / p.phases) + (int)rank, r % p.phases, tn, tiles_x, tiles_per_image);
  };

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
      mbar_init(tempty_bar(s), 2 * kEpiWarps);
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2sm(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();

  if (warp == 0 && lane == 0) {
    uint32_t stage = 0, phase = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
      const ConvTileCoord t = coord(tile);  // an invalid patch has b == B: every box is out of bounds = zeros
      const int n_cur = min(BN, p.Cout - t.n0);
      const int wrow = t.ph * p.Cout + t.n0 + (int)rank * (n_cur / 2);  // this CTA's half of the block's weight rows
      int kb = 0;
      for (int tap = 0; tap < p.taps; ++tap) {
        const int ay = t.y0 + p.dy[t.ph][tap], ax = t.x0 + p.dx[t.ph][tap];
        for (int cb = 0; cb < cpb; ++cb, ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t lead_full = map_to_rank(full_bar(stage), 0);
          if (rank == 0) mbar_arrive_expect_tx(full_bar(stage), 2 * kStageBytes);
          tma_load_4d_2sm(sA + stage * kABytes, &map_a, lead_full, cb * BK, ax, ay, t.b);
          tma_load_2d_2sm(sB + stage * kBBytes, &map_w, lead_full, kb * BK, wrow);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1 && lane == 0 && rank == 0) {
    uint32_t stage = 0, phase = 0, it = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters, ++it) {
      const int n0 = (tile % tiles_n) * BN;
      const int n_cur = min(BN, p.Cout - n0);
      const uint32_t idesc = umma_idesc_bf16(2 * BM, (uint32_t)n_cur);
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
        for (int k = 0; k < BK / 16; ++k) tc_mma_bf16_2sm(d_tmem, da + 2 * k, db + 2 * k, idesc, (kb | k) != 0);
        tc_commit_2sm(empty_bar(stage));
        if (kb == nkb - 1) tc_commit_2sm(tfull_bar(as));
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    const int q = warp & 3;
    const int h = (warp - 4) >> 2;
    uint8_t* stg = smem + kStages * kStageBytes + 256 + (warp - 4) * kStageOutBytes;
    const uint32_t lead_tempty[2] = {map_to_rank(tempty_bar(0), 0), map_to_rank(tempty_bar(1), 0)};
    uint32_t it = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters, ++it) {
      const ConvTileCoord t = coord(tile);
      const ConvLanePixels px = conv_lane_pixels(p, t, q, lane);
      const uint32_t as = it & 1, aph = (it >> 1) & 1;
      mbar_wait_sleep(tfull_bar(as), aph, 32);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + h * 128;
      conv_epilogue_tile<EPI>(p, t, px, t_addr, h, lane, stg, [&] {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(lead_tempty[as]);
      });
    }
  }

  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) tmem_dealloc_2sm(tmem_base, kTmemCols);
}

template <int EPI>
cudaError_t launch_one2(const CUtensorMap& map_a, const CUtensorMap& map_w, const ConvParams& p, int clusters,
                        cudaStream_t stream) {
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(conv_gemm2_kernel<EPI>, kSmemBytes, opt_in); e != cudaSuccess) return e;
  return launch_pdl(conv_gemm2_kernel<EPI>, dim3(2 * clusters), dim3(kThreads), kSmemBytes, stream, map_a, map_w, p);
}

}  // namespace

// map_w must have a {64, 128} box.  Every 256-channel block must split into two halves of a multiple of 16 channels
// (Cout % 32 == 0 guarantees it).
cudaError_t launch_conv2(const CUtensorMap& map_a, const CUtensorMap& map_w, const ConvParams& p, int num_sms,
                         cudaStream_t stream) {
  if (p.B <= 0 || p.H <= 0 || p.W <= 0) return cudaSuccess;
  if (p.Cin <= 0 || p.Cin % BK != 0 || p.Cout <= 0 || p.Cout % 32 != 0) return cudaErrorInvalidValue;
  if (p.phases < 1 || p.phases > kConvMaxPhases || p.taps < 1 || p.taps > kConvMaxTaps) return cudaErrorInvalidValue;
  if (p.split_n % 64 != 0) return cudaErrorInvalidValue;
  const long long patches = (long long)p.B * ((p.H + kConvTileH - 1) / kConvTileH) * ((p.W + kConvTileW - 1) / kConvTileW);
  const long long tiles = ((patches + 1) / 2) * p.phases * ((p.Cout + BN - 1) / BN);
  if (tiles > 0x7fffffffLL) return cudaErrorInvalidValue;
  const int clusters = tiles < num_sms / 2 ? (int)tiles : num_sms / 2;
  switch (p.epilogue) {
    case kConvBiasBf16: return launch_one2<kConvBiasBf16>(map_a, map_w, p, clusters, stream);
    case kConvBnGeluBf16: return launch_one2<kConvBnGeluBf16>(map_a, map_w, p, clusters, stream);
    case kConvBiasF32: return launch_one2<kConvBiasF32>(map_a, map_w, p, clusters, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace vc
