// Two-SM (cta_group::2) form of the persistent ViT GEMM in vit_gemm.cu: D[M,N] = A[M,K] * W[N,K]^T.
// A cluster of two CTAs on one TPC works on one 256 x 256 tile of D: each CTA stages its own 128 rows of
// A and HALF of the 256 weight rows, the leader CTA issues one tcgen05.mma.cta_group::2 (M = 256, N = 256)
// per k-step, and each CTA finds the fp32 accumulators of its own 128 rows x all 256 columns in its TMEM.
// Per 512 tensor cycles a CTA pulls 32 KB through L2 instead of 48 KB -- the one-SM kernel's tensor pipe is
// 67-70 % active with the L2 -> SM operand feed next in line -- and the shared-memory ring
// holds 6 stages instead of 4.  Same synchronisation scheme as match_fused2.cu: both CTAs' TMA completions
// are counted on the leader's barrier, tcgen05.commit multicasts to both CTAs' barriers, and the accumulator
// buffer returns to the leader once the epilogue warps of BOTH CTAs have read it (remote mbarrier arrive).
// The per-tile epilogue is the one-SM kernel's (vit_gemm_epilogue.cuh): results are bit-identical.
#include "vit_kernels.cuh"

#include "vc_launch.cuh"
#include "vc_ptx.cuh"
#include "vit_gemm_epilogue.cuh"

namespace vc {

namespace {

constexpr int BM = kGemmBM, BN = kGemmBN, BK = 64, kStages = 6;
constexpr int kABytes = BM * BK * 2, kBBytes = (BN / 2) * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kThreads = 384;  // warp 0 TMA, warp 1 MMA (leader only), warp 2 TMEM alloc, warps 4-11 epilogue
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
// default semantics (.release.cta): the TMEM hand-over is ordered by the tcgen05 fences on either side
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
// TMA load whose completion bytes are counted on the LEADER CTA's mbarrier; the data lands in this CTA
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst_smem, const CUtensorMap* m, uint32_t bar_leader, int c0,
                                                int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(m)), "r"(bar_leader), "r"(c0), "r"(c1)
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
vit_gemm2_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                 const GemmParams p) {
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
  const int nkb = p.K / BK;
  const int tiles_n = (p.N + BN - 1) / BN;
  const int tiles_m2 = (p.M + 2 * BM - 1) / (2 * BM);  // 256-row tiles: CTA rank r takes rows [256 t + 128 r, +128)
  const int n_tiles = tiles_m2 * tiles_n;
  const uint32_t rank = cluster_rank();
  const int cluster_id = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

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
      mbar_init(tempty_bar(s), 2 * kEpiWarps);  // one arrival per epilogue warp of either CTA
    }
    fence_mbar_init();
  }
  if (warp == 2) {
    tmem_alloc_2sm(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish_2sm();
  }
  tc_fence_before();
  cluster_sync_all();  // barriers of both CTAs initialised before any remote arrival / multicast
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_wait();  // everything above overlaps the previous kernel's tail

  if (warp == 0 && lane == 0) {
    uint32_t stage = 0, phase = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters) {
      const int tm2 = tile / tiles_n, tn = tile - tm2 * tiles_n;
      const int row_a = (2 * tm2 + (int)rank) * BM;          // may lie beyond M: TMA zero-fills, rows are masked
      const int row_w = tn * BN + (int)rank * (BN / 2);      // may lie beyond N likewise
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1);
        const uint32_t lead_full = map_to_rank(full_bar(stage), 0);
        if (rank == 0) mbar_arrive_expect_tx(full_bar(stage), 2 * kStageBytes);  // both CTAs' bytes
        tma_load_2d_2sm(sA + stage * kABytes, &map_a, lead_full, kb * BK, row_a);
        tma_load_2d_2sm(sB + stage * kBBytes, &map_w, lead_full, kb * BK, row_w);
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1 && lane == 0 && rank == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN);
    uint32_t stage = 0, phase = 0, it = 0;
    for (int tile = cluster_id; tile < n_tiles; tile += n_clusters, ++it) {
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
      const int tm2 = tile / tiles_n, tn = tile - tm2 * tiles_n;
      const int tm = 2 * tm2 + (int)rank;
      const uint32_t as = it & 1, aph = (it >> 1) & 1;
      float ln_mean = 0.f, ln_rstd = 1.f;
      if (gemm_is_fold(EPI)) gemm_row_stats(p, tm * BM + q * 32 + lane, ln_mean, ln_rstd);
      mbar_wait_sleep(tfull_bar(as), aph, 32);
      tc_fence_after();
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + h * 128;
      gemm_epilogue_tile<EPI>(p, t_addr, tm, tn, q, h, lane, stg, [&] {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(lead_tempty[as]);
      }, ln_mean, ln_rstd);
    }
  }

  tc_fence_before();
  cluster_sync_all();  // the peer may still multicast into this CTA's barriers / read its shared memory
  if (warp == 2) tmem_dealloc_2sm(tmem_base, kTmemCols);
}

template <int EPI>
cudaError_t launch_one2(const CUtensorMap& map_a, const CUtensorMap& map_w, const GemmParams& p, int clusters,
                        cudaStream_t stream) {
  static SmemOptIn opt_in;  // one per EPI instantiation
  if (cudaError_t e = ensure_dynamic_smem(vit_gemm2_kernel<EPI>, kSmemBytes, opt_in); e != cudaSuccess) return e;
  return launch_pdl(vit_gemm2_kernel<EPI>, dim3(2 * clusters), dim3(kThreads), kSmemBytes, stream, map_a, map_w, p);
}

}  // namespace

bool gemm2_supports(int epilogue) {
  return epilogue == kEpiBiasBf16 || epilogue == kEpiBiasGeluBf16 || epilogue == kEpiResidualF32 ||
         epilogue == kEpiPatchEmbed || epilogue == kEpiFoldBiasBf16 || epilogue == kEpiFoldBiasGeluBf16 ||
         epilogue == kEpiResidualStatsF32;
}

// map_w must have a {64, 128} box (half of the 256 weight rows per CTA); map_a the usual {64, 128}.
cudaError_t launch_gemm2(const CUtensorMap& map_a, const CUtensorMap& map_w, const GemmParams& p, int num_sms,
                         cudaStream_t stream) {
  if (p.M <= 0 || p.N <= 0) return cudaSuccess;
  if (p.K <= 0 || p.K % BK != 0 || p.N % 32 != 0 || !gemm2_supports(p.epilogue)) return cudaErrorInvalidValue;
  const int n_tiles = ((p.M + 2 * BM - 1) / (2 * BM)) * ((p.N + BN - 1) / BN);
  const int clusters = n_tiles < num_sms / 2 ? n_tiles : num_sms / 2;
  switch (p.epilogue) {
    case kEpiBiasBf16: return launch_one2<kEpiBiasBf16>(map_a, map_w, p, clusters, stream);
    case kEpiBiasGeluBf16: return launch_one2<kEpiBiasGeluBf16>(map_a, map_w, p, clusters, stream);
    case kEpiResidualF32: return launch_one2<kEpiResidualF32>(map_a, map_w, p, clusters, stream);
    case kEpiFoldBiasBf16:
    case kEpiFoldBiasGeluBf16:
      if (!p.ln_part || !p.colsum || p.ln_slices <= 0 || p.ln_slices * 128 != p.K) return cudaErrorInvalidValue;
      return p.epilogue == kEpiFoldBiasBf16 ? launch_one2<kEpiFoldBiasBf16>(map_a, map_w, p, clusters, stream)
                                            : launch_one2<kEpiFoldBiasGeluBf16>(map_a, map_w, p, clusters, stream);
    case kEpiResidualStatsF32:
      if (!p.xb_out || !p.part_out || p.N % 128 != 0 || p.N != p.ldo) return cudaErrorInvalidValue;
      return launch_one2<kEpiResidualStatsF32>(map_a, map_w, p, clusters, stream);
    default: return launch_one2<kEpiPatchEmbed>(map_a, map_w, p, clusters, stream);
  }
}

}  // namespace vc
