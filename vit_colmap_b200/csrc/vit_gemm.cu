// Persistent tcgen05 GEMM for the ViT linear layers: D[M,N] = A[M,K] * W[N,K]^T with the
// bias / GELU / LayerScale+residual / position-embedding work fused into the epilogue.
//   - one CTA per SM, tiles of 128 x 256 x 64 handed out round-robin (n fastest, so the CTAs that
//     run together share the same rows of A in L2; W stays L2-resident: <= 9.4 MB)
//   - warp 0: TMA producer (4-stage ring, 128-byte swizzle), warp 1: tcgen05.mma issuer
//     (M=128, N=256, K=16 per instruction), warp 2: TMEM allocator
//   - fp32 accumulators double-buffered in TMEM (2 x 256 columns): the 8 epilogue warps drain tile
//     i with tcgen05.ld while the tensor core fills tile i+1
//   - epilogue: tcgen05.ld hands thread t row t of the warp's 32 x 128-byte chunk; the chunk is
//     transposed through a swizzled shared-memory buffer so that every global access of a warp
//     covers 4 whole 128-byte lines (8 lanes x 16 B per row) instead of 32 different lines with
//     16 B each -- the L1/LSU line rate, not HBM, bounded the K = 768 GEMMs before
//   - kEpiResidualLnF32: the LayerNorm that follows a residual update runs in this kernel.  Every tile bumps
//     the counter of its 128-row block after its stores are visible device-wide.  The CTA that processes
//     the block's LAST column tile owns the block: it waits until the other column tiles have been counted
//     (they have lower tile indices, so they run in the same round or earlier and nobody waits in a cycle),
//     then its 8 epilogue warps read the 128 complete rows back from L2 (just written: no HBM read), compute
//     mean / variance exactly as the stand-alone kernel does (row in registers, two passes) and store the
//     bf16 rows the next GEMM loads.  A fixed owner keeps the work balanced: giving the block to whichever
//     CTA finishes last (first version) fed back on itself -- a CTA that had normalised one block was late
//     for all its later tiles, collected their blocks too and the GEMM waited for it (+ 170 us per GEMM).
//     MEASURED (ViT-B/14, 32 x 518^2, tools/time_vit.py): 13.95 ms per forward against 11.96 ms with one
//     launch per LayerNorm -- the 128 rows x 3 KB a block owner reads sit on its epilogue's critical path
//     (8 warps per SM cover far less L2 latency than the stand-alone kernel's 64) and owners wait on their
//     neighbours' tiles, so delays chain through a round.  vc_vit_forward therefore keeps the stand-alone
//     LayerNorm launches; this epilogue stays available (vc_vit_linear_ln, VC_VIT_FUSED_LN=1), bit-identical.
//   - the per-tile epilogue lives in vit_gemm_epilogue.cuh and is shared with the two-SM kernel (vit_gemm2.cu,
//     the default for K >= 768); this kernel serves K < 768, VC_VIT_2CTA=0 and the LayerNorm-placement options
//     (kEpiResidualLnF32, kEpiResidualPublishF32 + layernorm_follow_kernel below)
// Activations are bf16, accumulation and the residual stream are fp32.
#include "vit_kernels.cuh"

#include <math.h>

#include "vc_ptx.cuh"
#include "vc_launch.cuh"
#include "vc_epilogue.cuh"
#include "vit_gemm_epilogue.cuh"

namespace vc {

namespace {

constexpr int BM = 128, BN = 256, BK = 64, kStages = 4;
constexpr int kABytes = BM * BK * 2, kBBytes = BN * BK * 2, kStageBytes = kABytes + kBBytes;
constexpr int kThreads = 384;
constexpr int kTmemCols = 512;
constexpr int kStageOutBytes = 32 * 128;  // per epilogue warp: 32 rows x 128 B transposition buffer
constexpr int kSmemBytes = kStages * kStageBytes + 1024 + 256 + 8 * kStageOutBytes;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// LayerNorm (eps 1e-6, biased variance around the mean: torch.nn.LayerNorm) of RIF fp32 rows at a time by one
// warp, the rows read through L2 (ld.global.cg: other SMs wrote most of them), bf16 out.  Same arithmetic
// per row as layernorm_kernel in vit_pointwise.cu (bit-identical results); RIF rows are loaded before the
// first reduction so that the L2 latency is paid once per batch of rows, not once per row.
// KV = float4 per lane (C <= 128 * KV).
template <int KV, int RIF>
__device__ __forceinline__ void ln_rows_bf16(const float* __restrict__ x, size_t ldx, const float* __restrict__ w,
                                             const float* __restrict__ b, __nv_bfloat16* __restrict__ out, int C,
                                             int row_first, int row_step, int n_rows, int lane) {
  const int nvec = C >> 2;
  for (int r0 = row_first; r0 < n_rows; r0 += RIF * row_step) {
    float4 v[RIF][KV];
#pragma unroll
    for (int j = 0; j < RIF; ++j) {
      const int r = r0 + j * row_step;
#pragma unroll
      for (int k = 0; k < KV; ++k) {
        const int i = lane + 32 * k;
        v[j][k] = (r < n_rows && i < nvec) ? __ldcg(reinterpret_cast<const float4*>(x + (size_t)r * ldx) + i)
                                           : make_float4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int j = 0; j < RIF; ++j) {
      const int r = r0 + j * row_step;
      if (r >= n_rows) break;  // warp-uniform
      float s = 0.f;
#pragma unroll
      for (int k = 0; k < KV; ++k) s += v[j][k].x + v[j][k].y + v[j][k].z + v[j][k].w;
      const float mean = warp_sum(s) / (float)C;
      float ss = 0.f;
#pragma unroll
      for (int k = 0; k < KV; ++k) {
        if (lane + 32 * k < nvec) {
          const float a = v[j][k].x - mean, c = v[j][k].y - mean, d = v[j][k].z - mean, e = v[j][k].w - mean;
          ss += a * a + c * c + d * d + e * e;
        }
      }
      const float rstd = rsqrtf(warp_sum(ss) / (float)C + 1e-6f);
#pragma unroll
      for (int k = 0; k < KV; ++k) {
        const int i = lane + 32 * k;
        if (i < nvec) {
          const float4 g = __ldg(reinterpret_cast<const float4*>(w) + i);
          const float4 be = __ldg(reinterpret_cast<const float4*>(b) + i);
          uint2 pk;
          pk.x = pack_bf16((v[j][k].x - mean) * rstd * g.x + be.x, (v[j][k].y - mean) * rstd * g.y + be.y);
          pk.y = pack_bf16((v[j][k].z - mean) * rstd * g.z + be.z, (v[j][k].w - mean) * rstd * g.w + be.w);
          reinterpret_cast<uint2*>(out + (size_t)r * C)[i] = pk;
        }
      }
    }
  }
}

template <int EPI>
__device__ __forceinline__ void gemm_body(const CUtensorMap& map_a, const CUtensorMap& map_w, const GemmParams& p) {
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
  const int tiles_m = (p.M + BM - 1) / BM;
  const int n_tiles = tiles_m * tiles_n;

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
  // everything above overlaps the previous kernel's tail; operands and the residual stream are only
  // touched below
  griddep_wait();
  // the row-block consumer (launch_layernorm_follow) is the next kernel of the stream: let it become resident now
  if (EPI == kEpiResidualPublishF32) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

  if (warp == 0 && lane == 0) {
    uint32_t stage = 0, phase = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
      const int tm = tile / tiles_n, tn = tile - tm * tiles_n;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait(empty_bar(stage), phase ^ 1);
        mbar_arrive_expect_tx(full_bar(stage), kStageBytes);
        tma_load_2d(sA + stage * kABytes, &map_a, full_bar(stage), kb * BK, tm * BM);
        tma_load_2d(sB + stage * kBBytes, &map_w, full_bar(stage), kb * BK, tn * BN);
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    constexpr uint32_t idesc = umma_idesc_bf16(BM, BN);
    uint32_t stage = 0, phase = 0, it = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
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
    // this warp's transposition buffer: 32 rows x 128 B, 16-byte chunks XOR-swizzled with (row & 7)
    uint8_t* stg = smem + kStages * kStageBytes + 256 + (warp - 4) * kStageOutBytes;
    uint32_t it = 0;
    int pending_tm = -1;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
      const int tm = tile / tiles_n, tn = tile - tm * tiles_n;
      const uint32_t as = it & 1, aph = (it >> 1) & 1;
      float ln_mean = 0.f, ln_rstd = 1.f;
      if (gemm_is_fold(EPI)) gemm_row_stats(p, tm * BM + q * 32 + lane, ln_mean, ln_rstd);
      mbar_wait_sleep(tfull_bar(as), aph, 32);
      tc_fence_after();
      if (EPI == kEpiResidualPublishF32 && pending_tm >= 0) {
        // the previous tile's rows left this warp a whole main loop ago: the fence returns at once.  Every
        // epilogue warp counts for itself (8 per tile), so no CTA-wide barrier is needed.
        __threadfence();
        if (lane == 0) atomicAdd(p.tile_counters + pending_tm, 1);
        pending_tm = -1;
      }
      const uint32_t t_addr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN + h * 128;
      gemm_epilogue_tile<EPI>(p, t_addr, tm, tn, q, h, lane, stg, [&] {
        tc_fence_before();
        mbar_arrive(tempty_bar(as));
      }, ln_mean, ln_rstd);
      if (EPI == kEpiResidualPublishF32) pending_tm = tm;  // published one tile later, when the fence is free
      if (EPI == kEpiResidualLnF32) {
        // tile stored: publish it; the CTA holding the block's last column tile normalises the block
        __threadfence();                                          // this thread's stores, device-wide
        asm volatile("bar.sync 1, 256;" ::: "memory");            // all 8 epilogue warps have stored
        const bool owner = tn == tiles_n - 1;                      // CTA-uniform
        if (threadIdx.x == 128) {
          if (!owner) {
            atomicAdd(p.tile_counters + tm, 1);
          } else {
            volatile int* cnt = p.tile_counters + tm;
            while (*cnt < tiles_n - 1) __nanosleep(64);            // the block's other column tiles
            *cnt = 0;                                              // left zero for the next launch
          }
        }
        if (owner) {
          asm volatile("bar.sync 1, 256;" ::: "memory");          // thread 128 has seen the full count
          __threadfence();                                        // order the reads after the counter
          const float* xb = static_cast<const float*>(p.out) + (size_t)tm * BM * p.ldo;
          __nv_bfloat16* ob = p.ln_out + (size_t)tm * BM * p.N;
          const int n_rows = min(BM, p.M - tm * BM);
          if (p.N <= 384) ln_rows_bf16<3, 8>(xb, p.ldo, p.ln_w, p.ln_b, ob, p.N, warp - 4, 8, n_rows, lane);
          else if (p.N <= 768) ln_rows_bf16<6, 4>(xb, p.ldo, p.ln_w, p.ln_b, ob, p.N, warp - 4, 8, n_rows, lane);
          else ln_rows_bf16<8, 2>(xb, p.ldo, p.ln_w, p.ln_b, ob, p.N, warp - 4, 8, n_rows, lane);
        }
      }
    }
    if (EPI == kEpiResidualPublishF32 && pending_tm >= 0) {  // the CTA's last tile
      __threadfence();
      if (lane == 0) atomicAdd(p.tile_counters + pending_tm, 1);
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, kTmemCols);
}

template <int EPI>
__global__ void __launch_bounds__(kThreads, 1)
vit_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                const GemmParams p) {
  gemm_body<EPI>(map_a, map_w, p);
}
// The publishing form shares its SM with one CTA of layernorm_follow_kernel (128 threads x 80 registers):
// 384 x 144 + 128 x 80 registers fit the 64 K register file, 384 x 152 (what ptxas picks unconstrained) do not.
__global__ void __maxnreg__(144)
vit_gemm_publish_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_w,
                        const GemmParams p) {
  gemm_body<kEpiResidualPublishF32>(map_a, map_w, p);
}

// Row-block consumer of a kEpiResidualPublishF32 GEMM (see vit_kernels.cuh).  Small on purpose: 128 threads and
// 80 registers, so that one CTA fits on an SM next to the GEMM's CTA (384 threads x 144 registers).  A unit is
// 32 rows (a quarter of a row block): after the GEMM's last round 148 CTAs share the ~50 blocks that are left.
// It must NOT wait for the producer grid at its start -- the counters are the dependency -- and does so at its
// end, which orders every later kernel of the stream after the GEMM as well.
template <int KV, int RIF>
__global__ void __launch_bounds__(128, 6)
layernorm_follow_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                        __nv_bfloat16* __restrict__ out, int rows, int C, int tiles_n, int* counters, int* done) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units = ((rows + BM - 1) / BM) * 4;
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int tm = u >> 2, row0 = tm * BM + (u & 3) * 32;
    if (threadIdx.x == 0) {
      volatile int* cnt = counters + tm;
      while (*cnt < 8 * tiles_n) __nanosleep(128);  // eight epilogue warps count each tile
    }
    __syncthreads();
    __threadfence();  // order the row reads after the counter
    const int n_rows = min(32, rows - row0);
    if (n_rows > 0) ln_rows_bf16<KV, RIF>(x + (size_t)row0 * C, C, w, b, out + (size_t)row0 * C, C, warp, 4, n_rows, lane);
    __syncthreads();
    if (threadIdx.x == 0 && atomicAdd(done + tm, 1) == 3) {  // the block's last unit leaves both counters zero
      done[tm] = 0;
      counters[tm] = 0;
    }
  }
  griddep_wait();
}

template <int EPI>
cudaError_t launch_one(const CUtensorMap& map_a, const CUtensorMap& map_w, const GemmParams& p, int grid,
                       cudaStream_t stream) {
  static SmemOptIn opt_in;  // one per EPI instantiation
  if (cudaError_t e = ensure_dynamic_smem(vit_gemm_kernel<EPI>, kSmemBytes, opt_in); e != cudaSuccess) return e;
  return launch_pdl(vit_gemm_kernel<EPI>, dim3(grid), dim3(kThreads), kSmemBytes, stream, map_a, map_w, p);
}

}  // namespace

int gemm_smem_bytes() { return kSmemBytes; }

cudaError_t launch_layernorm_follow(const float* x, const float* w, const float* b, __nv_bfloat16* out, int rows,
                                    int C, int tiles_n, int* counters, int* done, int num_sms, cudaStream_t stream) {
  if (rows <= 0) return cudaSuccess;
  if (C % 4 != 0 || C > 1024 || !counters || !done) return cudaErrorInvalidValue;
  const int units = ((rows + BM - 1) / BM) * 4;
  const dim3 grid(units < num_sms ? units : num_sms), block(128);
  if (C <= 384) return launch_pdl(layernorm_follow_kernel<3, 4>, grid, block, 0, stream, x, w, b, out, rows, C, tiles_n, counters, done);
  if (C <= 768) return launch_pdl(layernorm_follow_kernel<6, 2>, grid, block, 0, stream, x, w, b, out, rows, C, tiles_n, counters, done);
  return launch_pdl(layernorm_follow_kernel<8, 1>, grid, block, 0, stream, x, w, b, out, rows, C, tiles_n, counters, done);
}

cudaError_t launch_gemm(const CUtensorMap& map_a, const CUtensorMap& map_w, const GemmParams& p,
                        int num_sms, cudaStream_t stream) {
  if (p.M <= 0 || p.N <= 0) return cudaSuccess;
  if (p.K <= 0 || p.K % BK != 0 || p.N % 32 != 0) return cudaErrorInvalidValue;
  const int n_tiles = ((p.M + BM - 1) / BM) * ((p.N + BN - 1) / BN);
  const int grid = n_tiles < num_sms ? n_tiles : num_sms;
  switch (p.epilogue) {
    case kEpiBiasBf16: return launch_one<kEpiBiasBf16>(map_a, map_w, p, grid, stream);
    case kEpiBiasGeluBf16: return launch_one<kEpiBiasGeluBf16>(map_a, map_w, p, grid, stream);
    case kEpiResidualF32: return launch_one<kEpiResidualF32>(map_a, map_w, p, grid, stream);
    case kEpiPatchEmbed: return launch_one<kEpiPatchEmbed>(map_a, map_w, p, grid, stream);
    case kEpiFoldBiasBf16:
    case kEpiFoldBiasGeluBf16:
      if (!p.ln_part || !p.colsum || p.ln_slices <= 0 || p.ln_slices * 128 != p.K) return cudaErrorInvalidValue;
      return p.epilogue == kEpiFoldBiasBf16 ? launch_one<kEpiFoldBiasBf16>(map_a, map_w, p, grid, stream)
                                            : launch_one<kEpiFoldBiasGeluBf16>(map_a, map_w, p, grid, stream);
    case kEpiResidualStatsF32:
      if (!p.xb_out || !p.part_out || p.N % 128 != 0 || p.N != p.ldo) return cudaErrorInvalidValue;
      return launch_one<kEpiResidualStatsF32>(map_a, map_w, p, grid, stream);
    case kEpiResidualPublishF32: {
      if (!p.tile_counters) return cudaErrorInvalidValue;
      static SmemOptIn opt_in;
      if (cudaError_t e = ensure_dynamic_smem(vit_gemm_publish_kernel, kSmemBytes, opt_in); e != cudaSuccess) return e;
      return launch_pdl(vit_gemm_publish_kernel, dim3(grid), dim3(kThreads), kSmemBytes, stream, map_a, map_w, p);
    }
    case kEpiResidualLnF32:
      if (!p.ln_w || !p.ln_b || !p.ln_out || !p.tile_counters || p.N != p.ldo || p.N > 1024 || p.N % 4 != 0)
        return cudaErrorInvalidValue;
      return launch_one<kEpiResidualLnF32>(map_a, map_w, p, grid, stream);
    default: return cudaErrorInvalidValue;
  }
}

}  // namespace vc
