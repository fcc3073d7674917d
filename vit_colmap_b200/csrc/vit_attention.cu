// Multi-head self-attention of the ViT blocks (head_dim 64) on tcgen05: softmax(Q K^T / 8) V per
// (image, head), straight from the packed QKV GEMM output, without materialising the T x T scores.
//
// CTA = (image, head, pair of 128-row query tiles).  Roles:
//   warp 0  TMA: both Q tiles once, then a 3-stage ring of (K_j, V_j) 128-key tiles (128B swizzle)
//   warp 1 / 2  tcgen05.mma issuer of query tile 0 / 1 (warp 1 also allocates TMEM; one issuer per
//           tile so every wait is a blocking mbarrier wait in its natural order, no polling):
//           S_w = Q_w K_j^T (M128 N128 K64) into TMEM as soon as the softmax threads have pulled the
//           previous S_w into registers (so Q K^T of tile j+1 overlaps the softmax of tile j), and,
//           once a softmax half has published its P (bf16, written to shared memory in the K-major
//           swizzled operand layout), O_{w,h} += P_{w,h} V_{j,h} (M128 N64 K64; V is consumed as an
//           MN-major operand exactly as TMA delivered it, so no transposed copy of V exists).
//           TMEM: S_0, S_1 (2 x 128 columns), O_{w,h} (4 x 64 columns).
//   warps 4-11 / 12-19  softmax warpgroup of query tile 0 / 1, 256 threads: thread = one query row x
//           one 64-key half of the tile (warp%4 = TMEM lane quarter, (warp-4)/4%2 = key half), so
//           every SM sub-partition holds four softmax warps to hide the MUFU / TMEM latencies.
//           The two key halves of a row are independent split-K streams: each keeps its own lazily
//           updated reference maximum, row sum and O accumulator (rescaled only when the maximum
//           grew by more than 2^8, so the TMEM read-modify-write of O is rare) and never
//           synchronises with its sibling until the end, where the two partial results are merged
//           as in split-K flash attention: O = (a0 O_0 + a1 O_1) / (a0 l_0 + a1 l_1), a_h = 2^(m_h - m).
// The two warpgroups share every K/V tile and keep the tensor core busy for each other.
#include "vit_kernels.cuh"

#include <math.h>
#include <stdlib.h>

#include "vc_ptx.cuh"

namespace vc {

namespace {

constexpr int kTile = 128, kHd = 64, kKvStages = 3;
constexpr int kQBytes = kTile * kHd * 2;          // 16 KB
constexpr int kKvBytes = 2 * kTile * kHd * 2;     // K + V of one stage, 32 KB
constexpr int kPBytes = kTile * kTile * 2;        // 32 KB per warpgroup
constexpr int kThreads = 640;  // warps 0-3 control (TMA, two MMA issuers, spare), warps 4-19 softmax
constexpr int kTmemCols = 512;
constexpr int kXchBytes = 2 * 2 * kTile * 8;      // (m, l) of every (warpgroup, half, row)
constexpr int kSmemBytes = 2 * kQBytes + kKvStages * kKvBytes + 2 * kPBytes + kXchBytes + 1024 + 256;

// K-major A (bits 15 = 0), MN-major B (bit 16 = 1): B = V tile [keys, 64] with the head dim contiguous
__host__ __device__ constexpr uint32_t idesc_pv() { return umma_idesc_bf16(128, 64) | (1u << 16); }

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Blackwell packed fp32 pairs (FFMA2 / FADD2 / FMUL2) and the 3-input maximum (FMNMX3)
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)),
        "l"(*reinterpret_cast<uint64_t*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float fmax3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}


// 2^y for a pair on the FMA pipe instead of the MUFU (the kernel is bound by the 16 exp2 per clock of
// an SM, so a share of the exponentials is moved over, as in FlashAttention-4): Cody-Waite split with
// the 1.5*2^23 magic constant (n = round(y) lands in the low mantissa bits of t, f = y - n in
// [-0.5, 0.5]), degree-3 minimax polynomial for 2^f (max relative error 7.5e-5, 50x below the bf16
// rounding of P), n added into the exponent field.  y is clamped at -125 so the result never
// leaves the normal range; y <= +8 by construction of the lazy reference maximum.
__device__ __forceinline__ float2 exp2_poly2(float2 y) {
  y.x = fmaxf(y.x, -125.0f);
  y.y = fmaxf(y.y, -125.0f);
  const float2 magic = make_float2(12582912.0f, 12582912.0f);
  const float2 t = fadd2(y, magic);
  const float2 n = fadd2(t, make_float2(-12582912.0f, -12582912.0f));
  const float2 f = fadd2(y, make_float2(-n.x, -n.y));
  float2 q = ffma2(make_float2(5.517145991e-02f, 5.517145991e-02f), f, make_float2(2.426108569e-01f, 2.426108569e-01f));
  q = ffma2(q, f, make_float2(6.932609677e-01f, 6.932609677e-01f));
  q = ffma2(q, f, make_float2(9.999281168e-01f, 9.999281168e-01f));
  return make_float2(__uint_as_float(__float_as_uint(q.x) + (__float_as_uint(t.x) << 23)),
                     __uint_as_float(__float_as_uint(q.y) + (__float_as_uint(t.y) << 23)));
}

// One 64-key half tile of the online softmax for one query row (thread = row x key half).  The 64
// scores are pulled into registers with one round of tcgen05.ld and never re-read.  kMasked is only
// instantiated for the last, partial key tile, so full tiles carry no per-element selects.
template <bool kMasked, int kPolyPairs>
__device__ __forceinline__ void softmax_half(uint32_t t_s, uint32_t t_o, uint8_t* prow, int r, int valid,
                                             float sc, bool have_o, uint32_t bar_s_free, uint32_t bar_o_ready,
                                             uint32_t o_parity, float& m_ref, float& l_run) {
  uint32_t a[32], b[32];
  tmem_ld_32x32(t_s, a);
  tmem_ld_32x32(t_s + 32, b);
  tmem_wait_ld();
  // the scores now live in registers: hand S back so Q K^T of the next key tile overlaps this softmax
  tc_fence_before();
  mbar_arrive(bar_s_free);
  if (kMasked) {
#pragma unroll
    for (int e = 0; e < 32; ++e) {
      if (e >= valid) a[e] = 0xFF800000u;  // -inf
      if (32 + e >= valid) b[e] = 0xFF800000u;
    }
  }
  float mx[4] = {-1e30f, -1e30f, -1e30f, -1e30f};
#pragma unroll
  for (int e = 0; e < 32; e += 8) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      mx[q] = fmax3(mx[q], __uint_as_float(a[e + 2 * q]), __uint_as_float(a[e + 2 * q + 1]));
      mx[q] = fmax3(mx[q], __uint_as_float(b[e + 2 * q]), __uint_as_float(b[e + 2 * q + 1]));
    }
  }
  const float m_tile = fmaxf(fmaxf(mx[0], mx[1]), fmaxf(mx[2], mx[3])) * sc;
  if (have_o) {
    // P V of the previous key tile must have completed before P is overwritten / O is rescaled
    mbar_wait_sleep(bar_o_ready, o_parity, 20);
    tc_fence_after();
  }
  // lazy reference maximum: move it (and rescale O, l) only when it is more than 2^8 behind
  if (__any_sync(0xffffffffu, m_tile > m_ref + 8.0f)) {
    const float m_new = fmaxf(m_ref, m_tile);
    const float alpha = ex2(m_ref - m_new);
    m_ref = m_new;
    l_run *= alpha;
    if (have_o) {
      const float2 a2 = make_float2(alpha, alpha);
#pragma unroll 1
      for (int c = 0; c < 4; ++c) {
        uint32_t o[16];
        tmem_ld_32x16(t_o + c * 16, o);
        tmem_wait_ld();
#pragma unroll
        for (int e = 0; e < 16; e += 2) {
          const float2 t = fmul2(make_float2(__uint_as_float(o[e]), __uint_as_float(o[e + 1])), a2);
          o[e] = __float_as_uint(t.x);
          o[e + 1] = __float_as_uint(t.y);
        }
        tmem_st_32x16(t_o + c * 16, o);
      }
      tmem_wait_st();
    }
  }
  const float2 sc2 = make_float2(sc, sc), nm2 = make_float2(-m_ref, -m_ref);
  float2 sum[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
  auto emit = [&](uint32_t (&x)[32], int c) {
    // 16 keys at a time keeps the live set small: 8 packed registers, two 16-byte stores
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      uint32_t pk[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        const int i0 = 16 * q + 2 * e;
        float2 y = ffma2(make_float2(__uint_as_float(x[i0]), __uint_as_float(x[i0 + 1])), sc2, nm2);
        if (!kMasked && e < kPolyPairs) {
          y = exp2_poly2(y);
        } else {
          y.x = ex2(y.x);  // masked keys: exp2(-inf) = 0
          y.y = ex2(y.y);
        }
        sum[e & 1] = fadd2(sum[e & 1], y);
        pk[e] = pack_bf16(y.x, y.y);
      }
      // chunk c covers keys [32c, 32c+32) of this half: 16-byte chunks 4c .. 4c+3 of the swizzled row
#pragma unroll
      for (int g = 0; g < 2; ++g) {
        const int chunk = (c * 4 + 2 * q + g) ^ (r & 7);
        *reinterpret_cast<uint4*>(prow + chunk * 16) =
            make_uint4(pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
      }
    }
  };
  emit(a, 0);
  emit(b, 1);
  l_run += (sum[0].x + sum[0].y) + (sum[1].x + sum[1].y);
}

}  // namespace

template <int kPolyPairs>
__global__ void __launch_bounds__(kThreads, 1)
vit_attention_kernel(const __grid_constant__ CUtensorMap map_qkv, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_addr = smem_u32(smem_raw);
  const uint32_t base = (raw_addr + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_addr);
  const uint32_t sQ = base;                                    // 2 x 16 KB
  const uint32_t sKV = sQ + 2 * kQBytes;                       // stages x (K 16 KB, V 16 KB)
  const uint32_t sP = sKV + kKvStages * kKvBytes;              // 2 x 32 KB
  uint8_t* sP_gen = smem + 2 * kQBytes + kKvStages * kKvBytes;
  float2* xch = reinterpret_cast<float2*>(sP_gen + 2 * kPBytes);  // [2 wg][2 halves][128 rows] (m, l)
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP_gen + 2 * kPBytes + kXchBytes);
  const uint32_t bar0 = smem_u32(bars);
  const uint32_t q_full = bar0;
  auto kv_full = [&](int s) { return bar0 + 8u * (1 + s); };
  auto kv_empty = [&](int s) { return bar0 + 8u * (1 + kKvStages + s); };
  auto s_full = [&](int w) { return bar0 + 8u * (1 + 2 * kKvStages + w); };
  auto s_free = [&](int w) { return bar0 + 8u * (3 + 2 * kKvStages + w); };
  auto p_full = [&](int w, int h) { return bar0 + 8u * (5 + 2 * kKvStages + 2 * w + h); };
  auto o_ready = [&](int w, int h) { return bar0 + 8u * (9 + 2 * kKvStages + 2 * w + h); };
  auto o_full = [&](int w) { return bar0 + 8u * (13 + 2 * kKvStages + w); };
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15 + 2 * kKvStages);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int C = p.heads * kHd;
  const int q_pairs = (p.T + 2 * kTile - 1) / (2 * kTile);
  const int qp = blockIdx.x % q_pairs;
  const int bh = blockIdx.x / q_pairs;
  const int head = bh % p.heads, img = bh / p.heads;
  const int q0 = qp * 2 * kTile;
  const int nw = (q0 + kTile < p.T) ? 2 : 1;
  const int nkv = (p.T + kTile - 1) / kTile;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&map_qkv);
    mbar_init(q_full, 1);
    for (int s = 0; s < kKvStages; ++s) {
      mbar_init(kv_full(s), 1);
      mbar_init(kv_empty(s), nw);
    }
    for (int w = 0; w < 2; ++w) {
      mbar_init(s_full(w), 1);
      mbar_init(s_free(w), 256);
      mbar_init(o_full(w), 1);
      for (int h = 0; h < 2; ++h) {
        mbar_init(p_full(w, h), 128);
        mbar_init(o_ready(w, h), 1);
      }
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    tmem_alloc(smem_u32(tmem_slot), kTmemCols);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ------------------------------------------------------------ TMA producer
    mbar_arrive_expect_tx(q_full, 2 * kQBytes);
    tma_load_3d(sQ, &map_qkv, q_full, head * kHd, q0, img);
    tma_load_3d(sQ + kQBytes, &map_qkv, q_full, head * kHd, q0 + kTile, img);
    for (int j = 0; j < nkv; ++j) {
      const int s = j % kKvStages;
      const uint32_t ph = (j / kKvStages) & 1;
      mbar_wait_sleep(kv_empty(s), ph ^ 1, 256);
      mbar_arrive_expect_tx(kv_full(s), kKvBytes);
      tma_load_3d(sKV + s * kKvBytes, &map_qkv, kv_full(s), C + head * kHd, j * kTile, img);
      tma_load_3d(sKV + s * kKvBytes + kKvBytes / 2, &map_qkv, kv_full(s), 2 * C + head * kHd, j * kTile,
                  img);
    }
  } else if ((warp == 1 || warp == 2) && lane == 0 && warp - 1 < nw) {
    // ------------------------------------------------------------ MMA issuer of query tile w
    const int w = warp - 1;
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128);
    constexpr uint32_t idesc_o = idesc_pv();
    const uint64_t dq = umma_desc_k_sw128(sQ + w * kQBytes);
    auto issue_s = [&](int s) {
      const uint64_t db = umma_desc_k_sw128(sKV + s * kKvBytes);
#pragma unroll
      for (int k = 0; k < kHd / 16; ++k)
        tc_mma_bf16(tmem_base + w * 128, dq + 2 * k, db + 2 * k, idesc_s, k != 0);
      tc_commit(s_full(w));
    };
    mbar_wait(q_full, 0);
    mbar_wait(kv_full(0), 0);
    tc_fence_after();
    issue_s(0);
    for (int j = 0; j < nkv; ++j) {
      const int s = j % kKvStages;
      if (j + 1 < nkv) {
        const int s1 = (j + 1) % kKvStages;
        mbar_wait_sleep(kv_full(s1), ((j + 1) / kKvStages) & 1, 32);
        mbar_wait_sleep(s_free(w), j & 1, 32);   // S_w(j) is in the softmax threads' registers
        tc_fence_after();
        issue_s(s1);
      }
      const uint64_t dv = umma_desc_k_sw128(sKV + s * kKvBytes + kKvBytes / 2);
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        mbar_wait_sleep(p_full(w, h), j & 1, 32);
        tc_fence_after();
        // O_{w,h} += P_{w,h} (128 x 64 keys, K-major) * V_j rows [64h, 64h+64) (MN-major)
        const uint64_t dp = umma_desc_k_sw128(sP + w * kPBytes + h * (kPBytes / 2));
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc_mma_bf16(tmem_base + 256 + (2 * w + h) * 64, dp + 2 * k, dv + 128 * (4 * h + k), idesc_o,
                      (j | k) != 0);
        tc_commit(o_ready(w, h));
      }
      if (j + 1 == nkv) tc_commit(o_full(w));
      tc_commit(kv_empty(s));
    }
  } else if (warp >= 4 && (warp - 4) / 8 < nw) {
    // ------------------------------------------------------------ softmax warpgroups
    const int w = (warp - 4) >> 3;
    const int h = ((warp - 4) >> 2) & 1;  // key half
    const int quarter = warp & 3;         // TMEM lane quarter a warp may touch = warp id % 4
    const int r = quarter * 32 + lane;    // row inside the query tile == TMEM lane
    const uint32_t lane_base = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16);
    const uint32_t t_s = lane_base + w * 128 + h * 64;
    const uint32_t t_o = lane_base + 256 + (2 * w + h) * 64;
    const float sc = p.scale * 1.4426950408889634f;  // scores are handled in the exp2 domain
    uint8_t* prow = sP_gen + w * kPBytes + h * (kPBytes / 2) + r * 128;
    float m_ref = -1e30f, l_run = 0.f;
    for (int j = 0; j < nkv; ++j) {
      // keys of this half tile inside the image
      const int valid = min(kTile / 2, p.T - j * kTile - h * (kTile / 2));
      mbar_wait_sleep(s_full(w), j & 1, 32);
      tc_fence_after();
      if (valid == kTile / 2)
        softmax_half<false, kPolyPairs>(t_s, t_o, prow, r, valid, sc, j > 0, s_free(w), o_ready(w, h), (j - 1) & 1, m_ref, l_run);
      else
        softmax_half<true, kPolyPairs>(t_s, t_o, prow, r, valid, sc, j > 0, s_free(w), o_ready(w, h), (j - 1) & 1, m_ref, l_run);
      fence_proxy_async_smem();
      tc_fence_before();
      mbar_arrive(p_full(w, h));
    }
    // merge the two key halves of the row (split-K): exchange (m, l) through shared memory
    xch[(w * 2 + h) * kTile + r] = make_float2(m_ref, l_run);
    asm volatile("bar.sync %0, 256;" ::"r"(1 + w) : "memory");
    const float2 other = xch[(w * 2 + (h ^ 1)) * kTile + r];
    const float m_all = fmaxf(m_ref, other.x);
    const float a_me = ex2(m_ref - m_all), a_ot = ex2(other.x - m_all);
    const float inv = 1.0f / (l_run * a_me + other.y * a_ot);
    const float c0 = (h == 0 ? a_me : a_ot) * inv, c1 = (h == 0 ? a_ot : a_me) * inv;
    mbar_wait_sleep(o_full(w), 0, 32);
    tc_fence_after();
    // this thread finishes output columns [32h, 32h+32) of its row from both partial accumulators
    uint32_t o0[32], o1[32];
    tmem_ld_32x32(lane_base + 256 + (2 * w) * 64 + h * 32, o0);
    tmem_ld_32x32(lane_base + 256 + (2 * w + 1) * 64 + h * 32, o1);
    tmem_wait_ld();
    // The 32 rows x 64 columns of a lane quarter (both halves) meet in a swizzled shared-memory tile
    // -- a slice of this warpgroup's P buffer, free now that every P V has completed -- and leave as
    // whole 128-byte lines (8 lanes per row): a row-per-thread store touches 32 lines per instruction
    uint8_t* stg = sP_gen + w * kPBytes + quarter * 4096;
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      uint32_t pk[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i0 = 8 * g + 2 * e;
        pk[e] = pack_bf16(fmaf(__uint_as_float(o0[i0]), c0, __uint_as_float(o1[i0]) * c1),
                          fmaf(__uint_as_float(o0[i0 + 1]), c0, __uint_as_float(o1[i0 + 1]) * c1));
      }
      *reinterpret_cast<uint4*>(stg + lane * 128 + (((h * 4 + g) ^ (lane & 7)) << 4)) =
          make_uint4(pk[0], pk[1], pk[2], pk[3]);
    }
    asm volatile("bar.sync %0, 256;" ::"r"(1 + w) : "memory");
    {
      const int tc = lane & 7;
      __nv_bfloat16* dst = p.out + (size_t)img * p.T * C + head * kHd + tc * 8;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int rr = 16 * h + (lane >> 3) + 4 * i;   // this warp stores rows [16h, 16h+16) of the quarter
        const int tt = q0 + w * kTile + quarter * 32 + rr;
        if (tt < p.T) {
          const uint4 v = *reinterpret_cast<const uint4*>(stg + rr * 128 + ((tc ^ (rr & 7)) << 4));
          *reinterpret_cast<uint4*>(dst + (size_t)tt * C) = v;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, kTmemCols);
}

int attn_smem_bytes() { return kSmemBytes; }

template <int kPolyPairs>
static cudaError_t launch_variant(const CUtensorMap& map_qkv, const AttnParams& p, int grid, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(vit_attention_kernel<kPolyPairs>,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  vit_attention_kernel<kPolyPairs><<<grid, kThreads, kSmemBytes, stream>>>(map_qkv, p);
  return cudaGetLastError();
}

cudaError_t launch_attention(const CUtensorMap& map_qkv, const AttnParams& p, cudaStream_t stream) {
  // pairs of every 8 (16 keys) whose exp2 runs on the FMA pipe; VC_ATTN_POLY overrides (tuning knob)
  static int poly = -1;
  if (poly < 0) {
    const char* env = getenv("VC_ATTN_POLY");
    poly = env ? atoi(env) : 0;
  }
  if (p.B <= 0 || p.T <= 0) return cudaSuccess;
  const int q_pairs = (p.T + 2 * kTile - 1) / (2 * kTile);
  const int grid = p.B * p.heads * q_pairs;
  switch (poly) {
    case 1: return launch_variant<1>(map_qkv, p, grid, stream);
    case 2: return launch_variant<2>(map_qkv, p, grid, stream);
    case 4: return launch_variant<4>(map_qkv, p, grid, stream);
    case 3: return launch_variant<3>(map_qkv, p, grid, stream);
    default: return launch_variant<0>(map_qkv, p, grid, stream);
  }
}

}  // namespace vc
