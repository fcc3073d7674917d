// Training-time similarity and hard-negative mining on patch tokens.
// Reference: vit_colmap/dataloader/homography_utils.py warp_patch_tokens :10-92,
// compute_feature_similarity :219-253; vit_colmap/dataloader/training_sampler.py
// select_invariant_points :56-181, sample_features_at_coords :183-211,
// generate_hard_negatives :452-534.
//
// Feature maps stay in the backbone's token-major layout [B, H*W, C] fp32 (the reference's NCHW view
// is converted once by nchw_to_tokens_kernel when a caller holds one).  All arithmetic is fp32: the
// selections are index lists, and an index list only equals the reference's when the similarities
// agree to ~1e-6, which bf16 operands would not give.
//   warp_similarity_kernel : warp per target cell; source position through the inverse homography,
//                            bilinear gather with zero padding (grid_sample align_corners=False),
//                            cosine with the target cell -- both maps are read once (HBM-bound).
//   topk_rows_kernel       : CTA per row, bitonic sort of 64-bit keys (value desc, index asc) in
//                            shared memory, plus the ascending list of entries >= threshold.
//   hard_negatives_kernel  : CTA = 32 positives of one image against every cell: fp32 SIMT GEMM tiles
//                            (8 x 4 per thread) through shared memory, distance mask, per-row top-n kept
//                            in registers across the lanes of a warp -- the K x HW matrix is never stored.
//   sample_bilinear_kernel : warp per point, the reference's (x / (W-1)) * 2 - 1 normalisation fed to
//                            an align_corners=False grid_sample with zero padding.
#include "sampler_kernels.cuh"
#include "vc_launch.cuh"
#include "vc_ptx.cuh"

namespace vc {

namespace {

__device__ __forceinline__ float4 ldf4(const float* row, int v) {
  return __ldg(reinterpret_cast<const float4*>(row) + v);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Blackwell packed fp32 pair FMA (FFMA2): two multiply-adds per instruction -- a 3-register FFMA issues
// every other cycle per SM sub-partition, so the full fp32 rate needs the packed form.
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)),
        "l"(*reinterpret_cast<uint64_t*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}

__device__ __forceinline__ unsigned orderable(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// Bilinear taps of F.grid_sample(mode="bilinear", padding_mode="zeros", align_corners=False) at the
// un-normalised source position (ix, iy): weights as ATen forms them (w = x - floor x, e = 1 - w,
// n = y - floor y, s = 1 - n; nw = e s, ne = w s, sw = e n, se = w n); taps outside the map are dropped.
struct Taps {
  int idx[4];
  float w[4];
};
__device__ __forceinline__ Taps zero_pad_taps(float ix, float iy, int H, int W) {
  const float x0 = floorf(ix), y0 = floorf(iy);
  const float ww = __fsub_rn(ix, x0), ee = __fsub_rn(1.0f, ww);
  const float nn = __fsub_rn(iy, y0), ss = __fsub_rn(1.0f, nn);
  Taps t;
  t.w[0] = __fmul_rn(ee, ss);
  t.w[1] = __fmul_rn(ww, ss);
  t.w[2] = __fmul_rn(ee, nn);
  t.w[3] = __fmul_rn(ww, nn);
  // positions far outside the map (or NaN) have no tap; the clamp keeps the int conversion defined
  const float xc = fminf(fmaxf(x0, -2.0f), (float)W + 1.0f), yc = fminf(fmaxf(y0, -2.0f), (float)H + 1.0f);
  const bool finite = (ix == ix) && (iy == iy);
  const int xi[2] = {(int)xc, (int)xc + 1}, yi[2] = {(int)yc, (int)yc + 1};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int yy = yi[k >> 1], xx = xi[k & 1];
    const bool in = finite && xc == x0 && yc == y0 && yy >= 0 && yy < H && xx >= 0 && xx < W;
    t.idx[k] = in ? yy * W + xx : -1;
  }
  return t;
}

// align_corners=False un-normalisation of ATen: ((g + 1) * size - 1) / 2
__device__ __forceinline__ float unnormalize(float g, int size) {
  return __fdiv_rn(__fsub_rn(__fmul_rn(__fadd_rn(g, 1.0f), (float)size), 1.0f), 2.0f);
}

template <int KV>
__device__ __forceinline__ void gather_taps(const float* img, const Taps& t, int C, int lane, float4 (&acc)[KV]) {
  const int nvec = C >> 2;
#pragma unroll
  for (int j = 0; j < KV; ++j) acc[j] = make_float4(0, 0, 0, 0);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    if (t.idx[s] < 0) continue;  // warp-uniform
    const float* row = img + (size_t)t.idx[s] * C;
    const float w = t.w[s];
#pragma unroll
    for (int j = 0; j < KV; ++j) {
      const int v = lane + 32 * j;
      if (v < nvec) {
        const float4 x = ldf4(row, v);
        acc[j].x = __fadd_rn(acc[j].x, __fmul_rn(x.x, w));
        acc[j].y = __fadd_rn(acc[j].y, __fmul_rn(x.y, w));
        acc[j].z = __fadd_rn(acc[j].z, __fmul_rn(x.z, w));
        acc[j].w = __fadd_rn(acc[j].w, __fmul_rn(x.w, w));
      }
    }
  }
}

// ------------------------------------------------------------------ warp + cosine similarity
template <int KV>
__global__ void __launch_bounds__(256) warp_similarity_kernel(const WarpSimParams p) {
  const int lane = threadIdx.x & 31;
  const long long wg = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int HW = p.H * p.W;
  griddep_wait();
  if (wg >= (long long)p.B * HW) return;
  const int b = (int)(wg / HW), cell = (int)(wg - (long long)b * HW);
  const int cy = cell / p.W, cx = cell - cy * p.W;
  // target patch centre in pixels, (i + 0.5) * patch_size (homography_utils.py:43-48)
  const float px = __fmul_rn(__fadd_rn((float)cx, 0.5f), p.patch_size);
  const float py = __fmul_rn(__fadd_rn((float)cy, 0.5f), p.patch_size);
  const float* h = p.h_inv + (size_t)b * 9;
  // H_inv @ (x, y, 1) (:61-63), homogeneous divide with w clamped at 1e-8 (:66-68)
  const float sx = __fadd_rn(__fadd_rn(__fmul_rn(h[0], px), __fmul_rn(h[1], py)), h[2]);
  const float sy = __fadd_rn(__fadd_rn(__fmul_rn(h[3], px), __fmul_rn(h[4], py)), h[5]);
  const float sw = fmaxf(__fadd_rn(__fadd_rn(__fmul_rn(h[6], px), __fmul_rn(h[7], py)), h[8]), 1e-8f);
  const float x = __fdiv_rn(sx, sw), y = __fdiv_rn(sy, sw);
  // normalise by the image size (:76-78), then grid_sample's own un-normalisation
  const float gx = __fsub_rn(__fmul_rn(__fdiv_rn(x, p.img_w), 2.0f), 1.0f);
  const float gy = __fsub_rn(__fmul_rn(__fdiv_rn(y, p.img_h), 2.0f), 1.0f);
  const Taps t = zero_pad_taps(unnormalize(gx, p.W), unnormalize(gy, p.H), p.H, p.W);
  float4 wv[KV];
  gather_taps<KV>(p.tokens1 + (int64_t)b * p.image_stride, t, p.C, lane, wv);
  const float* row2 = p.tokens2 + (int64_t)b * p.image_stride + (size_t)cell * p.C;
  const int nvec = p.C >> 2;
  float dot = 0.f, n1 = 0.f, n2 = 0.f;
#pragma unroll
  for (int j = 0; j < KV; ++j) {
    const int v = lane + 32 * j;
    if (v < nvec) {
      const float4 f = ldf4(row2, v);
      const float4 w = wv[j];
      dot += w.x * f.x + w.y * f.y + w.z * f.z + w.w * f.w;
      n1 += w.x * w.x + w.y * w.y + w.z * w.z + w.w * w.w;
      n2 += f.x * f.x + f.y * f.y + f.z * f.z + f.w * f.w;
    }
  }
  dot = warp_sum(dot);
  n1 = warp_sum(n1);
  n2 = warp_sum(n2);
  if (lane == 0) {
    // F.normalize(p=2, dim=1) of both maps (eps 1e-12), then the channel sum of products (:246-251)
    const float den = __fmul_rn(fmaxf(sqrtf(n1), 1e-12f), fmaxf(sqrtf(n2), 1e-12f));
    p.similarity[(size_t)b * HW + cell] = __fdiv_rn(dot, den);
  }
}

// ------------------------------------------------------------------ bilinear sampling at (x, y)
template <int KV>
__global__ void __launch_bounds__(256) sample_bilinear_kernel(const SampleBilinearParams p) {
  const int lane = threadIdx.x & 31;
  const long long wg = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  griddep_wait();
  if (wg >= (long long)p.B * p.n_pts) return;
  const int b = (int)(wg / p.n_pts);
  const float2 c = reinterpret_cast<const float2*>(p.xy)[wg];
  // training_sampler.py:199-201: x / (W - 1) * 2 - 1, handed to an align_corners=False grid_sample (:207)
  const float gx = __fsub_rn(__fmul_rn(__fdiv_rn(c.x, (float)(p.W - 1)), 2.0f), 1.0f);
  const float gy = __fsub_rn(__fmul_rn(__fdiv_rn(c.y, (float)(p.H - 1)), 2.0f), 1.0f);
  const Taps t = zero_pad_taps(unnormalize(gx, p.W), unnormalize(gy, p.H), p.H, p.W);
  float4 acc[KV];
  gather_taps<KV>(p.tokens + (int64_t)b * p.image_stride, t, p.C, lane, acc);
  float* out = p.out + (size_t)wg * p.C;
  const int nvec = p.C >> 2;
#pragma unroll
  for (int j = 0; j < KV; ++j) {
    const int v = lane + 32 * j;
    if (v < nvec) reinterpret_cast<float4*>(out)[v] = acc[j];
  }
}

// ------------------------------------------------------------------ row-wise ordered top-k
constexpr int kTopkThreads = 1024;

__global__ void __launch_bounds__(kTopkThreads, 1) topk_rows_kernel(const TopkParams p) {
  extern __shared__ __align__(16) unsigned long long keys[];
  __shared__ int s_scan[32];
  __shared__ int s_base;
  griddep_wait();
  const int r = blockIdx.x;
  const float* val = p.values + (size_t)r * p.n;
  int npad = 32;
  while (npad < p.n) npad <<= 1;

  // ascending list of the entries >= threshold (torch.nonzero order, training_sampler.py:108-110)
  if (p.above != nullptr || p.n_above != nullptr) {
    if (threadIdx.x == 0) s_base = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i0 = 0; i0 < p.n; i0 += kTopkThreads) {
      const int i = i0 + threadIdx.x;
      const bool hit = i < p.n && val[i] >= p.threshold;
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (lane == 0) s_scan[warp] = __popc(m);
      __syncthreads();
      int off = s_base;
      for (int w = 0; w < warp; ++w) off += s_scan[w];
      if (hit && p.above != nullptr) p.above[(size_t)r * p.n + off + __popc(m & ((1u << lane) - 1))] = i;
      __syncthreads();
      if (threadIdx.x == 0) {
        int tot = 0;
        for (int w = 0; w < kTopkThreads / 32; ++w) tot += s_scan[w];
        s_base += tot;
      }
      __syncthreads();
    }
    if (threadIdx.x == 0 && p.n_above != nullptr) p.n_above[r] = s_base;
  }
  if (p.order == nullptr || p.k <= 0) return;

  // value descending, lower index first among equal values
  for (int i = threadIdx.x; i < npad; i += kTopkThreads)
    keys[i] = i < p.n ? (((unsigned long long)orderable(val[i]) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)i))
                      : 0ull;
  __syncthreads();
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < npad; i += kTopkThreads) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const unsigned long long a = keys[i], c = keys[ixj];
          const bool desc = (i & k) == 0;
          if (desc ? (a < c) : (a > c)) {
            keys[i] = c;
            keys[ixj] = a;
          }
        }
      }
      __syncthreads();
    }
  }
  const int kk = p.k < p.n ? p.k : p.n;
  for (int i = threadIdx.x; i < p.k; i += kTopkThreads)
    p.order[(size_t)r * p.k + i] = i < kk ? (int)(0xFFFFFFFFu - (unsigned)(keys[i] & 0xFFFFFFFFull)) : -1;
}

// ------------------------------------------------------------------ hard negatives
// fp32 SIMT GEMM tile: CTA = 32 positives x 128 cells, 4 warps; warp = 8 positives, lane = 4 cells (lane + 32 j), so a
// thread holds an 8 x 4 accumulator tile and per 4 channels issues 128 FMAs against 12 shared-memory loads
// (8 broadcast reads of the positives, 4 vector reads of the cells), as 64 packed FFMA2: FMA-bound.  Both
// operands are staged per 32-channel chunk (20 KB of shared memory per CTA, several CTAs per SM).
constexpr int kHnRows = 32;      // positives per CTA
constexpr int kHnRowsW = 8;      // positives per warp
constexpr int kHnCells = 128;    // cells per tile (4 per lane)
constexpr int kHnChunk = 32;     // channels staged per step
constexpr int kHnThreads = 128;

// 16-byte global -> shared copy that bypasses the registers (LDGSTS); src_bytes = 0 writes zeros
__device__ __forceinline__ void cp_async16(void* dst_smem, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst_smem)), "l"(src), "r"(src_bytes)
               : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

__global__ void __launch_bounds__(kHnThreads, 4) hard_negatives_kernel(const HardNegParams p) {
  // both operand chunks are double-buffered: the copies of chunk s + 1 (cp.async, no registers) run under the
  // FMAs of chunk s -- with plain loads 29 % of the kernel's samples sat on the staging stores' scoreboard
  __shared__ __align__(16) float s_pos[2][kHnRows][kHnChunk];
  __shared__ __align__(16) float s_cell[2][kHnCells][kHnChunk + 4];  // +4: a quarter-warp's 16-byte reads hit 32 banks
  __shared__ float s_ss[4][kHnCells];   // per-warp partial squared norms of the tile's cells
  __shared__ float s_pn[kHnRows];       // norms of the positives
  __shared__ float2 s_pxy[kHnRows];     // their (x, y)
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int tiles_k = (p.K + kHnRows - 1) / kHnRows;
  const int b = blockIdx.x / tiles_k;
  const int k0 = (blockIdx.x - b * tiles_k) * kHnRows;
  const int HW = p.H * p.W;
  griddep_wait();
  const float* img = p.tokens + (int64_t)b * p.image_stride;
  const float* pos = p.pos_feat + (size_t)b * p.K * p.C;

  // norms of this CTA's positives (F.normalize eps 1e-12, training_sampler.py:479-481): warp per 8 rows
  for (int q = 0; q < kHnRowsW; ++q) {
    const int k = k0 + kHnRowsW * warp + q;
    float ss = 0.f;
    if (k < p.K)
      for (int v = lane; v < (p.C >> 2); v += 32) {
        const float4 x = ldf4(pos + (size_t)k * p.C, v);
        ss += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
      }
    ss = warp_sum(ss);
    if (lane == 0) s_pn[kHnRowsW * warp + q] = fmaxf(sqrtf(ss), 1e-12f);
  }
  if (threadIdx.x < kHnRows) {
    const int k = k0 + threadIdx.x;
    s_pxy[threadIdx.x] =
        k < p.K ? reinterpret_cast<const float2*>(p.pos_xy)[(size_t)b * p.K + k] : make_float2(0.f, 0.f);
  }
  unsigned long long best[kHnRowsW];  // lane i holds entry i of the row's descending top-n list
#pragma unroll
  for (int q = 0; q < kHnRowsW; ++q) best[q] = 0ull;

  auto stage = [&](int buf, int j0, int c0) {
    for (int i = threadIdx.x; i < kHnRows * (kHnChunk >> 2); i += kHnThreads) {
      const int r = i / (kHnChunk >> 2), v = i - r * (kHnChunk >> 2);
      const bool ok = k0 + r < p.K && c0 + 4 * v < p.C;
      cp_async16(&s_pos[buf][r][4 * v], ok ? pos + (size_t)(k0 + r) * p.C + c0 + 4 * v : pos, ok ? 16 : 0);
    }
    // cells: eight lanes copy the 128 contiguous bytes of one cell's chunk into one row
    for (int i = threadIdx.x; i < kHnCells * (kHnChunk >> 2); i += kHnThreads) {
      const int jc = i / (kHnChunk >> 2), v = i - jc * (kHnChunk >> 2);
      const bool ok = j0 + jc < HW && c0 + 4 * v < p.C;
      cp_async16(&s_cell[buf][jc][4 * v], ok ? img + (size_t)(j0 + jc) * p.C + c0 + 4 * v : img, ok ? 16 : 0);
    }
    cp_async_commit();
  };
  int buf = 0;
  stage(0, 0, 0);
  for (int j0 = 0; j0 < HW; j0 += kHnCells) {
    float2 acc[kHnRowsW][4];  // (even-channel sum, odd-channel sum) of positive q x cell j
#pragma unroll
    for (int q = 0; q < kHnRowsW; ++q)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[q][e] = make_float2(0.f, 0.f);
    float ssf[4] = {0.f, 0.f, 0.f, 0.f};
    for (int c0 = 0; c0 < p.C; c0 += kHnChunk, buf ^= 1) {
      // request the next chunk (of this cell tile, or the first of the next tile) into the other buffer, which
      // every thread left at the barrier that ended the previous iteration
      const bool tile_ends = c0 + kHnChunk >= p.C;
      const int nj0 = tile_ends ? j0 + kHnCells : j0, nc0 = tile_ends ? 0 : c0 + kHnChunk;
      if (nj0 < HW) {
        stage(buf ^ 1, nj0, nc0);
        cp_async_wait<1>();  // all but the newest group: this chunk has landed
      } else {
        cp_async_wait<0>();
      }
      __syncthreads();
#pragma unroll 1
      for (int c = 0; c < kHnChunk; c += 4) {
        float4 f[4];  // f[j] = channels c .. c+3 of this lane's cell j (cells lane, lane + 32, lane + 64, lane + 96)
#pragma unroll
        for (int j = 0; j < 4; ++j) f[j] = *reinterpret_cast<const float4*>(&s_cell[buf][lane + 32 * j][c]);
        if (((c >> 2) & 3) == warp) {  // warp-uniform: the four warps share the norm work by channel group
#pragma unroll
          for (int j = 0; j < 4; ++j)
            ssf[j] = fmaf(f[j].x, f[j].x, fmaf(f[j].y, f[j].y, fmaf(f[j].z, f[j].z, fmaf(f[j].w, f[j].w, ssf[j]))));
        }
#pragma unroll
        for (int q = 0; q < kHnRowsW; ++q) {
          const float4 pv = *reinterpret_cast<const float4*>(&s_pos[buf][kHnRowsW * warp + q][c]);
#pragma unroll
          for (int j = 0; j < 4; ++j)  // operand pairs are the adjacent registers of the 16-byte loads
            acc[q][j] = ffma2(make_float2(pv.z, pv.w), make_float2(f[j].z, f[j].w),
                              ffma2(make_float2(pv.x, pv.y), make_float2(f[j].x, f[j].y), acc[q][j]));
        }
      }
      __syncthreads();  // this buffer is the target of the next iteration's copies
    }
    // cell norms: sum of the four warps' partials
#pragma unroll
    for (int e = 0; e < 4; ++e) s_ss[warp][lane + 32 * e] = ssf[e];
    __syncthreads();
    float nf[4], cxf[4], cyf[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int jc = lane + 32 * e;
      nf[e] = fmaxf(sqrtf(s_ss[0][jc] + s_ss[1][jc] + s_ss[2][jc] + s_ss[3][jc]), 1e-12f);
      const int cell = j0 + jc, cy = cell / p.W;
      cxf[e] = (float)(cell - cy * p.W);
      cyf[e] = (float)cy;
    }
#pragma unroll
    for (int q = 0; q < kHnRowsW; ++q) {
      const bool row_ok = (k0 + kHnRowsW * warp + q) < p.K;
      const float np_ = s_pn[kHnRowsW * warp + q];
      const float2 pxy = s_pxy[kHnRowsW * warp + q];
      unsigned long long thr = __shfl_sync(0xffffffffu, best[q], p.n_hard - 1);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int cell = j0 + lane + 32 * e;
        // cosine with the positive; cells closer than the radius are not candidates (:497-503)
        const float sim = __fdiv_rn(__fadd_rn(acc[q][e].x, acc[q][e].y), __fmul_rn(nf[e], np_));
        const float dx = __fsub_rn(cxf[e], pxy.x), dy = __fsub_rn(cyf[e], pxy.y);
        const float dist = __fsqrt_rn(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
        const bool cand = row_ok && cell < HW && dist >= p.radius && sim == sim;
        const unsigned long long key =
            cand ? (((unsigned long long)orderable(sim) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)cell)) : 0ull;
        unsigned m = __ballot_sync(0xffffffffu, key > thr);
        while (m) {
          const int src = __ffs(m) - 1;
          m &= m - 1;
          const unsigned long long kin = __shfl_sync(0xffffffffu, key, src);
          if (kin > thr) {  // warp-uniform
            const int at = __popc(__ballot_sync(0xffffffffu, best[q] > kin && lane < p.n_hard));
            const unsigned long long up = __shfl_up_sync(0xffffffffu, best[q], 1);
            if (lane == at) best[q] = kin;
            else if (lane > at) best[q] = up;
            thr = __shfl_sync(0xffffffffu, best[q], p.n_hard - 1);
          }
        }
      }
    }
  }
#pragma unroll
  for (int q = 0; q < kHnRowsW; ++q) {
    const int k = k0 + kHnRowsW * warp + q;
    if (k >= p.K) continue;  // warp-uniform
    const bool have = lane < p.n_hard && best[q] != 0ull;
    if (lane < p.n_hard)
      p.idx[((size_t)b * p.K + k) * p.n_hard + lane] = have ? (int)(0xFFFFFFFFu - (unsigned)(best[q] & 0xFFFFFFFFull)) : -1;
    const unsigned cnt = __ballot_sync(0xffffffffu, have);
    if (lane == 0) p.counts[(size_t)b * p.K + k] = __popc(cnt);
  }
}

// ------------------------------------------------------------------ NCHW -> token-major
__global__ void __launch_bounds__(256) nchw_to_tokens_kernel(const float* __restrict__ in, float* __restrict__ out,
                                                             int C, int HW) {
  __shared__ float tile[32][33];
  griddep_wait();
  const int b = blockIdx.z;
  const int c0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  const float* src = in + (size_t)b * C * HW;
  float* dst = out + (size_t)b * C * HW;
  for (int i = ty; i < 32; i += 8) {
    const int c = c0 + i, j = j0 + tx;
    tile[i][tx] = (c < C && j < HW) ? src[(size_t)c * HW + j] : 0.f;
  }
  __syncthreads();
  for (int i = ty; i < 32; i += 8) {
    const int j = j0 + i, c = c0 + tx;
    if (j < HW && c < C) dst[(size_t)j * C + c] = tile[tx][i];
  }
}

template <typename P, typename F1, typename F2, typename F3, typename F4, typename F5, typename F6>
cudaError_t dispatch_kv(int C, F1 f1, F2 f2, F3 f3, F4 f4, F5 f6, F6 f8) {
  const int kv = (C / 4 + 31) / 32;
  if (kv <= 1) return f1();
  if (kv <= 2) return f2();
  if (kv <= 3) return f3();
  if (kv <= 4) return f4();
  if (kv <= 6) return f6();
  return f8();
}

}  // namespace

cudaError_t launch_warp_similarity(const WarpSimParams& p, cudaStream_t stream) {
  const long long warps = (long long)p.B * p.H * p.W;
  if (warps <= 0) return cudaSuccess;
  const dim3 grid((unsigned)((warps * 32 + 255) / 256)), block(256);
#define VC_WS(KV) [&] { return launch_pdl(warp_similarity_kernel<KV>, grid, block, 0, stream, p); }
  return dispatch_kv<WarpSimParams>(p.C, VC_WS(1), VC_WS(2), VC_WS(3), VC_WS(4), VC_WS(6), VC_WS(8));
#undef VC_WS
}

cudaError_t launch_sample_bilinear(const SampleBilinearParams& p, cudaStream_t stream) {
  const long long warps = (long long)p.B * p.n_pts;
  if (warps <= 0) return cudaSuccess;
  const dim3 grid((unsigned)((warps * 32 + 255) / 256)), block(256);
#define VC_SB(KV) [&] { return launch_pdl(sample_bilinear_kernel<KV>, grid, block, 0, stream, p); }
  return dispatch_kv<SampleBilinearParams>(p.C, VC_SB(1), VC_SB(2), VC_SB(3), VC_SB(4), VC_SB(6), VC_SB(8));
#undef VC_SB
}

int topk_smem_bytes(int n) {
  int npad = 32;
  while (npad < n) npad <<= 1;
  return npad * 8;
}

cudaError_t launch_topk_rows(const TopkParams& p, cudaStream_t stream) {
  if (p.n_rows <= 0 || p.n <= 0) return cudaSuccess;
  const int smem = topk_smem_bytes(p.n);
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(topk_rows_kernel, smem, opt_in); e != cudaSuccess) return e;
  return launch_pdl(topk_rows_kernel, dim3(p.n_rows), dim3(kTopkThreads), smem, stream, p);
}

cudaError_t launch_hard_negatives(const HardNegParams& p, cudaStream_t stream) {
  if (p.B <= 0 || p.K <= 0) return cudaSuccess;
  const int tiles_k = (p.K + kHnRows - 1) / kHnRows;
  return launch_pdl(hard_negatives_kernel, dim3(p.B * tiles_k), dim3(kHnThreads), 0, stream, p);
}

cudaError_t launch_nchw_to_tokens(const float* in, float* out, int B, int C, int HW, cudaStream_t stream) {
  if (B <= 0 || C <= 0 || HW <= 0) return cudaSuccess;
  const dim3 grid((HW + 31) / 32, (C + 31) / 32, B);
  return launch_pdl(nchw_to_tokens_kernel, grid, dim3(256), 0, stream, in, out, C, HW);
}

}  // namespace vc
