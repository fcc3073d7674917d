// Keypoint selection on the score map: per-bin top-k, global top-K, greedy NMS, ordered output.
// Reference: vit_colmap/features/vit_extractor.py _spatial_binning_selection :404-485,
// _apply_nms :500-543 (python loops with one device sync per kept keypoint).
//
// One CTA per image, the whole problem lives in shared memory (<= 16384 grid cells):
//   1. rank of every cell inside its bin (score descending, lower in-bin index wins ties); the
//      top `per_bin` become candidates at concat position bin * cnt + rank -- the order in which
//      the reference concatenates its per-bin torch.topk results.
//   2. bitonic sort of 64-bit keys (orderable score | ~concat position | cell): the reference's
//      argsort(descending) with ties by lower position; truncation to K is its global topk.
//   3. greedy NMS as a fixed point on the sorted list: a candidate is kept when every
//      higher-ranked neighbour within the radius is suppressed, suppressed when one is kept.
//      Decisions are final once taken, so the result equals the sequential greedy loop.
//   4. block-wide exclusive scan -> (y, x), score in descending-score order.
#include "extract_kernels.cuh"
#include "vc_launch.cuh"
#include "vc_ptx.cuh"

namespace vc {

namespace {

constexpr int kThreads = 1024;

__device__ __forceinline__ unsigned orderable(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float from_orderable(unsigned u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

struct Geometry {
  int nbh, nbw, bh, bw, cells, cnt, ncand, npad;
};

__host__ __device__ inline Geometry make_geometry(int H, int W, int K, int bin) {
  Geometry g;
  g.nbh = H / bin > 0 ? H / bin : 1;
  g.nbw = W / bin > 0 ? W / bin : 1;
  g.bh = H < bin ? H : bin;
  g.bw = W < bin ? W : bin;
  g.cells = g.bh * g.bw;
  int per_bin = K / (g.nbh * g.nbw);
  if (per_bin < 1) per_bin = 1;
  g.cnt = per_bin < g.cells ? per_bin : g.cells;
  g.ncand = g.nbh * g.nbw * g.cnt;
  g.npad = 32;
  while (g.npad < g.ncand) g.npad <<= 1;
  return g;
}

__host__ __device__ inline size_t region_a_bytes(int HW, int npad) {
  size_t a = (size_t)HW * 4;
  const size_t b = (size_t)HW * 2 + (size_t)npad;  // posmap (u16) + state (u8)
  if (b > a) a = b;
  return (a + 15) / 16 * 16;
}

__global__ void __launch_bounds__(kThreads, 1) select_kernel(const SelectParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  __shared__ int s_scan[32];
  __shared__ int s_total;
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  const int b = blockIdx.x;
  const int H = p.H, W = p.W, HW = H * W;
  const Geometry g = make_geometry(H, W, p.num_keypoints, p.bin_size);
  float* s_score = reinterpret_cast<float*>(smem);
  unsigned long long* keys = reinterpret_cast<unsigned long long*>(smem + region_a_bytes(HW, g.npad));
  const float* score = p.score + (size_t)b * HW;

  for (int i = threadIdx.x; i < HW; i += kThreads) s_score[i] = score[i];
  for (int i = threadIdx.x; i < g.npad; i += kThreads) keys[i] = 0ull;
  __syncthreads();

  // ---- 1. per-bin ranks
  const int nbinned = g.nbh * g.nbw * g.cells;
  for (int t = threadIdx.x; t < nbinned; t += kThreads) {
    const int bin = t / g.cells, local = t - bin * g.cells;
    const int bi = bin / g.nbw, bj = bin - bi * g.nbw;
    const int y0 = bi * p.bin_size, x0 = bj * p.bin_size;
    const int ly = local / g.bw, lx = local - ly * g.bw;
    const int cell = (y0 + ly) * W + x0 + lx;
    const float s = s_score[cell];
    int rank = 0;
    for (int yy = 0; yy < g.bh; ++yy) {
      const float* row = s_score + (y0 + yy) * W + x0;
      for (int xx = 0; xx < g.bw; ++xx) {
        const float o = row[xx];
        const int ol = yy * g.bw + xx;
        rank += (o > s || (o == s && ol < local)) ? 1 : 0;
      }
    }
    if (rank < g.cnt) {
      const int pos = bin * g.cnt + rank;
      keys[pos] = ((unsigned long long)orderable(s) << 32) |
                  ((unsigned long long)(0x3FFF - pos) << 14) | (unsigned long long)cell;
    }
  }
  __syncthreads();

  // ---- 2. bitonic sort, descending
  for (int k = 2; k <= g.npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < g.npad; i += kThreads) {
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
  const int n = g.ncand < p.num_keypoints ? g.ncand : p.num_keypoints;  // global topk :480-483

  // ---- 3. greedy NMS (region A is recycled: the scores now live in the keys)
  unsigned short* posmap = reinterpret_cast<unsigned short*>(smem);
  volatile uint8_t* state = smem + (size_t)HW * 2;
  for (int i = threadIdx.x; i < HW; i += kThreads) posmap[i] = 0xFFFF;
  for (int i = threadIdx.x; i < g.npad; i += kThreads) state[i] = 0;
  __syncthreads();
  for (int i = threadIdx.x; i < n; i += kThreads) posmap[(int)(keys[i] & 0x3FFF)] = (unsigned short)i;
  __syncthreads();
  const float r2 = p.nms_radius;
  int R = 0;
  while ((float)((R + 1) * (R + 1)) < r2 * r2 && R < 8) ++R;
  // distance test of the reference: 0 < sqrt(dy^2 + dx^2) < radius in fp32
  for (;;) {
    int pending = 0;
    for (int i = threadIdx.x; i < n; i += kThreads) {
      if (state[i] != 0) continue;
      const int cell = (int)(keys[i] & 0x3FFF);
      const int y = cell / W, x = cell - y * W;
      bool suppressed = false, wait = false;
      for (int dy = -R; dy <= R && !suppressed; ++dy) {
        const int yy = y + dy;
        if (yy < 0 || yy >= H) continue;
        for (int dx = -R; dx <= R; ++dx) {
          const int xx = x + dx;
          if ((dy | dx) == 0 || xx < 0 || xx >= W) continue;
          if (!(sqrtf((float)(dy * dy + dx * dx)) < p.nms_radius)) continue;
          const int q = posmap[yy * W + xx];
          if (q < i) {
            const uint8_t st = state[q];
            if (st == 1) {
              suppressed = true;
              break;
            }
            if (st == 0) wait = true;
          }
        }
      }
      if (suppressed)
        state[i] = 2;
      else if (!wait)
        state[i] = 1;
      else
        pending = 1;
    }
    if (!__syncthreads_or(pending)) break;
  }

  // ---- 4. ordered compaction
  const int per = (g.npad + kThreads - 1) / kThreads;
  const int i0 = threadIdx.x * per;
  int local = 0;
  for (int i = i0; i < i0 + per && i < n; ++i) local += state[i] == 1;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_scan[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int v = s_scan[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    s_scan[lane] = v;
    if (lane == 31) s_total = v;
  }
  __syncthreads();
  int out = incl - local + (warp > 0 ? s_scan[warp - 1] : 0);
  int32_t* oc = p.coords + (size_t)b * p.cap * 2;
  float* os = p.scores + (size_t)b * p.cap;
  for (int i = i0; i < i0 + per && i < n; ++i) {
    if (state[i] == 1) {
      if (out < p.cap) {
        const unsigned long long key = keys[i];
        const int cell = (int)(key & 0x3FFF);
        oc[2 * out] = cell / W;
        oc[2 * out + 1] = cell % W;
        os[out] = from_orderable((unsigned)(key >> 32));
      }
      ++out;
    }
  }
  if (threadIdx.x == 0) p.counts[b] = s_total < p.cap ? s_total : p.cap;
}

}  // namespace

int select_smem_bytes(int H, int W, int K, int bin_size) {
  const Geometry g = make_geometry(H, W, K, bin_size);
  return (int)(region_a_bytes(H * W, g.npad) + (size_t)g.npad * 8);
}

cudaError_t launch_select_keypoints(const SelectParams& p, cudaStream_t stream) {
  const int smem = select_smem_bytes(p.H, p.W, p.num_keypoints, p.bin_size);
  static SmemOptIn opt_in;
  if (cudaError_t e = ensure_dynamic_smem(select_kernel, smem, opt_in); e != cudaSuccess) return e;
  return launch_pdl(select_kernel, dim3(p.B), dim3(kThreads), smem, stream, p);
}

}  // namespace vc
