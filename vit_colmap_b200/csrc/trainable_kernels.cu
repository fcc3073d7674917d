// Keypoint selection of the `trainable_vit` extractor (reference:
// vit_colmap/features/trainable_vit_extractor.py _simple_nms :114-138, _run_inference :173-254):
// sigmoid score, (2r+1)^2 max-pool NMS with ties kept, threshold, top-K by score, sub-pixel
// offsets -> 6-column keypoints.  Descriptors are then gathered by sample_pack (quant mode 2).
//
// One CTA per image, any grid size:
//   0. sigmoid of the logits -> score plane (scratch, stays in L2)
//   A. each cell above the threshold compares itself with its window; survivors append a 64-bit key
//      (score bits << 32 | ~cell) to the candidate list (warp-aggregated shared-memory counter)
//   B. bitonic sort of the keys, descending: score order, ties by the lower row-major cell index
//      (the order torch.nonzero feeds torch.topk); the list lives in global memory but is touched by
//      this CTA only, so it is L2-resident and block barriers suffice
//   C. the first min(K, n) entries become keypoint rows.
#include "extract_kernels.cuh"
#include "vc_launch.cuh"
#include "vc_ptx.cuh"

#include <math.h>

namespace vc {

namespace {

constexpr int kThreads = 1024;

__global__ void __launch_bounds__(kThreads, 1) trainable_select_kernel(const TrainableSelectParams p) {
  __shared__ int s_count;
  griddep_wait();  // programmatic dependent launch: inputs belong to earlier kernels of the stream
  const int b = blockIdx.x;
  const int n = p.H * p.W;
  const float* kp = p.kp_map + (size_t)b * n * 4;
  float* score = p.score_scratch + (size_t)b * n;
  unsigned long long* keys = p.key_scratch + (size_t)b * p.key_capacity;
  if (threadIdx.x == 0) s_count = 0;

  // ---- 0. sigmoid plane
  for (int i = threadIdx.x; i < n; i += kThreads) {
    const float x = kp[4 * (size_t)i];
    score[i] = __fdiv_rn(1.0f, __fadd_rn(1.0f, expf(-x)));
  }
  __syncthreads();

  // ---- A. NMS + threshold -> candidate keys
  const int R = p.nms_radius;
  const int lane = threadIdx.x & 31;
  for (int i0 = 0; i0 < n; i0 += kThreads) {
    const int i = i0 + threadIdx.x;
    bool keep = false;
    float s = 0.f;
    if (i < n) {
      s = score[i];
      if (s > p.score_threshold) {
        const int y = i / p.W, x = i - y * p.W;
        float mx = s;
        for (int dy = -R; dy <= R; ++dy) {
          const int yy = y + dy;
          if (yy < 0 || yy >= p.H) continue;  // -inf padding of max_pool2d
          for (int dx = -R; dx <= R; ++dx) {
            const int xx = x + dx;
            if (xx < 0 || xx >= p.W) continue;
            mx = fmaxf(mx, score[yy * p.W + xx]);
          }
        }
        keep = (s == mx);
      }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    int base = 0;
    if (lane == 0 && bal) base = atomicAdd(&s_count, __popc(bal));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (keep) {
      const int pos = base + __popc(bal & ((1u << lane) - 1u));
      keys[pos] = ((unsigned long long)__float_as_uint(s) << 32) | (unsigned long long)(0xFFFFFFFFu - (unsigned)i);
    }
  }
  __syncthreads();
  const int count = s_count;
  int npad = 2;
  while (npad < count) npad <<= 1;
  for (int i = count + threadIdx.x; i < npad; i += kThreads) keys[i] = 0ull;
  __syncthreads();

  // ---- B. bitonic sort, descending
  for (int k = 2; k <= npad; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int t = threadIdx.x; t < (npad >> 1); t += kThreads) {
        // t-th compare-exchange of this stage: i has bit j clear
        const int i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        const int ixj = i | j;
        const unsigned long long a = keys[i], c = keys[ixj];
        const bool desc = (i & k) == 0;
        if (desc ? (a < c) : (a > c)) {
          keys[i] = c;
          keys[ixj] = a;
        }
      }
      __syncthreads();
    }
  }

  // ---- C. emit
  const int m = min(min(count, p.num_keypoints), p.cap);
  int32_t* oc = p.coords + (size_t)b * p.cap * 2;
  float* ok = p.keypoints + (size_t)b * p.cap * 6;
  for (int r = threadIdx.x; r < m; r += kThreads) {
    const unsigned long long key = keys[r];
    const int cell = (int)(0xFFFFFFFFu - (unsigned)(key & 0xFFFFFFFFull));
    const int y = cell / p.W, x = cell - y * p.W;
    const float4 v = *reinterpret_cast<const float4*>(kp + 4 * (size_t)cell);
    float kx = __fmul_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)x, v.y), 0.5f), 4.0f), p.scale_x);
    float ky = __fmul_rn(__fmul_rn(__fadd_rn(__fadd_rn((float)y, v.z), 0.5f), 4.0f), p.scale_y);
    kx = fminf(fmaxf(kx, 0.f), p.max_x);
    ky = fminf(fmaxf(ky, 0.f), p.max_y);
    const float ori = p.apply_tanh ? __fmul_rn(tanhf(v.w), 3.14159265358979323846f) : v.w;
    float* o = ok + 6 * (size_t)r;
    o[0] = kx;
    o[1] = ky;
    o[2] = 1.0f;
    o[3] = ori;
    o[4] = __uint_as_float((unsigned)(key >> 32));
    o[5] = 0.0f;
    oc[2 * r] = y;
    oc[2 * r + 1] = x;
  }
  if (threadIdx.x == 0) p.counts[b] = m;
}

}  // namespace

cudaError_t launch_trainable_select(const TrainableSelectParams& p, cudaStream_t stream) {
  if (p.B <= 0) return cudaSuccess;
  return launch_pdl(trainable_select_kernel, dim3(p.B), dim3(kThreads), 0, stream, p);
}

}  // namespace vc
