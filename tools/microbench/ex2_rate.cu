// Microbenchmark: MUFU ex2 throughput per SM for f32, f16x2 and bf16x2 operands (sm_100a).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ex2_rate ex2_rate.cu && ./ex2_rate
// Each thread runs ILP independent dependency chains of ex2 so that the MUFU pipe, not latency, limits.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(512) k(float* out, int iters) {
  constexpr int ILP = 8;
  uint32_t v[ILP];
  for (int i = 0; i < ILP; ++i) v[i] = (MODE == 0) ? __float_as_uint(-0.001f * (threadIdx.x + i + 1)) : 0x80018001u + threadIdx.x + i;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < ILP; ++i) {
      if (MODE == 0) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+r"(v[i]));
      if (MODE == 1) asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(v[i]));
      if (MODE == 2) asm volatile("ex2.approx.ftz.bf16x2 %0, %0;" : "+r"(v[i]));
      if (MODE == 3) {  // f32 FMA chain as a control for the FMA pipe
        float f = __uint_as_float(v[i]);
        asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(f));
        v[i] = __float_as_uint(f);
      }
    }
  }
  uint32_t s = 0;
  for (int i = 0; i < ILP; ++i) s ^= v[i];
  if (s == 0x12345678u) out[0] = 1.f;
}

template <int MODE>
void run(const char* name, int elems_per_op) {
  float* out;
  cudaMalloc(&out, 4);
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  const int iters = 20000, blocks = p.multiProcessorCount * 4, threads = 512;
  k<MODE><<<blocks, threads>>>(out, 10);
  cudaDeviceSynchronize();
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a);
  k<MODE><<<blocks, threads>>>(out, iters);
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms;
  cudaEventElapsedTime(&ms, a, b);
  int khz;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  const double ops = (double)blocks * threads * iters * 8;
  printf("%-10s %8.3f ms  %7.2f Tops/s (instr lanes)  %7.2f Telem/s  ~%.1f lane-ops/clk/SM at %d MHz nominal\n", name, ms,
         ops / ms / 1e9, ops * elems_per_op / ms / 1e9, ops / (ms * 1e-3) / p.multiProcessorCount / (khz * 1e3), khz / 1000);
  cudaFree(out);
}

int main() {
  run<0>("f32", 1);
  run<1>("f16x2", 2);
  run<2>("bf16x2", 2);
  run<3>("fma.f32", 1);
  return 0;
}
