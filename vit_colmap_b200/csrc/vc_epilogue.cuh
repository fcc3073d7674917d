// Epilogue arithmetic shared by the tcgen05 GEMM kernels (vit_gemm.cu, conv_gemm.cu): packed fp32x2
// helpers, the exact-erf GELU and bf16 packing.
#pragma once
#include <cstdint>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace vc {

__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)),
        "l"(*reinterpret_cast<uint64_t*>(&c)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float2 fmul2(float2 a, float2 b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;"
      : "=l"(d)
      : "l"(*reinterpret_cast<uint64_t*>(&a)), "l"(*reinterpret_cast<uint64_t*>(&b)));
  return *reinterpret_cast<float2*>(&d);
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Exact-erf GELU, x * Phi(x), for two values at once on the packed fp32 pipe.
// Phi(-|x|) = erfc(|x|/sqrt2)/2 = 2^(P(|x|/sqrt2) - 1) with P a degree-6 weighted least-squares
// fit of log2(erfc(t)) on [0, 3.95] (|t| is clamped there: erfc(3.95) = 2e-8); then
// gelu(x) = x * e for x < 0 and x - x * e for x >= 0, so no cancellation on either side.
// Max abs error against the fp64 erf formula: 4.9e-7 over [-12, 12] (tools/fit_gelu.py), two orders
// below the bf16 rounding of the stored activation.
__device__ __forceinline__ float2 gelu_erf2(float2 x) {
  float2 z = fmul2(make_float2(fabsf(x.x), fabsf(x.y)), make_float2(0.70710678118654752440f, 0.70710678118654752440f));
  z.x = fminf(z.x, 3.95f);
  z.y = fminf(z.y, 3.95f);
  float2 q = make_float2(1.689791825e-04f, 1.689791825e-04f);
  q = ffma2(q, z, make_float2(-3.802316729e-03f, -3.802316729e-03f));
  q = ffma2(q, z, make_float2(3.115207143e-02f, 3.115207143e-02f));
  q = ffma2(q, z, make_float2(-1.499114484e-01f, -1.499114484e-01f));
  q = ffma2(q, z, make_float2(-9.180871844e-01f, -9.180871844e-01f));
  q = ffma2(q, z, make_float2(-1.627936125e+00f, -1.627936125e+00f));
  q = ffma2(q, z, make_float2(3.678871110e-07f - 1.0f, 3.678871110e-07f - 1.0f));
  const float2 t = fmul2(x, make_float2(ex2(q.x), ex2(q.y)));
  return make_float2(x.x >= 0.f ? x.x - t.x : t.x, x.y >= 0.f ? x.y - t.y : t.y);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

}  // namespace vc
