// Shared device helpers for the eegspec kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace eegspec {

constexpr int kWarp = 32;

// NaN-propagating min / max (torch.min / torch.max semantics; fminf/fmaxf drop NaNs).
__device__ __forceinline__ float min_nan(float a, float b) {
  float r;
  asm("min.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}
__device__ __forceinline__ float max_nan(float a, float b) {
  float r;
  asm("max.NaN.f32 %0, %1, %2;" : "=f"(r) : "f"(a), "f"(b));
  return r;
}

__device__ __forceinline__ float warp_min_nan(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = min_nan(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_max_nan(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = max_nan(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Streaming 128-bit store: the image is written once and never re-read by this library.
__device__ __forceinline__ void st_stream_f4(float* p, float4 v) {
  asm volatile("st.global.cs.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}

// log1p(m) for m >= 0 in ~11 instructions.  m < 1/16: alternating series to m^6 (relative error
// < 2e-7); otherwise ln2 * lg2.approx(1 + m), whose absolute error (< 1.2e-7 for 1+m < 2, relative
// 2^-22 above) is far below what survives the per-row min-max normalisation (tests: <= 1e-4).
__device__ __forceinline__ float log1p_pos(float m) {
  float p = fmaf(m, -1.0f / 6.0f, 0.2f);
  p = fmaf(m, p, -0.25f);
  p = fmaf(m, p, 1.0f / 3.0f);
  p = fmaf(m, p, -0.5f);
  p = fmaf(m, p, 1.0f);
  const float big = 0.69314718055994531f * __log2f(1.0f + m);
  return m < 0.0625f ? m * p : big;
}

// sqrt(p) for p >= 0 via rsqrt.approx (2 ulp); p == 0 and NaN handled (0 * inf would be NaN).
__device__ __forceinline__ float sqrt_pos(float p) {
  const float r = p * rsqrtf(p);
  return p > 0.f ? r : p;
}

// sqrt.approx (one MUFU, relative error 2^-23, sqrt(0) = 0, NaN propagates)
__device__ __forceinline__ float sqrt_mufu(float p) {
  float r;
  asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(p));
  return r;
}

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}

}  // namespace eegspec
