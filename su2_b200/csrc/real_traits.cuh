// Element-type specifics of the data-polling kernels (precond_poll_impl.inc, precond_factor_impl.inc): the sentinel NaN that
// marks "not produced yet", relaxed loads/stores at gpu scope and the asynchronous shared-memory copy of ONE element,
// for double (8 bytes) and float (4 bytes; the reference's su2mixedfloat).
#pragma once
#include <cuda_runtime.h>

namespace su2b200 {
template <class T> struct RealBits;
template <> struct RealBits<double> {
  static constexpr unsigned long long SENTINEL = 0xFFF8C0DEC0DE5A5AULL;  // quiet NaN with a private payload
  __device__ __forceinline__ static bool is_sentinel(double v) { return (unsigned long long)__double_as_longlong(v) == SENTINEL; }
  // every double with this high word is a NaN that arithmetic cannot produce: the high word identifies the sentinel
  __device__ __forceinline__ static bool is_sentinel_hi(double v) { return (unsigned)__double2hiint(v) == (unsigned)(SENTINEL >> 32); }
  __device__ __forceinline__ static double sentinel() { return __longlong_as_double((long long)SENTINEL); }
  __device__ __forceinline__ static double ld_relaxed(const double* p) {
    double v;
    asm volatile("ld.relaxed.gpu.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
  }
  __device__ __forceinline__ static void st_relaxed(double* p, double v) {
    asm volatile("st.relaxed.gpu.global.f64 [%0], %1;" ::"l"(p), "d"(v) : "memory");
  }
  __device__ __forceinline__ static void cp_async(void* smem, const double* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s), "l"(gmem) : "memory");
  }
};
template <> struct RealBits<float> {
  static constexpr unsigned SENTINEL = 0xFFC0DE5Au;  // quiet NaN with a private payload (arithmetic yields 0x7FFFFFFF)
  __device__ __forceinline__ static bool is_sentinel(float v) { return __float_as_uint(v) == SENTINEL; }
  __device__ __forceinline__ static bool is_sentinel_hi(float v) { return __float_as_uint(v) == SENTINEL; }
  __device__ __forceinline__ static float sentinel() { return __uint_as_float(SENTINEL); }
  __device__ __forceinline__ static float ld_relaxed(const float* p) {
    float v;
    asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
  }
  __device__ __forceinline__ static void st_relaxed(float* p, float v) {
    asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
  }
  __device__ __forceinline__ static void cp_async(void* smem, const float* gmem) {
    const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(s), "l"(gmem) : "memory");
  }
};
}  // namespace su2b200
