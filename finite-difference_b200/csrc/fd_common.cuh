// Shared helpers for libfd_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>

#include "fd_b200.h"

namespace fdb {

extern thread_local char g_err[512];
extern std::atomic<long long> g_launches;

int fail(const char* fmt, ...);

// fd_last_kernel(): which kernel a fused entry point chose, and its geometry
extern thread_local char g_kernel[192];
void note_kernel(const char* fmt, ...);

inline int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail("%s: %s", what, cudaGetErrorString(e));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

// ---- exact (numpy-order) arithmetic: every product and sum is rounded on its
// own; the library is also compiled with -fmad=false.
__device__ __forceinline__ double mul_rn(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ float mul_rn(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ double add_rn(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ float add_rn(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ double sub_rn(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ float sub_rn(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ double div_rn(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ float div_rn(float a, float b) { return __fdiv_rn(a, b); }

template <typename T>
__device__ __forceinline__ T quiet_nan();
template <>
__device__ __forceinline__ double quiet_nan<double>() {
  return __longlong_as_double(0x7ff8000000000000LL);
}
template <>
__device__ __forceinline__ float quiet_nan<float>() {
  return __int_as_float(0x7fc00000);
}

template <typename T>
__device__ __forceinline__ bool is_nan(T v) {
  return v != v;
}
__device__ __forceinline__ bool is_inf(double v) {
  return fabs(v) == __longlong_as_double(0x7ff0000000000000LL);
}
__device__ __forceinline__ bool is_inf(float v) { return fabsf(v) == __int_as_float(0x7f800000); }

// Device-side stencil: weights in the compute type, divisor pre-analysed.
//   div_mode 0: scale == 1 (x/1 == x exactly, skipped)
//   div_mode 1: multiply by inv (exact when scale is a power of two, or the
//               caller asked for reciprocal scaling)
//   div_mode 2: IEEE division
//   div_mode 3: division by the invariant `scale` through its correctly rounded
//               reciprocal and one FMA correction step (Markstein): q = x*inv,
//               r = fma(-q, scale, x) (exact), result = fma(r, inv, q).  This is
//               the correctly rounded quotient whenever inv = RN(1/scale) and the
//               significand of scale is not all ones (checked on the host) and
//               the quotient is a normal number (checked per value; the rare
//               huge / tiny / non-finite quotients take the IEEE division).
//               3 FP instructions instead of ~30: keeps bit-exactness for the
//               everyday dx = 0.1 without making the kernels compute-bound.
template <typename T>
struct DevStencil {
  int32_t npts;
  int32_t div_mode;
  int32_t shifts[FD_MAX_POINTS];
  T weights[FD_MAX_POINTS];
  T scale;
  T inv;
};

template <typename T>
inline int make_dev_stencil(const fd_stencil_t* st, DevStencil<T>* out) {
  if (st == nullptr) return fail("stencil is NULL");
  if (st->npts < 1 || st->npts > FD_MAX_POINTS)
    return fail("stencil has %d points, supported range is 1..%d", st->npts, FD_MAX_POINTS);
  out->npts = st->npts;
  for (int i = 0; i < FD_MAX_POINTS; ++i) {
    out->shifts[i] = i < st->npts ? st->shifts[i] : 0;
    out->weights[i] = i < st->npts ? (T)st->weights[i] : (T)0;
  }
  T scale = (T)st->scale;
  out->scale = scale;
  out->inv = (T)1 / scale;
  int exp2 = 0;
  double mant = frexp((double)scale, &exp2);
  bool pow2 = (mant == 0.5 || mant == -0.5);
  // all-ones significand: the one case the correction step does not cover
  bool ones = false;
  if (sizeof(T) == 8) {
    uint64_t bits;
    double sd = (double)scale;
    memcpy(&bits, &sd, 8);
    ones = (bits & 0xFFFFFFFFFFFFFull) == 0xFFFFFFFFFFFFFull;
  } else {
    uint32_t bits;
    float sf = (float)scale;
    memcpy(&bits, &sf, 4);
    ones = (bits & 0x7FFFFFu) == 0x7FFFFFu;
  }
  const double lim = sizeof(T) == 8 ? 1e30 : 1e10;   // keeps q*scale and the residual inside the normal range
  const bool usable = scale == scale && fabs((double)scale) > 1.0 / lim && fabs((double)scale) < lim;
  if (scale == (T)1)
    out->div_mode = 0;
  else if (pow2 || !st->exact_div)
    out->div_mode = 1;
  else if (st->exact_div == 2 || ones || !usable)
    out->div_mode = 2;
  else
    out->div_mode = 3;
  return 0;
}

__device__ __forceinline__ double fma_rn(double a, double b, double c) { return __fma_rn(a, b, c); }
__device__ __forceinline__ float fma_rn(float a, float b, float c) { return __fmaf_rn(a, b, c); }

// Quotient-magnitude window inside which the FMA-corrected reciprocal division
// is exact (normal quotient, exact residual): biased exponent well away from
// both ends.  Tested on the integer pipe -- the FP64 pipe is the scarce one.
__device__ __forceinline__ bool quotient_is_safe(double q) {
  const unsigned e = ((unsigned)__double2hiint(q) >> 20) & 0x7ffu;
  return e - 200u < 1647u;                 // 2^-823 <= |q| < 2^824
}
__device__ __forceinline__ bool quotient_is_safe(float q) {
  const unsigned e = ((unsigned)__float_as_int(q) >> 23) & 0xffu;
  return e - 44u < 167u;                   // 2^-83 <= |q| < 2^84
}

// x / scale, correctly rounded, for a loop-invariant scale (div_mode 3 above).
template <typename T>
__device__ __forceinline__ T div_invariant(T x, T scale, T inv) {
  const T q = mul_rn(x, inv);
  if (quotient_is_safe(q)) {
    const T r = fma_rn(-q, scale, x);
    return fma_rn(r, inv, q);
  }
  if (x == (T)0) return q;                 // +-0 with the quotient's sign
  return div_rn(x, scale);                 // tiny, huge, inf, NaN: rare
}

template <typename T>
__device__ __forceinline__ T apply_scale(T v, int div_mode, T scale, T inv) {
  if (div_mode == 0) return v;
  if (div_mode == 1) return mul_rn(v, inv);
  if (div_mode == 3) return div_invariant(v, scale, inv);
  return div_rn(v, scale);
}

__device__ __forceinline__ void raise_flag(long long* flags, int which) {
  if (flags) atomicOr(reinterpret_cast<unsigned long long*>(flags) + which, 1ULL);
}

// Adds the per-thread count over the warp and issues at most one atomic.
__device__ __forceinline__ void warp_add_count(long long* counter, unsigned local) {
  unsigned total = __reduce_add_sync(0xffffffffu, local);
  if (counter && total && (threadIdx.x + threadIdx.y * blockDim.x) % 32 == 0)
    atomicAdd(reinterpret_cast<unsigned long long*>(counter), (unsigned long long)total);
}

}  // namespace fdb
