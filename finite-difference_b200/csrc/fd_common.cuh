// Shared helpers for libfd_b200.so (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <atomic>

#include "fd_b200.h"

namespace fdb {

extern thread_local char g_err[512];
extern std::atomic<long long> g_launches;

int fail(const char* fmt, ...);

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
  if (scale == (T)1)
    out->div_mode = 0;
  else if (pow2 || !st->exact_div)
    out->div_mode = 1;
  else
    out->div_mode = 2;
  return 0;
}

template <typename T>
__device__ __forceinline__ T apply_scale(T v, int div_mode, T scale, T inv) {
  if (div_mode == 0) return v;
  if (div_mode == 1) return mul_rn(v, inv);
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
