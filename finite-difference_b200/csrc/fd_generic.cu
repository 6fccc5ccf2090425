// Unfused operators: one kernel per reference array operation.  They execute
// any expression DAG in any number of dimensions and are the correctness
// backbone; the fused statement kernel (fd_fused.cu) is the fast path.
//
// Reference lines restated (src/finite_difference/FD.py):
//   der           990-1122      ewise   794-869
//   merge-store   541-576       Euler   515-516     NaN counts 528-529

#include <math.h>
#include <stdarg.h>

#include "fd_common.cuh"

namespace fdb {

thread_local char g_err[512] = "";
std::atomic<long long> g_launches{0};

int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return 1;
}

thread_local char g_kernel[192] = "";

void note_kernel(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_kernel, sizeof(g_kernel), fmt, ap);
  va_end(ap);
}

constexpr int kThreads = 256;

static inline unsigned blocks_for(int64_t n, int per_block) {
  int64_t b = (n + per_block - 1) / per_block;
  const int64_t cap = 148LL * 64;  // grid-stride beyond 64 waves of the 148 SMs
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (unsigned)b;
}

// ---------------------------------------------------------------- der ------
template <typename T>
__global__ void __launch_bounds__(kThreads)
der_kernel(T* __restrict__ dst, const T* __restrict__ src, int64_t outer, int64_t n_src,
           int64_t n_dst, int64_t inner, const __grid_constant__ DevStencil<T> st, int periodic,
           int64_t lo, int64_t hi, long long* flags) {
  const int64_t total = outer * n_dst * inner;
  bool saw_inf = false;
  for (int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; idx < total;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t i = idx % inner;
    const int64_t oj = idx / inner;
    const int64_t j = oj % n_dst;
    const int64_t o = oj / n_dst;
    const T* row = src + o * n_src * inner + i;
    T v;
    if (!periodic && (j < lo || j >= hi)) {
      v = quiet_nan<T>();
    } else {
      T acc = (T)0;
      for (int p = 0; p < st.npts; ++p) {
        int64_t jj = j + st.shifts[p];
        if (periodic) {
          jj %= n_src;
          if (jj < 0) jj += n_src;
        }
        acc = add_rn(acc, mul_rn(st.weights[p], row[jj * inner]));
      }
      v = apply_scale(acc, st.div_mode, st.scale, st.inv);
      saw_inf |= is_inf(v);
    }
    dst[idx] = v;
  }
  if (saw_inf) raise_flag(flags, FD_FLAG_INF);
}

template <typename T>
static int der_impl(void* dst, const void* src, int64_t outer, int64_t n_src, int64_t n_dst,
                    int64_t inner, const fd_stencil_t* st, int periodic, int64_t lo, int64_t hi,
                    void* flags, void* stream) {
  if (!dst || !src) return fail("fd_der: NULL data pointer");
  if (outer < 0 || n_src < 1 || n_dst < 0 || inner < 1) return fail("fd_der: bad extents");
  DevStencil<T> ds;
  if (make_dev_stencil<T>(st, &ds)) return 1;
  if (periodic && n_dst != n_src) return fail("fd_der: periodic axis must keep its length");
  if (!periodic) {
    if (lo < 0) lo = 0;
    if (hi > n_dst) hi = n_dst;
    for (int p = 0; p < ds.npts && lo < hi; ++p) {
      if (lo + ds.shifts[p] < 0 || hi - 1 + ds.shifts[p] >= n_src)
        return fail("fd_der: range [%lld,%lld) with shift %d leaves the source (n_src=%lld)",
                    (long long)lo, (long long)hi, ds.shifts[p], (long long)n_src);
    }
  }
  const int64_t total = outer * n_dst * inner;
  if (total == 0) return 0;
  der_kernel<T><<<blocks_for(total, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      (T*)dst, (const T*)src, outer, n_src, n_dst, inner, ds, periodic, lo, hi, (long long*)flags);
  return check_launch("fd_der");
}

// -------------------------------------------------------------- ewise ------
template <typename T>
__device__ __forceinline__ T pow_t(T a, T b);
template <>
__device__ __forceinline__ double pow_t<double>(double a, double b) { return pow(a, b); }
template <>
__device__ __forceinline__ float pow_t<float>(float a, float b) { return powf(a, b); }
template <typename T>
__device__ __forceinline__ T sqrt_t(T a);
template <>
__device__ __forceinline__ double sqrt_t<double>(double a) { return __dsqrt_rn(a); }
template <>
__device__ __forceinline__ float sqrt_t<float>(float a) { return __fsqrt_rn(a); }

template <typename T, int OP>
__global__ void __launch_bounds__(kThreads)
ewise_kernel(T* dst, const T* a, const T* b, T s, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x) {
    T x = a[i];
    T y = (OP >= FD_OP_ADD_FF && OP <= FD_OP_POW_FF) || OP == FD_OP_EULER ? b[i] : s;
    T r;
    switch (OP) {
      case FD_OP_NEG: r = -x; break;
      case FD_OP_ADD_FF: case FD_OP_ADD_FS: r = add_rn(x, y); break;
      case FD_OP_SUB_FF: case FD_OP_SUB_FS: r = sub_rn(x, y); break;
      case FD_OP_MUL_FF: case FD_OP_MUL_FS: r = mul_rn(x, y); break;
      case FD_OP_DIV_FF: case FD_OP_DIV_FS: r = div_rn(x, y); break;
      case FD_OP_POW_FF: case FD_OP_POW_FS: r = pow_t<T>(x, y); break;
      case FD_OP_SQUARE: r = mul_rn(x, x); break;
      case FD_OP_RECIP: r = div_rn((T)1, x); break;
      case FD_OP_SQRT: r = sqrt_t<T>(x); break;
      case FD_OP_COPY: r = x; break;
      case FD_OP_EULER: r = add_rn(x, mul_rn(y, s)); break;
      default: r = x;
    }
    dst[i] = r;
  }
}

template <typename T>
static int ewise_impl(int op, void* dst, const void* a, const void* b, double s, int64_t n,
                      void* stream) {
  if (n == 0) return 0;
  if (!dst || !a) return fail("fd_ewise: NULL data pointer");
  const bool needs_b = (op >= FD_OP_ADD_FF && op <= FD_OP_POW_FF) || op == FD_OP_EULER;
  if (needs_b && !b) return fail("fd_ewise: op %d needs a second field", op);
  const unsigned g = blocks_for(n, kThreads);
  cudaStream_t st = (cudaStream_t)stream;
#define FD_EW(OPC)                                                                           \
  case OPC:                                                                                  \
    ewise_kernel<T, OPC><<<g, kThreads, 0, st>>>((T*)dst, (const T*)a, (const T*)b, (T)s, n); \
    break;
  switch (op) {
    FD_EW(FD_OP_NEG) FD_EW(FD_OP_ADD_FF) FD_EW(FD_OP_SUB_FF) FD_EW(FD_OP_MUL_FF)
    FD_EW(FD_OP_DIV_FF) FD_EW(FD_OP_POW_FF) FD_EW(FD_OP_ADD_FS) FD_EW(FD_OP_SUB_FS)
    FD_EW(FD_OP_MUL_FS) FD_EW(FD_OP_DIV_FS) FD_EW(FD_OP_POW_FS) FD_EW(FD_OP_SQUARE)
    FD_EW(FD_OP_RECIP) FD_EW(FD_OP_SQRT) FD_EW(FD_OP_COPY) FD_EW(FD_OP_EULER)
    default: return fail("fd_ewise: unknown op %d", op);
  }
#undef FD_EW
  return check_launch("fd_ewise");
}

// -------------------------------------------------------- merge-store ------
template <typename T>
__global__ void __launch_bounds__(kThreads)
merge_kernel(T* __restrict__ stored, const T* __restrict__ fresh, int64_t n, long long* nan_count,
             long long* flags) {
  unsigned nans = 0;
  bool clash = false;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t rounds = (n + stride - 1) / stride;  // uniform trip count: warp ops below
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (int64_t r = 0; r < rounds; ++r, i += stride) {
    if (i < n) {
      const T f = fresh[i];
      if (is_nan(f)) {
        nans += is_nan(stored[i]) ? 1u : 0u;
      } else {
        clash |= !is_nan(stored[i]);
        stored[i] = f;
      }
    }
  }
  if (clash) raise_flag(flags, FD_FLAG_OVERRIDE);
  warp_add_count(nan_count, nans);
}

template <typename T>
static int merge_impl(void* stored, const void* fresh, int64_t n, void* nan_count, void* flags,
                      void* stream) {
  if (n == 0) return 0;
  if (!stored || !fresh) return fail("fd_merge_store: NULL data pointer");
  merge_kernel<T><<<blocks_for(n, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      (T*)stored, (const T*)fresh, n, (long long*)nan_count, (long long*)flags);
  return check_launch("fd_merge_store");
}

// ---------------------------------------------------- fill / face / count ---
template <typename T>
__global__ void __launch_bounds__(kThreads) fill_kernel(T* dst, T v, int64_t n) {
  for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < n;
       i += (int64_t)gridDim.x * blockDim.x)
    dst[i] = v;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
face_kernel(T* __restrict__ stored, const T* __restrict__ face, int64_t outer, int64_t n,
            int64_t inner, int64_t index) {
  const int64_t total = outer * inner;
  for (int64_t k = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; k < total;
       k += (int64_t)gridDim.x * blockDim.x) {
    const T v = face[k];
    if (!is_nan(v)) {
      const int64_t o = k / inner, i = k % inner;
      stored[(o * n + index) * inner + i] = v;
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
count_nan_kernel(const T* __restrict__ src, int64_t n, long long* nan_count) {
  unsigned nans = 0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t rounds = (n + stride - 1) / stride;
  int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  for (int64_t r = 0; r < rounds; ++r, i += stride)
    if (i < n) nans += is_nan(src[i]) ? 1u : 0u;
  warp_add_count(nan_count, nans);
}

}  // namespace fdb

using namespace fdb;

extern "C" {

int fd_abi_version(void) { return FD_ABI_VERSION; }
const char* fd_last_error(void) { return g_err; }
long long fd_launch_count(void) { return g_launches.load(); }
const char* fd_last_kernel(void) { return g_kernel; }

int fd_der_f64(void* dst, const void* src, int64_t outer, int64_t n_src, int64_t n_dst,
               int64_t inner, const fd_stencil_t* st, int periodic, int64_t lo, int64_t hi,
               void* flags, void* stream) {
  return der_impl<double>(dst, src, outer, n_src, n_dst, inner, st, periodic, lo, hi, flags, stream);
}
int fd_der_f32(void* dst, const void* src, int64_t outer, int64_t n_src, int64_t n_dst,
               int64_t inner, const fd_stencil_t* st, int periodic, int64_t lo, int64_t hi,
               void* flags, void* stream) {
  return der_impl<float>(dst, src, outer, n_src, n_dst, inner, st, periodic, lo, hi, flags, stream);
}

int fd_ewise_f64(int op, void* dst, const void* a, const void* b, double s, int64_t n, void* stream) {
  return ewise_impl<double>(op, dst, a, b, s, n, stream);
}
int fd_ewise_f32(int op, void* dst, const void* a, const void* b, double s, int64_t n, void* stream) {
  return ewise_impl<float>(op, dst, a, b, s, n, stream);
}

int fd_merge_store_f64(void* stored, const void* fresh, int64_t n, void* nan_count, void* flags, void* stream) {
  return merge_impl<double>(stored, fresh, n, nan_count, flags, stream);
}
int fd_merge_store_f32(void* stored, const void* fresh, int64_t n, void* nan_count, void* flags, void* stream) {
  return merge_impl<float>(stored, fresh, n, nan_count, flags, stream);
}

int fd_fill_f64(void* dst, double value, int64_t n, void* stream) {
  if (n == 0) return 0;
  if (!dst) return fail("fd_fill: NULL data pointer");
  fill_kernel<double><<<blocks_for(n, kThreads), kThreads, 0, (cudaStream_t)stream>>>((double*)dst, value, n);
  return check_launch("fd_fill");
}
int fd_fill_f32(void* dst, double value, int64_t n, void* stream) {
  if (n == 0) return 0;
  if (!dst) return fail("fd_fill: NULL data pointer");
  fill_kernel<float><<<blocks_for(n, kThreads), kThreads, 0, (cudaStream_t)stream>>>((float*)dst, (float)value, n);
  return check_launch("fd_fill");
}

int fd_face_store_f64(void* stored, const void* face, int64_t outer, int64_t n, int64_t inner, int64_t index, void* stream) {
  if (outer * inner == 0) return 0;
  if (!stored || !face) return fail("fd_face_store: NULL data pointer");
  if (index < 0 || index >= n) return fail("fd_face_store: index %lld outside [0,%lld)", (long long)index, (long long)n);
  face_kernel<double><<<blocks_for(outer * inner, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      (double*)stored, (const double*)face, outer, n, inner, index);
  return check_launch("fd_face_store");
}
int fd_face_store_f32(void* stored, const void* face, int64_t outer, int64_t n, int64_t inner, int64_t index, void* stream) {
  if (outer * inner == 0) return 0;
  if (!stored || !face) return fail("fd_face_store: NULL data pointer");
  if (index < 0 || index >= n) return fail("fd_face_store: index %lld outside [0,%lld)", (long long)index, (long long)n);
  face_kernel<float><<<blocks_for(outer * inner, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      (float*)stored, (const float*)face, outer, n, inner, index);
  return check_launch("fd_face_store");
}

int fd_count_nan_f64(const void* src, int64_t n, void* nan_count, void* stream) {
  if (n == 0) return 0;
  if (!src || !nan_count) return fail("fd_count_nan: NULL pointer");
  count_nan_kernel<double><<<blocks_for(n, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      (const double*)src, n, (long long*)nan_count);
  return check_launch("fd_count_nan");
}
int fd_count_nan_f32(const void* src, int64_t n, void* nan_count, void* stream) {
  if (n == 0) return 0;
  if (!src || !nan_count) return fail("fd_count_nan: NULL pointer");
  count_nan_kernel<float><<<blocks_for(n, kThreads), kThreads, 0, (cudaStream_t)stream>>>(
      (const float*)src, n, (long long*)nan_count);
  return check_launch("fd_count_nan");
}

}  // extern "C"
