// A derivative of a computed expression -- d.der(kappa.always * d.der(T.prev, "x"), "x"),
// d.der(u.prev * u.prev, "x") -- in two launches instead of one per reference
// array operation: the inner expression is evaluated once into a scratch level
// (the array the reference's operators would have produced, FD.py:794-869 and
// 990-1122, NaN outside a fixed axis' range included), then the statement runs
// as usual with the scratch level as the source of its der terms.
//
// fd_fused_nested_*        the caller provides the scratch level;
// fd_fused_nested_static_* the scratch level is a fixed array inside the library.

#include <cuda_runtime.h>

#include "fd_common.cuh"

namespace fdb {

// Copies of the two descriptors with the scratch level filled in where the
// caller left NULL: the inner statement's destination, the outer statement's
// sources.  Returns 0 or an error (message in fd_last_error).
struct NestedPlan {
  fd_expr_t inner, outer;     // (an fd_fused_t occupies the `base` part)
  bool inner_is_expr, outer_is_expr;
};

static int plan_nested(int inner_is_expr, const void* inner, int outer_is_expr, const void* outer, void* scratch,
                       long long scratch_elems, NestedPlan* plan) {
  if (!inner || !outer) return fail("fd_fused_nested: descriptor is NULL");
  if (!scratch) return fail("fd_fused_nested: no scratch level");
  plan->inner_is_expr = inner_is_expr != 0;
  plan->outer_is_expr = outer_is_expr != 0;
  memset(&plan->inner, 0, sizeof plan->inner);
  memset(&plan->outer, 0, sizeof plan->outer);
  memcpy(&plan->inner, inner, inner_is_expr ? sizeof(fd_expr_t) : sizeof(fd_fused_t));
  memcpy(&plan->outer, outer, outer_is_expr ? sizeof(fd_expr_t) : sizeof(fd_fused_t));
  fd_fused_t& in = plan->inner.base;
  if (in.depth != 1 || in.prev[0] || in.out[0] || in.dot_out)
    return fail("fd_fused_nested: the inner statement must be a plain assignment (depth 1, prev[0], out[0] and dot_out NULL)");
  const long long elems = (long long)in.n[0] * in.n[1] * in.n[2];
  if (elems > scratch_elems)
    return fail("fd_fused_nested: the inner expression has %lld values, the scratch level holds %lld", elems, scratch_elems);
  in.out[0] = scratch;
  in.nan_count[0] = nullptr;
  int patched = 0;
  fd_fused_t& out = plan->outer.base;
  for (int k = 0; k < out.nterms && k < FD_MAX_TERMS; ++k)
    if (!out.term[k].src) {
      out.term[k].src = scratch;
      ++patched;
    }
  if (outer_is_expr)
    for (int k = 0; k < plan->outer.nfields && k < FD_EXPR_MAX_FIELDS; ++k)
      if (!plan->outer.field[k]) {
        plan->outer.field[k] = scratch;
        ++patched;
      }
  if (!patched) return fail("fd_fused_nested: the outer statement does not read the scratch level (no NULL source)");
  return 0;
}

}  // namespace fdb

// ---- scratch level inside the library (fd_fused_nested_static_f64) ---------------
//
// For hosts that cannot hand over device memory of their own at this point (see
// fd_b200.h).  The array is part of the loaded module: every GPU that loads the
// library holds these bytes until the process ends, whether the entry point is
// ever called or not -- hence the modest size.
static __device__ __align__(256) unsigned char g_static_scratch[FD_STATIC_SCRATCH_BYTES];

namespace fdb {

// In host code `g_static_scratch` is only a handle for the runtime, not an address
// a kernel could use: the device address is asked for once per device, with that
// device current, and kept.
static int static_scratch_address(void** out) {
  constexpr int kMaxDevices = 64;
  static void* address[kMaxDevices] = {};
  static std::atomic<int> ready[kMaxDevices] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices)
    return fail("fd_fused_nested_static: no current CUDA device (%s)", cudaGetErrorString(e));
  if (!ready[dev].load(std::memory_order_acquire)) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail("fd_fused_nested_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
    }
    address[dev] = p;          // (two threads racing here store the same value)
    ready[dev].store(1, std::memory_order_release);
  }
  *out = address[dev];
  return 0;
}

}  // namespace fdb

using namespace fdb;

#define FD_NESTED(SFX)                                                                                              \
  static int run_nested_##SFX(const NestedPlan& p, void* stream) {                                                  \
    int rc = p.inner_is_expr ? fd_fused_expr_##SFX(&p.inner, stream) : fd_fused_step_##SFX(&p.inner.base, stream);  \
    if (rc) return rc;                                                                                              \
    return p.outer_is_expr ? fd_fused_expr_##SFX(&p.outer, stream) : fd_fused_step_##SFX(&p.outer.base, stream);    \
  }
FD_NESTED(f64)
FD_NESTED(f32)
#undef FD_NESTED

extern "C" {

int fd_fused_nested_f64(int inner_is_expr, const void* inner, int outer_is_expr, const void* outer, void* scratch,
                        long long scratch_elems, void* stream) {
  NestedPlan p;
  if (plan_nested(inner_is_expr, inner, outer_is_expr, outer, scratch, scratch_elems, &p)) return 1;
  return run_nested_f64(p, stream);
}

int fd_fused_nested_f32(int inner_is_expr, const void* inner, int outer_is_expr, const void* outer, void* scratch,
                        long long scratch_elems, void* stream) {
  NestedPlan p;
  if (plan_nested(inner_is_expr, inner, outer_is_expr, outer, scratch, scratch_elems, &p)) return 1;
  return run_nested_f32(p, stream);
}

int fd_fused_nested_static_f64(int inner_is_expr, const void* inner, int outer_is_expr, const void* outer, void* stream) {
  // the size test comes first: an expression that does not fit is refused before
  // anything touches the device
  if (inner) {
    const fd_fused_t* in = inner_is_expr ? &((const fd_expr_t*)inner)->base : (const fd_fused_t*)inner;
    const long long elems = (long long)in->n[0] * in->n[1] * in->n[2];
    if (elems > (long long)(FD_STATIC_SCRATCH_BYTES / sizeof(double)))
      return fail("fd_fused_nested_static_f64: the inner expression has %lld values, the built-in scratch level holds %lld "
                  "(FD_STATIC_SCRATCH_BYTES = %d); use fd_fused_nested_f64 with a scratch level of your own",
                  elems, (long long)(FD_STATIC_SCRATCH_BYTES / sizeof(double)), (int)FD_STATIC_SCRATCH_BYTES);
  }
  void* scratch = nullptr;
  if (static_scratch_address(&scratch)) return 1;
  NestedPlan p;
  if (plan_nested(inner_is_expr, inner, outer_is_expr, outer, scratch, FD_STATIC_SCRATCH_BYTES / sizeof(double), &p)) return 1;
  return run_nested_f64(p, stream);
}

}  // extern "C"
