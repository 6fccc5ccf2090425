// Fused statement with a general point-wise expression (nonlinear terms such as
// u*du/dx, u**2, reaction terms): the expression is a postfix program over der
// terms, stored field levels and immediates, evaluated per point on a value
// stack that lives in shared memory ([slot][thread], conflict-free), followed by
// the usual assign_update / time_integrate_update chain.  One pass over HBM
// instead of one pass per reference array operation.
//
// Every operation is the IEEE operation the unfused ewise kernels perform
// (FD.py:794-869), in the same order, so the result is bit-identical to the
// operator-by-operator evaluation.

#include <math.h>

#include "fd_fused.cuh"

namespace fdb {

constexpr int kExprThreads = 128;

template <typename T>
struct DevExpr {
  DevFused<T> F;
  int32_t nfields, ncode;
  const T* field[FD_EXPR_MAX_FIELDS];
  int32_t op[FD_EXPR_MAX_CODE];
  int32_t arg[FD_EXPR_MAX_CODE];
  T imm[FD_EXPR_MAX_CODE];
};

template <typename T>
__device__ __forceinline__ T pow_dev(T a, T b);
template <>
__device__ __forceinline__ double pow_dev<double>(double a, double b) { return pow(a, b); }
template <>
__device__ __forceinline__ float pow_dev<float>(float a, float b) { return powf(a, b); }
template <typename T>
__device__ __forceinline__ T sqrt_dev(T a);
template <>
__device__ __forceinline__ double sqrt_dev<double>(double a) { return __dsqrt_rn(a); }
template <>
__device__ __forceinline__ float sqrt_dev<float>(float a) { return __fsqrt_rn(a); }

// One der term at one point (same code path as the scalar fused kernel).
template <typename T>
__device__ __forceinline__ T der_at(const DevTerm<T>& tm, int i0, int i1, int i2, bool& saw_inf) {
  const int j = term_index(tm, i0, i1, i2);
  if (!tm.periodic && (j < tm.lo || j >= tm.hi)) return quiet_nan<T>();
  const int64_t sa = tm.sa;
  const int64_t base = (int64_t)i0 * tm.stride[0] + (int64_t)i1 * tm.stride[1] + i2 - (int64_t)j * sa;
  T acc = (T)0;
  for (int p = 0; p < tm.st.npts; ++p) {
    int jj = j + tm.st.shifts[p];
    if (tm.periodic) {
      jj += (jj < 0) ? tm.n_src : 0;
      jj -= (jj >= tm.n_src) ? tm.n_src : 0;
    }
    acc = add_rn(acc, mul_rn(tm.st.weights[p], __ldg(tm.src + base + jj * sa)));
  }
  const T d = apply_scale(acc, tm.st.div_mode, tm.st.scale, tm.st.inv);
  saw_inf |= is_inf(d);
  return d;
}

// One der term at the PT points of a thread (i2_0 + p*step): the stencil point is
// the outer loop, so its shift and weight are fetched once for all PT points.
template <typename T, int PT>
__device__ __forceinline__ void ders_at(const DevTerm<T>& tm, int i0, int i1, int i2_0, int step, const bool (&active)[PT],
                                        T (&out)[PT], bool& saw_inf) {
  T acc[PT];
  bool bad[PT];
#pragma unroll
  for (int p = 0; p < PT; ++p) {
    acc[p] = (T)0;
    bad[p] = false;
  }
  if (tm.axis != 2) {          // (warp-uniform) the same index along the axis for all PT points
    const int j = term_index(tm, i0, i1, i2_0);
    const bool outside = !tm.periodic && (j < tm.lo || j >= tm.hi);
    if (!outside) {
      const int64_t sa = tm.sa;
      const T* base = tm.src + (int64_t)i0 * tm.stride[0] + (int64_t)i1 * tm.stride[1] + i2_0 - (int64_t)j * sa;
      for (int q = 0; q < tm.st.npts; ++q) {
        int jj = j + tm.st.shifts[q];
        if (tm.periodic) {
          jj += (jj < 0) ? tm.n_src : 0;
          jj -= (jj >= tm.n_src) ? tm.n_src : 0;
        }
        const T w = tm.st.weights[q];
        const T* row = base + (int64_t)jj * sa;
#pragma unroll
        for (int p = 0; p < PT; ++p)
          if (active[p]) acc[p] = add_rn(acc[p], mul_rn(w, __ldg(row + p * step)));
      }
    }
#pragma unroll
    for (int p = 0; p < PT; ++p) bad[p] = outside;
  } else {
    const T* row = tm.src + (int64_t)i0 * tm.stride[0] + (int64_t)i1 * tm.stride[1];
#pragma unroll
    for (int p = 0; p < PT; ++p) {
      const int j = i2_0 + p * step;
      bad[p] = !tm.periodic && (j < tm.lo || j >= tm.hi);
    }
    for (int q = 0; q < tm.st.npts; ++q) {
      const int sh = tm.st.shifts[q];
      const T w = tm.st.weights[q];
#pragma unroll
      for (int p = 0; p < PT; ++p) {
        int jj = i2_0 + p * step + sh;
        if (tm.periodic) {
          jj += (jj < 0) ? tm.n_src : 0;
          jj -= (jj >= tm.n_src) ? tm.n_src : 0;
        }
        if (active[p] && !bad[p]) acc[p] = add_rn(acc[p], mul_rn(w, __ldg(row + jj)));
      }
    }
  }
#pragma unroll
  for (int p = 0; p < PT; ++p) {
    const T d = apply_scale(acc[p], tm.st.div_mode, tm.st.scale, tm.st.inv);
    saw_inf |= active[p] && !bad[p] && is_inf(d);
    out[p] = bad[p] ? quiet_nan<T>() : d;
  }
}

// PT points per thread along the contiguous axis (consecutive points of a thread
// lie blockDim.x apart, so every access of a warp stays one contiguous run): the
// program is decoded once per PT points instead of once per point, and the top of
// the stack lives in registers, so a binary operation reads one operand from
// shared memory instead of reading two and writing one.  (One point per thread
// and everything on the shared-memory stack: Burgers 2-D at 8192^2 took 2.55 ms,
// ~1000 instructions per point, issue-bound; profiles/r02_expr_burgers_before_*.)
template <typename T, int PT>
__global__ void __launch_bounds__(kExprThreads)
expr_kernel(const __grid_constant__ DevExpr<T> P) {
  __shared__ T stack[FD_EXPR_MAX_STACK - 1][PT][kExprThreads];
  const DevFused<T>& F = P.F;
  const int n1 = F.n[1], n2 = F.n[2];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const int i2_0 = F.box_lo[2] + blockIdx.x * (blockDim.x * PT) + threadIdx.x;
  const int i1 = F.box_lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
  const bool row_active = i1 < F.box_hi[1];
  bool active[PT];
#pragma unroll
  for (int p = 0; p < PT; ++p) active[p] = row_active && i2_0 + p * (int)blockDim.x < F.box_hi[2];
  unsigned nan_cnt[2] = {0u, 0u};
  bool saw_inf = false, clash = false;

  for (int i0 = F.box_lo[0] + blockIdx.z * F.x_chunk, i0_end = min(i0 + F.x_chunk, F.box_hi[0]); i0 < i0_end; ++i0) {
    if (!active[0]) continue;
    const int64_t idx0 = ((int64_t)i0 * n1 + i1) * n2 + i2_0;
    // every der term of the statement, once per point (the program may push a term
    // several times; terms are few, so they wait in registers)
    T dval[FD_MAX_TERMS][PT];
#pragma unroll
    for (int k = 0; k < FD_MAX_TERMS; ++k) {
#pragma unroll
      for (int p = 0; p < PT; ++p) dval[k][p] = (T)0;
      if (k < F.nterms) ders_at<T, PT>(F.term[k], i0, i1, i2_0, (int)blockDim.x, active, dval[k], saw_inf);
    }
    T top[PT];
#pragma unroll
    for (int p = 0; p < PT; ++p) top[p] = (T)0;
    int sp = 0;   // warp-uniform: every thread runs the same program
    // spill the register top before something new is pushed
#define FD_X_SPILL()                                        \
    if (sp > 0) {                                           \
      _Pragma("unroll") for (int p = 0; p < PT; ++p) stack[sp - 1][p][tid] = top[p]; \
    }
#define FD_X_EACH(EXPR)                                     \
    _Pragma("unroll") for (int p = 0; p < PT; ++p) {        \
      const T a = top[p];                                   \
      top[p] = (EXPR);                                      \
    }
#define FD_X_PAIR(EXPR)                                     \
    --sp;                                                   \
    _Pragma("unroll") for (int p = 0; p < PT; ++p) {        \
      const T a = stack[sp - 1][p][tid], b = top[p];        \
      top[p] = (EXPR);                                      \
    }
    for (int pc = 0; pc < P.ncode; ++pc) {
      const int op = P.op[pc];
      const T imm = P.imm[pc];
      switch (op) {                 // decoded once for the PT points of the thread
        case FD_X_PUSH_DER: {
          FD_X_SPILL()
          ++sp;
          const int k = P.arg[pc];
#pragma unroll
          for (int p = 0; p < PT; ++p) top[p] = k == 0 ? dval[0][p] : (k == 1 ? dval[1][p] : (k == 2 ? dval[2][p] : dval[3][p]));
          break;
        }
        case FD_X_PUSH_FIELD: {
          FD_X_SPILL()
          ++sp;
          const T* f = P.field[P.arg[pc]] + idx0;
#pragma unroll
          for (int p = 0; p < PT; ++p)
            if (active[p]) top[p] = __ldg(f + p * (int)blockDim.x);
          break;
        }
        case FD_X_PUSH_IMM: {
          FD_X_SPILL()
          ++sp;
#pragma unroll
          for (int p = 0; p < PT; ++p) top[p] = imm;
          break;
        }
        case FD_X_NEG: FD_X_EACH(-a) break;
        case FD_X_ADD: FD_X_PAIR(add_rn(a, b)) break;
        case FD_X_SUB: FD_X_PAIR(sub_rn(a, b)) break;
        case FD_X_MUL: FD_X_PAIR(mul_rn(a, b)) break;
        case FD_X_DIV: FD_X_PAIR(div_rn(a, b)) break;
        case FD_X_POW: FD_X_PAIR(pow_dev<T>(a, b)) break;
        case FD_X_ADD_IMM: FD_X_EACH(add_rn(a, imm)) break;
        case FD_X_SUB_IMM: FD_X_EACH(sub_rn(a, imm)) break;
        case FD_X_MUL_IMM: FD_X_EACH(mul_rn(a, imm)) break;
        case FD_X_DIV_IMM: FD_X_EACH(div_rn(a, imm)) break;
        case FD_X_POW_IMM: FD_X_EACH(pow_dev<T>(a, imm)) break;
        case FD_X_SQUARE: FD_X_EACH(mul_rn(a, a)) break;
        case FD_X_RECIP: FD_X_EACH(div_rn((T)1, a)) break;
        default: FD_X_EACH(sqrt_dev<T>(a)) break;
      }
    }
#undef FD_X_EACH
#undef FD_X_PAIR
#undef FD_X_SPILL
#pragma unroll
    for (int p = 0; p < PT; ++p)
      if (active[p])
        fused_epilogue(F, top[p], idx0 + p * (int)blockDim.x, i0, i1, i2_0 + p * (int)blockDim.x, nan_cnt, clash);
  }
  if (saw_inf) raise_flag(F.flags, FD_FLAG_INF);
  if (clash) raise_flag(F.flags, FD_FLAG_OVERRIDE);
  warp_add_count(F.nan_count[0], nan_cnt[0]);
  if (F.depth > 1) warp_add_count(F.nan_count[1], nan_cnt[1]);
}

template <typename T>
static int expr_impl(const fd_expr_t* d, void* stream) {
  if (!d) return fail("fd_fused_expr: descriptor is NULL");
  DevExpr<T> P;
  if (d->base.npost != 0) return fail("fd_fused_expr: base.npost must be 0");
  if (make_dev_fused<T>(&d->base, &P.F, 0)) return 1;
  if (d->nfields < 0 || d->nfields > FD_EXPR_MAX_FIELDS) return fail("fd_fused_expr: %d fields, supported 0..%d", d->nfields, FD_EXPR_MAX_FIELDS);
  if (d->ncode < 1 || d->ncode > FD_EXPR_MAX_CODE) return fail("fd_fused_expr: %d instructions, supported 1..%d", d->ncode, FD_EXPR_MAX_CODE);
  P.nfields = d->nfields;
  P.ncode = d->ncode;
  for (int k = 0; k < FD_EXPR_MAX_FIELDS; ++k) {
    P.field[k] = k < d->nfields ? (const T*)d->field[k] : nullptr;
    if (k < d->nfields && !d->field[k]) return fail("fd_fused_expr: field %d is NULL", k);
  }
  // verify the program: operand indices and stack discipline
  int sp = 0;
  for (int pc = 0; pc < FD_EXPR_MAX_CODE; ++pc) {
    P.op[pc] = pc < d->ncode ? d->op[pc] : FD_X_NEG;
    P.arg[pc] = pc < d->ncode ? d->arg[pc] : 0;
    P.imm[pc] = pc < d->ncode ? (T)d->imm[pc] : (T)0;
    if (pc >= d->ncode) continue;
    const int op = d->op[pc];
    if (op == FD_X_PUSH_DER) {
      if (d->arg[pc] < 0 || d->arg[pc] >= d->base.nterms) return fail("fd_fused_expr: der index %d out of range", d->arg[pc]);
      ++sp;
    } else if (op == FD_X_PUSH_FIELD) {
      if (d->arg[pc] < 0 || d->arg[pc] >= d->nfields) return fail("fd_fused_expr: field index %d out of range", d->arg[pc]);
      ++sp;
    } else if (op == FD_X_PUSH_IMM) {
      ++sp;
    } else if (op >= FD_X_ADD && op <= FD_X_POW) {
      if (sp < 2) return fail("fd_fused_expr: stack underflow at instruction %d", pc);
      --sp;
    } else if (op == FD_X_NEG || (op >= FD_X_ADD_IMM && op <= FD_X_SQRT)) {
      if (sp < 1) return fail("fd_fused_expr: stack underflow at instruction %d", pc);
    } else {
      return fail("fd_fused_expr: unknown operation %d", op);
    }
    if (sp > FD_EXPR_MAX_STACK) return fail("fd_fused_expr: expression needs more than %d stack slots", FD_EXPR_MAX_STACK);
  }
  if (sp != 1) return fail("fd_fused_expr: the program leaves %d values on the stack", sp);

  const int n1 = P.F.box_hi[1] - P.F.box_lo[1], n2 = P.F.box_hi[2] - P.F.box_lo[2], nx = P.F.box_hi[0] - P.F.box_lo[0];
  if (nx <= 0 || n1 <= 0 || n2 <= 0) return 0;
  int bx = 32;
  while (bx < 128 && bx < n2) bx <<= 1;
  int by = kExprThreads / bx;
  while (by > 1 && by / 2 >= n1) by >>= 1;
  // points per thread along the row: 4 on long rows, fewer while blocks would idle
  int pt = 4;
  while (pt > 1 && bx * pt > n2) pt >>= 1;
  const int64_t gy = (n1 + by - 1) / by;
  if (gy > 65535) return fail("fd_fused_expr: middle dimension %d too large", n1);
  const int gx = (n2 + bx * pt - 1) / (bx * pt);
  int chunk = 8;
  while (chunk > 1 && (int64_t)gx * gy * ((nx + chunk - 1) / chunk) < 148 * 8) chunk >>= 1;
  while ((nx + chunk - 1) / chunk > 65535) chunk <<= 1;
  P.F.x_chunk = chunk;
  dim3 grid(gx, (unsigned)gy, (nx + chunk - 1) / chunk), block(bx, by, 1);
  note_kernel("expr f%d code=%d pt=%d chunk=%d grid=%ux%ux%u block=%dx%d", (int)sizeof(T) * 8, P.ncode, pt, chunk, grid.x, grid.y,
              grid.z, bx, by);
  if (pt == 4)
    expr_kernel<T, 4><<<grid, block, 0, (cudaStream_t)stream>>>(P);
  else if (pt == 2)
    expr_kernel<T, 2><<<grid, block, 0, (cudaStream_t)stream>>>(P);
  else
    expr_kernel<T, 1><<<grid, block, 0, (cudaStream_t)stream>>>(P);
  return check_launch("fd_fused_expr");
}

}  // namespace fdb

extern "C" {
int fd_fused_expr_f64(const fd_expr_t* desc, void* stream) { return fdb::expr_impl<double>(desc, stream); }
int fd_fused_expr_f32(const fd_expr_t* desc, void* stream) { return fdb::expr_impl<float>(desc, stream); }
}
