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

constexpr int kExprThreads = 256;

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

template <typename T>
__global__ void __launch_bounds__(kExprThreads)
expr_kernel(const __grid_constant__ DevExpr<T> P) {
  __shared__ T stack[FD_EXPR_MAX_STACK][kExprThreads];
  const DevFused<T>& F = P.F;
  const int n1 = F.n[1], n2 = F.n[2];
  const int tid = threadIdx.y * blockDim.x + threadIdx.x;
  const int i2 = blockIdx.x * blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y * blockDim.y + threadIdx.y;
  const bool active = i2 >= F.box_lo[2] && i2 < F.box_hi[2] && i1 >= F.box_lo[1] && i1 < F.box_hi[1];
  unsigned nan_cnt[2] = {0u, 0u};
  bool saw_inf = false, clash = false;

  for (int i0 = F.box_lo[0] + blockIdx.z * F.x_chunk, i0_end = min(i0 + F.x_chunk, F.box_hi[0]); i0 < i0_end; ++i0) {
    if (!active) continue;
    const int64_t idx = ((int64_t)i0 * n1 + i1) * n2 + i2;
    int sp = 0;   // warp-uniform: every thread runs the same program
    for (int pc = 0; pc < P.ncode; ++pc) {
      const int op = P.op[pc];
      switch (op) {
        case FD_X_PUSH_DER: stack[sp++][tid] = der_at<T>(F.term[P.arg[pc]], i0, i1, i2, saw_inf); break;
        case FD_X_PUSH_FIELD: stack[sp++][tid] = __ldg(P.field[P.arg[pc]] + idx); break;
        case FD_X_PUSH_IMM: stack[sp++][tid] = P.imm[pc]; break;
        case FD_X_NEG: stack[sp - 1][tid] = -stack[sp - 1][tid]; break;
        case FD_X_ADD: --sp; stack[sp - 1][tid] = add_rn(stack[sp - 1][tid], stack[sp][tid]); break;
        case FD_X_SUB: --sp; stack[sp - 1][tid] = sub_rn(stack[sp - 1][tid], stack[sp][tid]); break;
        case FD_X_MUL: --sp; stack[sp - 1][tid] = mul_rn(stack[sp - 1][tid], stack[sp][tid]); break;
        case FD_X_DIV: --sp; stack[sp - 1][tid] = div_rn(stack[sp - 1][tid], stack[sp][tid]); break;
        case FD_X_POW: --sp; stack[sp - 1][tid] = pow_dev<T>(stack[sp - 1][tid], stack[sp][tid]); break;
        case FD_X_ADD_IMM: stack[sp - 1][tid] = add_rn(stack[sp - 1][tid], P.imm[pc]); break;
        case FD_X_SUB_IMM: stack[sp - 1][tid] = sub_rn(stack[sp - 1][tid], P.imm[pc]); break;
        case FD_X_MUL_IMM: stack[sp - 1][tid] = mul_rn(stack[sp - 1][tid], P.imm[pc]); break;
        case FD_X_DIV_IMM: stack[sp - 1][tid] = div_rn(stack[sp - 1][tid], P.imm[pc]); break;
        case FD_X_POW_IMM: stack[sp - 1][tid] = pow_dev<T>(stack[sp - 1][tid], P.imm[pc]); break;
        case FD_X_SQUARE: stack[sp - 1][tid] = mul_rn(stack[sp - 1][tid], stack[sp - 1][tid]); break;
        case FD_X_RECIP: stack[sp - 1][tid] = div_rn((T)1, stack[sp - 1][tid]); break;
        case FD_X_SQRT: stack[sp - 1][tid] = sqrt_dev<T>(stack[sp - 1][tid]); break;
      }
    }
    fused_epilogue(F, stack[0][tid], idx, i0, i1, i2, nan_cnt, clash);
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

  const int n1 = P.F.box_hi[1] - P.F.box_lo[1], n2 = P.F.n[2], nx = P.F.box_hi[0] - P.F.box_lo[0];
  if (nx <= 0 || n1 <= 0 || P.F.box_hi[2] <= P.F.box_lo[2]) return 0;
  int bx = 32;
  while (bx < 128 && bx < n2) bx <<= 1;
  int by = kExprThreads / bx;
  while (by > 1 && by / 2 >= P.F.n[1]) by >>= 1;
  const int64_t gy = (P.F.n[1] + by - 1) / by;
  if (gy > 65535) return fail("fd_fused_expr: middle dimension %d too large", P.F.n[1]);
  int chunk = 8;
  while (chunk > 1 && (int64_t)((n2 + bx - 1) / bx) * gy * ((nx + chunk - 1) / chunk) < 148 * 8) chunk >>= 1;
  while ((nx + chunk - 1) / chunk > 65535) chunk <<= 1;
  P.F.x_chunk = chunk;
  dim3 grid((n2 + bx - 1) / bx, (unsigned)gy, (nx + chunk - 1) / chunk), block(bx, by, 1);
  note_kernel("expr f%d code=%d chunk=%d grid=%ux%ux%u block=%dx%d", (int)sizeof(T) * 8, P.ncode, chunk, grid.x, grid.y, grid.z,
              bx, by);
  expr_kernel<T><<<grid, block, 0, (cudaStream_t)stream>>>(P);
  return check_launch("fd_fused_expr");
}

}  // namespace fdb

extern "C" {
int fd_fused_expr_f64(const fd_expr_t* desc, void* stream) { return fdb::expr_impl<double>(desc, stream); }
int fd_fused_expr_f32(const fd_expr_t* desc, void* stream) { return fdb::expr_impl<float>(desc, stream); }
}
