// Device-side descriptor of one fused statement and the pieces shared by the
// fused kernels (general and star-stencil): scalar post-operations, the Euler
// chain and the in-kernel boundary merge.
#pragma once

#include "fd_common.cuh"

namespace fdb {

// ---- 128-bit vector access along the contiguous axis --------------------------
template <typename T>
struct Vec;
template <>
struct Vec<double> {
  using type = double2;
  static constexpr int N = 2;
};
template <>
struct Vec<float> {
  using type = float4;
  static constexpr int N = 4;
};

template <typename T, int N>
struct Pack {
  T v[N];
};

template <typename T>
__device__ __forceinline__ Pack<T, Vec<T>::N> load_vec(const T* p) {
  using V = typename Vec<T>::type;
  union {
    V vec;
    Pack<T, Vec<T>::N> pk;
  } u;
  u.vec = __ldg(reinterpret_cast<const V*>(p));
  return u.pk;
}

template <typename T>
__device__ __forceinline__ void store_vec(T* p, const Pack<T, Vec<T>::N>& pk) {
  using V = typename Vec<T>::type;
  union {
    V vec;
    Pack<T, Vec<T>::N> pk;
  } u;
  u.pk = pk;
  *reinterpret_cast<V*>(p) = u.vec;
}

__device__ __forceinline__ int wrap_or_clamp(int i, int n, bool periodic) {
  if (periodic) {
    i += (i < 0) ? n : 0;
    i -= (i >= n) ? n : 0;
    return i;
  }
  return max(0, min(i, n - 1));
}

template <typename T>
struct DevTerm {
  const T* src;
  int32_t axis, periodic, lo, hi, n_src, subtract, has_pre, negate;
  int32_t fold_div, fold_mod;   // axis 0 is a merge of several leading axes (fd_term_t), fold_mod == 0: it is not
  T pre;
  int64_t stride[3];  // element strides of the source slice
  int64_t sa;         // element stride between neighbours along the differentiated axis
  DevStencil<T> st;
};

// index of the point along the term's differentiated axis
template <typename T>
__device__ __forceinline__ int term_index(const DevTerm<T>& tm, int i0, int i1, int i2) {
  if (tm.axis == 0) return tm.fold_mod ? (i0 / tm.fold_div) % tm.fold_mod : i0;
  return tm.axis == 1 ? i1 : i2;
}

template <typename T>
struct DevPost {
  int32_t op;
  T s;
  const T* coef;
};

template <typename T>
struct DevFused {
  int32_t n[3];
  int32_t nterms, npost, depth, x_chunk, any_face;
  DevTerm<T> term[FD_MAX_TERMS];
  DevPost<T> post[FD_MAX_POST];
  T dt;
  T* dot_out;
  const T* prev[2];
  T* out[2];
  const T* face[2][3][2];
  long long* nan_count[2];
  long long* flags;
  int32_t box_lo[3], box_hi[3];
  int32_t whole;   // the box is the whole slice
  int32_t rim_lo[3], rim_hi[3];
};

template <typename T>
inline int make_dev_fused(const fd_fused_t* d, DevFused<T>* P, int min_terms = 1) {
  if (!d) return fail("fd_fused_step: descriptor is NULL");
  memset((void*)&P->term[0], 0, sizeof(P->term[0]));
  if (d->nterms < min_terms || d->nterms > FD_MAX_TERMS)
    return fail("fd_fused_step: %d terms, supported 1..%d", d->nterms, FD_MAX_TERMS);
  if (d->npost < 0 || d->npost > FD_MAX_POST)
    return fail("fd_fused_step: %d post operations, supported 0..%d", d->npost, FD_MAX_POST);
  if (d->depth < 1 || d->depth > 2) return fail("fd_fused_step: depth %d, supported 1..2", d->depth);
  for (int a = 0; a < 3; ++a) {
    if (d->n[a] < 0) return fail("fd_fused_step: negative extent");
    P->n[a] = d->n[a];
  }
  P->nterms = d->nterms;
  P->npost = d->npost;
  P->depth = d->depth;
  P->x_chunk = 1;
  P->any_face = 0;
  for (int k = 0; k < d->nterms; ++k) {
    const fd_term_t& s = d->term[k];
    DevTerm<T>& t = P->term[k];
    if (!s.src) return fail("fd_fused_step: term %d has no source", k);
    if (s.axis < 0 || s.axis > 2) return fail("fd_fused_step: term %d axis %d", k, s.axis);
    if (make_dev_stencil<T>(&s.st, &t.st)) return 1;
    t.src = (const T*)s.src;
    t.axis = s.axis;
    t.periodic = s.periodic;
    t.n_src = s.n_src;
    t.subtract = s.subtract;
    t.has_pre = s.has_pre;
    t.negate = s.negate;
    t.pre = (T)s.pre;
    t.fold_div = t.fold_mod = 0;
    if (s.fold_mod != 0) {
      if (s.axis != 0 || s.fold_mod < 1 || s.fold_div < 1 || d->n[0] % ((int64_t)s.fold_div * s.fold_mod) != 0)
        return fail("fd_fused_step: term %d: fold (%d, %d) does not divide dimension 0 of %d points", k, s.fold_div,
                    s.fold_mod, d->n[0]);
      if (s.n_src != s.fold_mod) return fail("fd_fused_step: term %d: a folded axis must keep its extent", k);
      if (s.fold_div != 1 || s.fold_mod != d->n[0]) {
        t.fold_div = s.fold_div;
        t.fold_mod = s.fold_mod;
      }
    }
    const int n_dst = t.fold_mod ? t.fold_mod : d->n[s.axis];
    if (s.periodic) {
      if (s.n_src != n_dst) return fail("fd_fused_step: periodic term %d changes the axis length", k);
      for (int p = 0; p < t.st.npts; ++p)
        if (t.st.shifts[p] <= -n_dst || t.st.shifts[p] >= n_dst)
          return fail("fd_fused_step: shift %d does not fit the periodic axis of %d points", t.st.shifts[p], n_dst);
      t.lo = 0;
      t.hi = n_dst;
    } else {
      t.lo = s.lo < 0 ? 0 : s.lo;
      t.hi = s.hi > n_dst ? n_dst : s.hi;
      for (int p = 0; p < t.st.npts && t.lo < t.hi; ++p)
        if (t.lo + t.st.shifts[p] < 0 || t.hi - 1 + t.st.shifts[p] >= s.n_src)
          return fail("fd_fused_step: term %d range [%d,%d) with shift %d leaves the source (n_src=%d)", k,
                      t.lo, t.hi, t.st.shifts[p], s.n_src);
    }
    int dims[3] = {d->n[0], d->n[1], d->n[2]};
    if (!t.fold_mod) dims[s.axis] = s.n_src;
    t.stride[2] = 1;
    t.stride[1] = dims[2];
    t.stride[0] = (int64_t)dims[1] * dims[2];
    t.sa = t.fold_mod ? t.stride[0] * t.fold_div : t.stride[s.axis];
  }
  for (int k = 0; k < FD_MAX_TERMS; ++k)
    if (k >= d->nterms) P->term[k] = P->term[0];
  for (int k = 0; k < FD_MAX_POST; ++k) {
    P->post[k].op = k < d->npost ? d->post[k].op : FD_POST_ADD;
    P->post[k].s = k < d->npost ? (T)d->post[k].s : (T)0;
    P->post[k].coef = k < d->npost ? (const T*)d->post[k].coef : nullptr;
    if (k < d->npost && d->post[k].op == FD_POST_MUL_FIELD && !d->post[k].coef)
      return fail("fd_fused_step: coefficient post-op %d has no field", k);
    if (k < d->npost && (d->post[k].op < FD_POST_MUL || d->post[k].op > FD_POST_MUL_FIELD))
      return fail("fd_fused_step: unknown post-op %d", d->post[k].op);
  }
  P->dt = (T)d->dt;
  P->dot_out = (T*)d->dot_out;
  for (int l = 0; l < 2; ++l) {
    P->prev[l] = (const T*)d->prev[l];
    P->out[l] = (T*)d->out[l];
    P->nan_count[l] = (long long*)d->nan_count[l];
    // prev[0] == NULL: level 0 is the assigned field itself (no Euler step)
    if (l < d->depth && (!d->out[l] || (!d->prev[l] && l != 0)))
      return fail("fd_fused_step: level %d needs prev and out", l);
    for (int a = 0; a < 3; ++a)
      for (int s = 0; s < 2; ++s) {
        P->face[l][a][s] = (const T*)d->face[l][a][s];
        if (l < d->depth && d->face[l][a][s]) P->any_face |= (1 << a);
      }
  }
  if (P->any_face & 1)
    for (int k = 0; k < d->nterms; ++k)
      if (P->term[k].fold_mod) return fail("fd_fused_step: boundary faces on dimension 0 cannot be combined with folded axes");
  P->flags = (long long*)d->flags;
  const bool no_box = d->box_hi[0] == 0 && d->box_hi[1] == 0 && d->box_hi[2] == 0 && d->box_lo[0] == 0 &&
                      d->box_lo[1] == 0 && d->box_lo[2] == 0;
  P->whole = 1;
  for (int a = 0; a < 3; ++a) {
    P->box_lo[a] = no_box ? 0 : d->box_lo[a];
    P->box_hi[a] = no_box ? d->n[a] : d->box_hi[a];
    if (P->box_lo[a] < 0 || P->box_hi[a] > d->n[a] || P->box_lo[a] > P->box_hi[a])
      return fail("fd_fused_step: box [%d,%d) outside dimension %d of %d points", P->box_lo[a], P->box_hi[a], a,
                  d->n[a]);
    if (P->box_lo[a] != 0 || P->box_hi[a] != d->n[a]) P->whole = 0;
  }
  for (int a = 0; a < 3; ++a) {
    P->rim_lo[a] = d->rim_lo[a];
    P->rim_hi[a] = d->rim_hi[a];
    if (d->rim_lo[a] < 0 || d->rim_hi[a] >= d->n[a] + (d->n[a] == 0))
      return fail("fd_fused_step: rim indices of dimension %d outside the slice", a);
  }
  return 0;
}

template <typename T>
__device__ __forceinline__ T apply_post(const DevFused<T>& P, T v, int64_t idx) {
  for (int k = 0; k < P.npost; ++k) {
    const DevPost<T>& po = P.post[k];
    switch (po.op) {
      case FD_POST_MUL: v = mul_rn(v, po.s); break;
      case FD_POST_ADD: v = add_rn(v, po.s); break;
      case FD_POST_SUB: v = sub_rn(v, po.s); break;
      case FD_POST_DIV: v = div_rn(v, po.s); break;
      case FD_POST_NEG: v = -v; break;
      case FD_POST_MUL_FIELD: v = mul_rn(v, __ldg(po.coef + idx)); break;
    }
  }
  return v;
}

// Value the reference pre-populated at time t+1 on the boundary faces (set_BC),
// NaN when the point is on no registered face.
template <typename T>
__device__ __noinline__ T stored_face_value(const DevFused<T>& P, int l, int i0, int i1, int i2) {
  const int n1 = P.n[1], n2 = P.n[2];
  T v = quiet_nan<T>();
  const T* f;
  if (i2 == P.rim_lo[2] && (f = P.face[l][2][0])) v = f[(int64_t)i0 * n1 + i1];
  if (i2 == P.rim_hi[2] && (f = P.face[l][2][1]) && is_nan(v)) v = f[(int64_t)i0 * n1 + i1];
  if (i1 == P.rim_lo[1] && (f = P.face[l][1][0]) && is_nan(v)) v = f[(int64_t)i0 * n2 + i2];
  if (i1 == P.rim_hi[1] && (f = P.face[l][1][1]) && is_nan(v)) v = f[(int64_t)i0 * n2 + i2];
  if (i0 == P.rim_lo[0] && (f = P.face[l][0][0]) && is_nan(v)) v = f[(int64_t)i1 * n2 + i2];
  if (i0 == P.rim_hi[0] && (f = P.face[l][0][1]) && is_nan(v)) v = f[(int64_t)i1 * n2 + i2];
  return v;
}

// rhs -> optional .dot store -> Euler chain with the NaN-merge of FD.py:541-576.
template <typename T>
__device__ __forceinline__ void fused_epilogue(const DevFused<T>& P, T rhs, int64_t idx, int i0, int i1,
                                               int i2, unsigned (&nan_cnt)[2], bool& clash) {
  if (P.dot_out) P.dot_out[idx] = rhs;
  const int fm = P.any_face;  // bit d: some level has a face on dimension d
  const bool rim = fm && (((fm & 4) && (i2 == P.rim_lo[2] || i2 == P.rim_hi[2])) ||
                          ((fm & 2) && (i1 == P.rim_lo[1] || i1 == P.rim_hi[1])) ||
                          ((fm & 1) && (i0 == P.rim_lo[0] || i0 == P.rim_hi[0])));
  T rate = rhs;
#pragma unroll
  for (int l = 0; l < 2; ++l) {
    if (l < P.depth) {
      T v = (l == 0 && !P.prev[0]) ? rate : add_rn(__ldg(P.prev[l] + idx), mul_rn(rate, P.dt));
      if (rim) {
        const T stored = stored_face_value(P, l, i0, i1, i2);
        if (is_nan(v))
          v = stored;
        else
          clash |= !is_nan(stored);
      }
      nan_cnt[l] += is_nan(v) ? 1u : 0u;
      P.out[l][idx] = v;
      rate = v;
    }
  }
}

}  // namespace fdb
