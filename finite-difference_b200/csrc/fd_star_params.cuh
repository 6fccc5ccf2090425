// Star-stencil specialisation of the fused statement (the diffusion / wave
// family: every der term samples the SAME source level with the SAME integer
// shifts -R..R, one term per trailing axis).
//
//   * 128-bit vector loads/stores along the contiguous axis (2 doubles or
//     4 floats per thread);
//   * the x (slowest) axis is marched: the 2R+1 planes a point needs are kept
//     in registers and rotate, so every plane is fetched once per block;
//   * y and z neighbours come through L1 (they were fetched R iterations
//     earlier as the "new plane" by threads of the same block);
//   * periodic axes wrap the index in-kernel, fixed axes predicate the result
//     to NaN outside [lo, hi); boundary faces are merged in the epilogue;
//   * FAST instantiation: canonical expression  [c *] (d_x + d_y + d_z)  with
//     unit grid spacing -- no run-time dispatch inside the point loop.
//
// Arithmetic order is the reference's (FD.py:1054-1055, 1095, 1107-1110,
// 515-516): acc = 0; acc += w_p * f_p in point order; / dx**order; terms summed
// in the user's order; * scalars; prev + rhs*dt.  No FMA contraction.

#pragma once

#include "fd_fused.cuh"

namespace fdb {

constexpr int kMaxR = 4;

template <typename T>
struct StarAxis {
  T w[2 * kMaxR + 1];
  T scale, inv;
  int32_t div_mode, periodic, lo, hi;
};

template <typename T>
struct StarParams {
  DevFused<T> F;
  StarAxis<T> ax[3];        // by axis: 0 = x (slowest), 1 = y, 2 = z (contiguous)
  const T* src;
  int32_t order[3];         // axis of term k (user's summation order)
  int32_t subtract[3], negate[3], has_pre[3];
  T pre[3];
  int32_t nterms, x_chunk, has_mul;
  int32_t clean;            // no output of the box is outside a term's [lo, hi), no face touches it
  int32_t lean_expr;        // 0: not a lean expression; 1: c*(sum of ders in axis order); 2: same with per-term factors / minus signs
  int32_t only_z_bounded;   // the box is clean along x and y; only the contiguous axis has a rim in it
  int32_t scale_class;   // 0: every axis divides by 1; 1: exact reciprocal multiplies; 3: FMA-corrected; 2: IEEE division somewhere
  int32_t prev_is_src[2];
  T mul;
};

// Decide whether the statement has the star form; fill StarParams if so.
template <typename T>
static bool star_applicable(const DevFused<T>& F_in, StarParams<T>* P, int* radius, int* axes, int* mode) {
  constexpr int VN = Vec<T>::N;
  DevFused<T> F = F_in;
  {
    // a 2-D problem (n0 == 1, terms along y and z) is marched along y: view it
    // as (n1, 1, n2) with the y term on the marched axis
    bool y_term = false, x_term = false;
    for (int k = 0; k < F.nterms; ++k) {
      y_term |= F.term[k].axis == 1;
      x_term |= F.term[k].axis == 0;
    }
    if (F.n[0] == 1 && y_term && !x_term) {
      F.n[0] = F.n[1];
      F.n[1] = 1;
      for (int k = 0; k < F.nterms; ++k)
        if (F.term[k].axis == 1) F.term[k].axis = 0;
      for (int l = 0; l < 2; ++l)
        for (int sd = 0; sd < 2; ++sd) {
          F.face[l][0][sd] = F.face[l][1][sd];
          F.face[l][1][sd] = nullptr;
        }
      F.any_face = (F.any_face & 4) | ((F.any_face & 2) ? 1 : 0);
      F.rim_lo[0] = F.rim_lo[1];
      F.rim_hi[0] = F.rim_hi[1];
      F.rim_lo[1] = F.rim_hi[1] = 0;
      F.box_lo[0] = F.box_lo[1];
      F.box_hi[0] = F.box_hi[1];
      F.box_lo[1] = 0;
      F.box_hi[1] = 1;
    }
    // sub-ranges along the vectorised axis must fall on 128-bit vector boundaries
    if (F.box_lo[2] % VN != 0 || (F.box_hi[2] % VN != 0 && F.box_hi[2] != F.n[2])) return false;
  }
  if (F.nterms < 1 || F.nterms > 3) return false;
  for (int k = 0; k < F.nterms; ++k)
    if (F.term[k].fold_mod) return false;      // more than three spatial axes: general kernels
  const DevTerm<T>& t0 = F.term[0];
  const int npts = t0.st.npts;
  if (npts < 3 || npts % 2 == 0) return false;
  const int R = npts / 2;
  if (R > kMaxR) return false;
  int term_of_axis[3] = {-1, -1, -1};
  for (int k = 0; k < F.nterms; ++k) {
    const DevTerm<T>& t = F.term[k];
    if (t.src != t0.src || t.st.npts != npts || term_of_axis[t.axis] != -1) return false;
    for (int p = 0; p < npts; ++p)
      if (t.st.shifts[p] != p - R) return false;
    if (t.n_src != F.n[t.axis]) return false;
    if (F.n[t.axis] < npts) return false;   // wrap_or_clamp handles one wrap only
    term_of_axis[t.axis] = k;
  }
  // always a term along z; x and/or y optional
  if (term_of_axis[2] < 0) return false;
  const bool has_x = term_of_axis[0] >= 0, has_y = term_of_axis[1] >= 0;
  const int axes_code = has_x ? (has_y ? 3 : 4) : (has_y ? 2 : 1);
  const int canon[5][3] = {{2, 2, 2}, {2, 2, 2}, {1, 2, 2}, {0, 1, 2}, {0, 2, 2}};
  if (F.n[2] % VN != 0 || (F.n[2] / VN) < 2 * ((R + VN - 1) / VN) + 1) return false;
  if ((int64_t)F.n[1] * F.n[2] * (R + 1) >= (1LL << 31)) return false;
  auto aligned = [](const void* p) { return ((uintptr_t)p & 15u) == 0; };
  if (!aligned(t0.src) || (F.dot_out && !aligned(F.dot_out))) return false;
  for (int l = 0; l < F.depth; ++l)
    if (!aligned(F.prev[l]) || !aligned(F.out[l])) return false;

  P->F = F;
  P->src = t0.src;
  P->nterms = F.nterms;
  const int naxes = F.nterms;
  bool canonical = true, ordered = true;   // ordered: terms follow the axes, none negated
  int scale_class = 0;
  for (int k = 0; k < 3; ++k) {
    P->order[k] = 2;
    P->subtract[k] = P->negate[k] = P->has_pre[k] = 0;
    P->pre[k] = (T)0;
  }
  for (int a = 0; a < 3; ++a) {
    StarAxis<T>& A = P->ax[a];
    const int k = term_of_axis[a];
    const DevTerm<T>& t = F.term[k < 0 ? 0 : k];
    for (int p = 0; p < 2 * kMaxR + 1; ++p) A.w[p] = p < npts ? t.st.weights[p] : (T)0;
    A.scale = t.st.scale;
    A.inv = t.st.inv;
    A.div_mode = t.st.div_mode;
    A.periodic = t.periodic;
    A.lo = t.lo;
    A.hi = t.hi;
    if (k >= 0) {
      const int m = t.st.div_mode;
      if (m == 2 || scale_class == 2)
        scale_class = 2;
      else if (m > scale_class)
        scale_class = m;     // 0 < 1 < 3
    }
  }
  for (int k = 0; k < F.nterms; ++k) {
    const DevTerm<T>& t = F.term[k];
    P->order[k] = t.axis;
    P->subtract[k] = k > 0 ? t.subtract : 0;
    P->negate[k] = t.negate;
    P->has_pre[k] = t.has_pre;
    P->pre[k] = t.pre;
    if (t.axis != canon[axes_code][k] || t.negate || t.has_pre || (k > 0 && t.subtract)) canonical = false;
    if (t.axis != canon[axes_code][k] || t.negate) ordered = false;
    if (!t.has_pre) P->pre[k] = (T)1;        // multiplying by 1 is exact
  }
  P->has_mul = 0;
  P->mul = (T)1;
  bool simple_post = F.npost == 0;
  if (F.npost == 1 && F.post[0].op == FD_POST_MUL) {
    simple_post = true;
    P->has_mul = 1;
    P->mul = F.post[0].s;
  }
  for (int l = 0; l < 2; ++l) P->prev_is_src[l] = (l < F.depth && F.prev[l] == t0.src) ? 1 : 0;
  *radius = R;
  *axes = axes_code;
  // "clean" box: no output in it is outside a term's [lo, hi) and no boundary
  // face touches it, so the predicate-free instantiation is exact there
  bool clean = true, clean_xy = true;
  for (int k = 0; k < F.nterms; ++k) {
    const DevTerm<T>& t = F.term[k];
    if (!t.periodic && (F.box_lo[t.axis] < t.lo || F.box_hi[t.axis] > t.hi)) {
      clean = false;
      if (t.axis != 2) clean_xy = false;
    }
  }
  for (int a = 0; a < 3; ++a)
    for (int l = 0; l < F.depth; ++l) {
      const bool lo_in = F.face[l][a][0] && F.box_lo[a] <= F.rim_lo[a] && F.rim_lo[a] < F.box_hi[a];
      const bool hi_in = F.face[l][a][1] && F.box_lo[a] <= F.rim_hi[a] && F.rim_hi[a] < F.box_hi[a];
      if (lo_in || hi_in) {
        clean = false;
        if (a != 2) clean_xy = false;
      }
    }
  P->only_z_bounded = (!clean && clean_xy) ? 1 : 0;
  P->scale_class = scale_class;
  P->clean = clean ? 1 : 0;
  P->lean_expr = !simple_post ? 0 : (canonical ? 1 : (ordered ? 2 : 0));
  // the fast instantiations of the general kernel assume unit spacing (keeping
  // the scaling code out of them is worth 0.63 -> 0.46 ms on config 4); other
  // spacings take the lean kernel where it applies, else MODE 2
  *mode = (canonical && simple_post && scale_class == 0) ? (clean ? 0 : 1) : 2;
  return true;
}

}  // namespace fdb
