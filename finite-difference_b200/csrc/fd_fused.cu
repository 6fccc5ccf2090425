// Fused statement kernel: der terms + scalar arithmetic + Euler chain + BC
// merge in ONE pass over HBM (general form: any mix of sources, axes,
// staggered shifts and periodic/fixed axes; neighbours are read through
// L1/L2).  The star-stencil specialisation with shared-memory tile staging
// lives in fd_star.cu.
//
// Reference lines restated (src/finite_difference/FD.py):
//   der 990-1122 (periodic 1046-1055, fixed 1058-1105, scale 1107-1110,
//   inf check 1112-1113), arithmetic 794-869, assign_update 478-498,
//   time_integrate_update 500-538, merge 541-576.

#include <stdlib.h>

#include "fd_fused.cuh"

namespace fdb {

template <typename T>
__global__ void __launch_bounds__(256)
fused_kernel(const __grid_constant__ DevFused<T> P) {
  const int n1 = P.n[1], n2 = P.n[2];
  const int i2 = blockIdx.x * blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y * blockDim.y + threadIdx.y;
  const bool active = i2 >= P.box_lo[2] && i2 < P.box_hi[2] && i1 >= P.box_lo[1] && i1 < P.box_hi[1];
  const int x_lo = P.box_lo[0], x_hi = P.box_hi[0];
  unsigned nan_cnt[2] = {0u, 0u};
  bool saw_inf = false, clash = false;

  for (int i0 = x_lo + blockIdx.z * P.x_chunk, i0_end = min(i0 + P.x_chunk, x_hi); i0 < i0_end; ++i0) {
    if (!active) continue;
    T total = (T)0;
    for (int k = 0; k < P.nterms; ++k) {
      const DevTerm<T>& tm = P.term[k];
      const int j = term_index(tm, i0, i1, i2);
      T d;
      if (!tm.periodic && (j < tm.lo || j >= tm.hi)) {
        d = quiet_nan<T>();
      } else {
        // offset of the source element with the axis coordinate taken out
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
        d = apply_scale(acc, tm.st.div_mode, tm.st.scale, tm.st.inv);
        saw_inf |= is_inf(d);
      }
      if (tm.negate) d = -d;
      if (tm.has_pre) d = mul_rn(d, tm.pre);
      total = (k == 0) ? d : (tm.subtract ? sub_rn(total, d) : add_rn(total, d));
    }
    const int64_t idx = ((int64_t)i0 * n1 + i1) * n2 + i2;
    total = apply_post(P, total, idx);
    fused_epilogue(P, total, idx, i0, i1, i2, nan_cnt, clash);
  }

  if (saw_inf) raise_flag(P.flags, FD_FLAG_INF);
  if (clash) raise_flag(P.flags, FD_FLAG_OVERRIDE);
  warp_add_count(P.nan_count[0], nan_cnt[0]);
  if (P.depth > 1) warp_add_count(P.nan_count[1], nan_cnt[1]);
}


// ---------------------------------------------------------------------------
// Vectorised general kernel: every thread owns one 128-bit vector (2 doubles /
// 4 floats) along the contiguous axis.  Terms along x or y load whole vectors
// of the neighbouring rows / planes; terms along z read a three-vector window
// and pick the shifted elements out of registers.  Same arithmetic, same
// order as the scalar kernel.
// ---------------------------------------------------------------------------
template <typename T, int S>
__device__ __forceinline__ void pick_shift(const T (&win)[3 * Vec<T>::N], T (&out)[Vec<T>::N]) {
  constexpr int VN = Vec<T>::N;
#pragma unroll
  for (int e = 0; e < VN; ++e) out[e] = win[VN + e + S];
}

// out[e] = win[VN + e + s] for a run-time (warp-uniform) shift |s| <= VN
template <typename T>
__device__ __forceinline__ void pick(const T (&win)[3 * Vec<T>::N], int s, T (&out)[Vec<T>::N]) {
  switch (s) {
    case -4: if (Vec<T>::N >= 4) pick_shift<T, (Vec<T>::N >= 4 ? -4 : 0)>(win, out); break;
    case -3: if (Vec<T>::N >= 4) pick_shift<T, (Vec<T>::N >= 4 ? -3 : 0)>(win, out); break;
    case -2: pick_shift<T, -2>(win, out); break;
    case -1: pick_shift<T, -1>(win, out); break;
    case 0: pick_shift<T, 0>(win, out); break;
    case 1: pick_shift<T, 1>(win, out); break;
    case 2: pick_shift<T, 2>(win, out); break;
    case 3: if (Vec<T>::N >= 4) pick_shift<T, (Vec<T>::N >= 4 ? 3 : 0)>(win, out); break;
    case 4: if (Vec<T>::N >= 4) pick_shift<T, (Vec<T>::N >= 4 ? 4 : 0)>(win, out); break;
  }
}

template <typename T>
__global__ void __launch_bounds__(256, 2)
fused_vec_kernel(const __grid_constant__ DevFused<T> P) {
  constexpr int VN = Vec<T>::N;
  using PK = Pack<T, VN>;
  const int n1 = P.n[1], n2 = P.n[2];
  const int vz = P.box_lo[2] / VN + blockIdx.x * blockDim.x + threadIdx.x;
  const int i2 = vz * VN;
  const int i1 = P.box_lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
  const bool active = i2 < P.box_hi[2] && i1 < P.box_hi[1];
  const int x_lo = P.box_lo[0], x_hi = P.box_hi[0];
  unsigned nan_cnt[2] = {0u, 0u};
  bool saw_inf = false, clash = false;
  const int fm = P.any_face;

  for (int i0 = x_lo + blockIdx.z * P.x_chunk, i0_end = min(i0 + P.x_chunk, x_hi); i0 < i0_end; ++i0) {
    if (!active) continue;
    T total[VN];
    for (int k = 0; k < P.nterms; ++k) {
      const DevTerm<T>& tm = P.term[k];
      T acc[VN];
#pragma unroll
      for (int e = 0; e < VN; ++e) acc[e] = (T)0;
      if (tm.axis == 2) {
        // window: previous vector | own vector | next vector of the source row
        const int nv = tm.n_src / VN;
        const T* row = tm.src + (int64_t)i0 * tm.stride[0] + (int64_t)i1 * tm.stride[1];
        T win[3 * VN];
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const PK t = load_vec<T>(row + (int64_t)wrap_or_clamp(vz - 1 + q, nv, tm.periodic) * VN);
#pragma unroll
          for (int e = 0; e < VN; ++e) win[q * VN + e] = t.v[e];
        }
        for (int p = 0; p < tm.st.npts; ++p) {
          T val[VN];
          pick<T>(win, tm.st.shifts[p], val);
#pragma unroll
          for (int e = 0; e < VN; ++e) acc[e] = add_rn(acc[e], mul_rn(tm.st.weights[p], val[e]));
        }
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          acc[e] = apply_scale(acc[e], tm.st.div_mode, tm.st.scale, tm.st.inv);
          if (!tm.periodic && (i2 + e < tm.lo || i2 + e >= tm.hi)) acc[e] = quiet_nan<T>();
        }
      } else {
        const int j = term_index(tm, i0, i1, i2);
        const T* col = tm.src + (int64_t)i0 * tm.stride[0] + (int64_t)i1 * tm.stride[1] + i2 - (int64_t)j * tm.sa;
        const bool bad = !tm.periodic && (j < tm.lo || j >= tm.hi);
        for (int p = 0; p < tm.st.npts; ++p) {
          const int jj = wrap_or_clamp(j + tm.st.shifts[p], tm.n_src, tm.periodic);
          const PK t = load_vec<T>(col + (int64_t)jj * tm.sa);
#pragma unroll
          for (int e = 0; e < VN; ++e) acc[e] = add_rn(acc[e], mul_rn(tm.st.weights[p], t.v[e]));
        }
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          acc[e] = apply_scale(acc[e], tm.st.div_mode, tm.st.scale, tm.st.inv);
          if (bad) acc[e] = quiet_nan<T>();
        }
      }
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        T d = acc[e];
        saw_inf |= is_inf(d);
        if (tm.negate) d = -d;
        if (tm.has_pre) d = mul_rn(d, tm.pre);
        total[e] = (k == 0) ? d : (tm.subtract ? sub_rn(total[e], d) : add_rn(total[e], d));
      }
    }
    const int64_t idx0 = ((int64_t)i0 * n1 + i1) * n2 + i2;
    PK rhs;
#pragma unroll
    for (int e = 0; e < VN; ++e) rhs.v[e] = apply_post(P, total[e], idx0 + e);

    if (P.dot_out) store_vec<T>(P.dot_out + idx0, rhs);
    const bool rim_row = fm && (((fm & 2) && (i1 == P.rim_lo[1] || i1 == P.rim_hi[1])) ||
                                ((fm & 1) && (i0 == P.rim_lo[0] || i0 == P.rim_hi[0])));
    const bool rim_z = (fm & 4) && (i2 <= P.rim_lo[2] || i2 + VN > P.rim_hi[2]);
    PK rate = rhs;
#pragma unroll
    for (int l = 0; l < 2; ++l) {
      if (l < P.depth) {
        const bool assigned = l == 0 && !P.prev[0];     // level 0 is the assigned field itself
        PK pv = rate;
        if (!assigned) pv = load_vec<T>(P.prev[l] + idx0);
        PK nv;
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          T v = assigned ? rate.v[e] : add_rn(pv.v[e], mul_rn(rate.v[e], P.dt));
          if (rim_row || (rim_z && (i2 + e == P.rim_lo[2] || i2 + e == P.rim_hi[2]))) {
            const T stored = stored_face_value(P, l, i0, i1, i2 + e);
            if (is_nan(v))
              v = stored;
            else
              clash |= !is_nan(stored);
          }
          nan_cnt[l] += is_nan(v) ? 1u : 0u;
          nv.v[e] = v;
        }
        store_vec<T>(P.out[l] + idx0, nv);
        rate = nv;
      }
    }
  }

  if (saw_inf) raise_flag(P.flags, FD_FLAG_INF);
  if (clash) raise_flag(P.flags, FD_FLAG_OVERRIDE);
  warp_add_count(P.nan_count[0], nan_cnt[0]);
  if (P.depth > 1) warp_add_count(P.nan_count[1], nan_cnt[1]);
}

// Can the statement use one 128-bit vector per thread?
template <typename T>
static bool vec_applicable(const DevFused<T>& P) {
  constexpr int VN = Vec<T>::N;
  auto aligned = [](const void* p) { return ((uintptr_t)p & 15u) == 0; };
  if (P.n[2] % VN != 0 || P.n[2] < 3 * VN) return false;
  if (P.box_lo[2] % VN != 0 || (P.box_hi[2] % VN != 0 && P.box_hi[2] != P.n[2])) return false;
  for (int k = 0; k < P.nterms; ++k) {
    const DevTerm<T>& t = P.term[k];
    if (!aligned(t.src) || t.stride[1] % VN != 0 || t.stride[0] % VN != 0) return false;
    if (t.axis == 2) {
      if (t.n_src != P.n[2]) return false;     // staggered fixed rows have another pitch
      for (int p = 0; p < t.st.npts; ++p)
        if (t.st.shifts[p] < -VN || t.st.shifts[p] > VN) return false;
    } else if (t.periodic) {
      for (int p = 0; p < t.st.npts; ++p)
        if (t.st.shifts[p] <= -t.n_src || t.st.shifts[p] >= t.n_src) return false;
    }
  }
  for (int k = 0; k < P.npost; ++k)
    if (P.post[k].op == FD_POST_MUL_FIELD && !aligned(P.post[k].coef)) return false;
  if (P.dot_out && !aligned(P.dot_out)) return false;
  for (int l = 0; l < P.depth; ++l)
    if (!aligned(P.prev[l]) || !aligned(P.out[l])) return false;
  return true;
}

template <typename T>
int try_star(const DevFused<T>& F, cudaStream_t stream);  // fd_star.cu

static bool star_enabled() {
  static const bool on = [] {
    const char* v = getenv("FD_B200_NO_STAR");
    return !(v && v[0] == '1');
  }();
  return on;
}

template <typename T>
static int fused_impl(const fd_fused_t* d, void* stream) {
  DevFused<T> P;
  if (make_dev_fused<T>(d, &P)) return 1;
  const int n0 = P.n[0], n1 = P.n[1], n2 = P.n[2];
  if ((int64_t)n0 * n1 * n2 == 0) return 0;
  if (star_enabled()) {
    const int r = try_star<T>(P, (cudaStream_t)stream);
    if (r != 0) return r < 0 ? 1 : 0;
  }
  if (vec_applicable<T>(P)) {
    constexpr int VN = Vec<T>::N;
    const int nv = (P.box_hi[2] - P.box_lo[2]) / VN, rows = P.box_hi[1] - P.box_lo[1], nx = P.box_hi[0] - P.box_lo[0];
    if (nv <= 0 || rows <= 0 || nx <= 0) return 0;
    int bx = 32;
    while (bx < 64 && bx < nv) bx <<= 1;
    int by = 256 / bx;
    while (by > 1 && by / 2 >= rows) by >>= 1;
    const int64_t gy = (rows + by - 1) / by;
    if (gy <= 65535) {
      int chunk = 8;
      while (chunk > 1 && (int64_t)((nv + bx - 1) / bx) * gy * ((nx + chunk - 1) / chunk) < 148 * 8) chunk >>= 1;
      while ((nx + chunk - 1) / chunk > 65535) chunk <<= 1;
      P.x_chunk = chunk;
      dim3 grid((nv + bx - 1) / bx, (unsigned)gy, (nx + chunk - 1) / chunk), block(bx, by, 1);
      note_kernel("fused_vec f%d chunk=%d grid=%ux%ux%u block=%dx%d", (int)sizeof(T) * 8, chunk, grid.x, grid.y, grid.z, bx, by);
      fused_vec_kernel<T><<<grid, block, 0, (cudaStream_t)stream>>>(P);
      return check_launch("fd_fused_step(vector)");
    }
  }
  // threads along the contiguous axis first; 256 threads per block
  int bx = 32;
  while (bx < 128 && bx < n2) bx <<= 1;
  int by = 256 / bx;
  while (by > 1 && by / 2 >= n1) by >>= 1;
  dim3 block(bx, by, 1);
  int64_t gy = (n1 + by - 1) / by;
  if (gy > 65535) return fail("fd_fused_step: middle dimension %d too large", n1);
  // consecutive planes handled by one block: reuses neighbour planes through L1
  const int nx = P.box_hi[0] - P.box_lo[0];   // planes actually processed
  if (P.box_hi[1] <= P.box_lo[1] || P.box_hi[2] <= P.box_lo[2]) return 0;
  if (nx <= 0) return 0;
  int chunk = 8;
  while (chunk > 1 && (int64_t)((n2 + bx - 1) / bx) * gy * ((nx + chunk - 1) / chunk) < 148 * 8) chunk >>= 1;
  while ((nx + chunk - 1) / chunk > 65535) chunk <<= 1;
  P.x_chunk = chunk;
  dim3 grid((n2 + bx - 1) / bx, (unsigned)gy, (nx + chunk - 1) / chunk);
  note_kernel("fused f%d chunk=%d grid=%ux%ux%u block=%dx%d", (int)sizeof(T) * 8, chunk, grid.x, grid.y, grid.z, bx, by);
  fused_kernel<T><<<grid, block, 0, (cudaStream_t)stream>>>(P);
  return check_launch("fd_fused_step");
}

}  // namespace fdb

extern "C" {
int fd_fused_step_f64(const fd_fused_t* desc, void* stream) { return fdb::fused_impl<double>(desc, stream); }
int fd_fused_step_f32(const fd_fused_t* desc, void* stream) { return fdb::fused_impl<float>(desc, stream); }
}
