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
//     to NaN outside [lo, hi); boundary faces are merged in the epilogue.
//
// Arithmetic order is the reference's (FD.py:1054-1055, 1095, 1107-1110,
// 515-516): acc = 0; acc += w_p * f_p in point order; / dx**order; terms summed
// in the user's order; * scalars; prev + rhs*dt.  No FMA contraction.

#include "fd_fused.cuh"

namespace fdb {

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

template <typename T>
struct StarParams {
  DevFused<T> F;
  const T* src;
  int32_t term_of_axis[3];  // index into F.term, -1: no der along that axis
  int32_t x_chunk;
  int32_t prev_is_src[2];
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

template <typename T, int R, bool HAS_X, bool HAS_Y>
__global__ void __launch_bounds__(256)
star_kernel(const __grid_constant__ StarParams<T> P) {
  constexpr int VN = Vec<T>::N;
  constexpr int NS = (R + VN - 1) / VN;   // neighbour vectors needed on each side along z
  constexpr int NP = 2 * R + 1;
  using PK = Pack<T, VN>;
  const DevFused<T>& F = P.F;
  const int n0 = F.n[0], n1 = F.n[1], n2 = F.n[2];
  const int nvec = n2 / VN;
  const int vz = blockIdx.x * blockDim.x + threadIdx.x;
  const int i1 = blockIdx.y * blockDim.y + threadIdx.y;
  const bool active = (vz < nvec) && (i1 < n1);
  const int i2 = vz * VN;

  const DevTerm<T>& tz = F.term[P.term_of_axis[2]];
  const DevTerm<T>& ty = F.term[HAS_Y ? P.term_of_axis[1] : 0];
  const DevTerm<T>& tx = F.term[HAS_X ? P.term_of_axis[0] : 0];

  unsigned nan_cnt[2] = {0u, 0u};
  bool saw_inf = false, clash = false;

  const int x_begin = blockIdx.z * P.x_chunk;
  const int x_end = min(x_begin + P.x_chunk, n0);
  const int64_t plane = (int64_t)n1 * n2;

  if (active) {
    const T* col = P.src + (int64_t)i1 * n2 + i2;   // (x=0, i1, i2)
    PK win[HAS_X ? NP : 1];
    if (HAS_X) {
#pragma unroll
      for (int p = 0; p < NP - 1; ++p)
        win[p] = load_vec<T>(col + wrap_or_clamp(x_begin - R + p, n0, tx.periodic) * plane);
    }
    // row offsets of the y neighbours (do not depend on x)
    int64_t yoff[HAS_Y ? NP : 1];
    if (HAS_Y) {
#pragma unroll
      for (int p = 0; p < NP; ++p)
        yoff[p] = (int64_t)(wrap_or_clamp(i1 - R + p, n1, ty.periodic) - i1) * n2;
    }
    int64_t zoff[2 * NS + 1];
#pragma unroll
    for (int q = 0; q < 2 * NS + 1; ++q)
      zoff[q] = (int64_t)(wrap_or_clamp(vz - NS + q, nvec, tz.periodic) - vz) * VN;

    for (int i0 = x_begin; i0 < x_end; ++i0) {
      const T* here = col + (int64_t)i0 * plane;
      PK center;
      if (HAS_X) {
        win[NP - 1] = load_vec<T>(col + wrap_or_clamp(i0 + R, n0, tx.periodic) * plane);
        center = win[R];
      } else {
        center = load_vec<T>(here);
      }
      // z window: (2 NS + 1) vectors around the thread's own
      T zw[(2 * NS + 1) * VN];
#pragma unroll
      for (int q = 0; q < 2 * NS + 1; ++q) {
        PK t = (q == NS) ? center : load_vec<T>(here + zoff[q]);
#pragma unroll
        for (int e = 0; e < VN; ++e) zw[q * VN + e] = t.v[e];
      }
      PK yv[HAS_Y ? NP : 1];
      if (HAS_Y) {
#pragma unroll
        for (int p = 0; p < NP; ++p) yv[p] = (p == R) ? center : load_vec<T>(here + yoff[p]);
      }

      PK rhs;
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        T d[3];
        {
          T acc = (T)0;
#pragma unroll
          for (int p = 0; p < NP; ++p) acc = add_rn(acc, mul_rn(tz.st.weights[p], zw[NS * VN - R + e + p]));
          acc = apply_scale(acc, tz.st.div_mode, tz.st.scale, tz.st.inv);
          const int j = i2 + e;
          if (!tz.periodic && (j < tz.lo || j >= tz.hi)) acc = quiet_nan<T>();
          saw_inf |= is_inf(acc);
          d[2] = acc;
        }
        if (HAS_Y) {
          T acc = (T)0;
#pragma unroll
          for (int p = 0; p < NP; ++p) acc = add_rn(acc, mul_rn(ty.st.weights[p], yv[p].v[e]));
          acc = apply_scale(acc, ty.st.div_mode, ty.st.scale, ty.st.inv);
          if (!ty.periodic && (i1 < ty.lo || i1 >= ty.hi)) acc = quiet_nan<T>();
          saw_inf |= is_inf(acc);
          d[1] = acc;
        }
        if (HAS_X) {
          T acc = (T)0;
#pragma unroll
          for (int p = 0; p < NP; ++p) acc = add_rn(acc, mul_rn(tx.st.weights[p], win[p].v[e]));
          acc = apply_scale(acc, tx.st.div_mode, tx.st.scale, tx.st.inv);
          if (!tx.periodic && (i0 < tx.lo || i0 >= tx.hi)) acc = quiet_nan<T>();
          saw_inf |= is_inf(acc);
          d[0] = acc;
        }
        // terms in the user's order
        T total = (T)0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
          if (k < F.nterms) {
            const DevTerm<T>& tm = F.term[k];
            T v = tm.axis == 2 ? d[2] : (tm.axis == 1 ? (HAS_Y ? d[1] : d[2]) : (HAS_X ? d[0] : d[2]));
            if (tm.negate) v = -v;
            if (tm.has_pre) v = mul_rn(v, tm.pre);
            total = (k == 0) ? v : (tm.subtract ? sub_rn(total, v) : add_rn(total, v));
          }
        }
        const int64_t idx = ((int64_t)i0 * n1 + i1) * n2 + i2 + e;
        rhs.v[e] = apply_post(F, total, idx);
      }

      // epilogue on vectors: optional .dot store, Euler chain, boundary merge
      const int64_t idx0 = ((int64_t)i0 * n1 + i1) * n2 + i2;
      if (F.dot_out) store_vec<T>(F.dot_out + idx0, rhs);
      const int fm = F.any_face;
      const bool rim_row = fm && (((fm & 2) && (i1 == 0 || i1 == n1 - 1)) || ((fm & 1) && (i0 == 0 || i0 == n0 - 1)));
      const bool rim_z = fm && (fm & 4) && (i2 == 0 || i2 + VN == n2);
      PK rate = rhs;
#pragma unroll
      for (int l = 0; l < 2; ++l) {
        if (l < F.depth) {
          const PK pv = P.prev_is_src[l] ? center : load_vec<T>(F.prev[l] + idx0);
          PK nv;
#pragma unroll
          for (int e = 0; e < VN; ++e) {
            T v = add_rn(pv.v[e], mul_rn(rate.v[e], F.dt));
            if (rim_row || (rim_z && (i2 + e == 0 || i2 + e == n2 - 1))) {
              const T stored = stored_face_value(F, l, i0, i1, i2 + e);
              if (is_nan(v))
                v = stored;
              else
                clash |= !is_nan(stored);
            }
            nan_cnt[l] += is_nan(v) ? 1u : 0u;
            nv.v[e] = v;
          }
          store_vec<T>(F.out[l] + idx0, nv);
          rate = nv;
        }
      }
      if (HAS_X) {
#pragma unroll
        for (int p = 0; p < NP - 1; ++p) win[p] = win[p + 1];
      }
    }
  }

  if (saw_inf) raise_flag(F.flags, FD_FLAG_INF);
  if (clash) raise_flag(F.flags, FD_FLAG_OVERRIDE);
  warp_add_count(F.nan_count[0], nan_cnt[0]);
  if (F.depth > 1) warp_add_count(F.nan_count[1], nan_cnt[1]);
}

// Decide whether the statement has the star form; fill StarParams if so.
template <typename T>
static bool star_applicable(const DevFused<T>& F, StarParams<T>* P, int* radius) {
  constexpr int VN = Vec<T>::N;
  if (F.nterms < 1 || F.nterms > 3) return false;
  const DevTerm<T>& t0 = F.term[0];
  const int npts = t0.st.npts;
  if (npts < 3 || npts % 2 == 0) return false;
  const int R = npts / 2;
  if (R > 4) return false;
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
  // one term per trailing axis: z alone, y+z, or x+y+z
  if (term_of_axis[2] < 0) return false;
  if (term_of_axis[0] >= 0 && term_of_axis[1] < 0) return false;
  if (F.n[2] % VN != 0 || (F.n[2] / VN) < 2 * ((R + VN - 1) / VN) + 1) return false;
  auto aligned = [](const void* p) { return ((uintptr_t)p & 15u) == 0; };
  if (!aligned(t0.src) || (F.dot_out && !aligned(F.dot_out))) return false;
  for (int l = 0; l < F.depth; ++l)
    if (!aligned(F.prev[l]) || !aligned(F.out[l])) return false;
  P->F = F;
  P->src = t0.src;
  for (int a = 0; a < 3; ++a) P->term_of_axis[a] = term_of_axis[a];
  for (int l = 0; l < 2; ++l) P->prev_is_src[l] = (l < F.depth && F.prev[l] == t0.src) ? 1 : 0;
  *radius = R;
  return true;
}

template <typename T, int R, bool HAS_X, bool HAS_Y>
static void launch_star(StarParams<T>& P, cudaStream_t stream) {
  constexpr int VN = Vec<T>::N;
  const int n0 = P.F.n[0], n1 = P.F.n[1], nvec = P.F.n[2] / VN;
  int bx = 32;
  while (bx > 1 && bx / 2 >= nvec) bx >>= 1;
  int by = 256 / bx;
  while (by > 1 && by / 2 >= n1) by >>= 1;
  if (bx * by < 32) by = 32 / bx;   // whole warps: the count reductions use full-warp shuffles
  const int gx = (nvec + bx - 1) / bx, gy = (n1 + by - 1) / by;
  // planes marched by one block: long enough to amortise the 2R preloaded
  // planes, short enough to give every SM several blocks
  int chunk = HAS_X ? 64 : 16;
  while (chunk > 1 && (int64_t)gx * gy * ((n0 + chunk - 1) / chunk) < 148 * 16) chunk >>= 1;
  while ((n0 + chunk - 1) / chunk > 65535) chunk <<= 1;
  P.x_chunk = chunk;
  dim3 grid(gx, gy, (n0 + chunk - 1) / chunk), block(bx, by, 1);
  star_kernel<T, R, HAS_X, HAS_Y><<<grid, block, 0, stream>>>(P);
}

template <typename T, int R>
static void dispatch_axes(StarParams<T>& P, cudaStream_t stream) {
  const bool hx = P.term_of_axis[0] >= 0, hy = P.term_of_axis[1] >= 0;
  if (hx)
    launch_star<T, R, true, true>(P, stream);
  else if (hy)
    launch_star<T, R, false, true>(P, stream);
  else
    launch_star<T, R, false, false>(P, stream);
}

// Returns 1 if the star kernel was launched, 0 if the statement does not have
// the star form (the caller then uses the general fused kernel), -1 on error.
template <typename T>
int try_star(const DevFused<T>& F, cudaStream_t stream) {
  StarParams<T> P;
  int R = 0;
  if (!star_applicable<T>(F, &P, &R)) return 0;
  if ((int64_t)F.n[1] > 65535LL * 8) return 0;
  switch (R) {
    case 1: dispatch_axes<T, 1>(P, stream); break;
    case 2: dispatch_axes<T, 2>(P, stream); break;
    case 3: dispatch_axes<T, 3>(P, stream); break;
    case 4: dispatch_axes<T, 4>(P, stream); break;
    default: return 0;
  }
  return check_launch("fd_fused_step(star)") ? -1 : 1;
}

template int try_star<double>(const DevFused<double>&, cudaStream_t);
template int try_star<float>(const DevFused<float>&, cudaStream_t);

}  // namespace fdb
