// Star-stencil fused kernel: device code and launchers (see fd_star_params.cuh
// for the design notes).  Included by the per-(dtype, radius) instantiation
// files only.
#pragma once

#include <stdlib.h>

#include "fd_star_params.cuh"

namespace fdb {

// MODE 0: canonical expression, unit spacing, every axis periodic (no range
//         predicates, no faces)   1: canonical expression, unit spacing,
//         any boundaries          2: everything decided at run time
// 3 blocks per SM (<= 80 registers) where that does not spill: 3-point
// stencils, and the 1-D / 2-D (y-marched) forms of the wider ones
template <typename T, int R, int AXES>
constexpr int star_min_blocks() {
  return (R <= 1 || AXES == 4 || AXES == 1) ? 3 : 2;
}

// VPT: 128-bit vectors per thread along z (1 or 2).  Two vectors halve the
// z-halo loads and the per-iteration bookkeeping per point.
template <typename T, int R, int AXES, int MODE, int VPT = 1, int OCC = star_min_blocks<T, R, AXES>()>
__global__ void __launch_bounds__(256, OCC)
star_kernel(const __grid_constant__ StarParams<T> P) {
  // AXES: 1 = {z}, 2 = {y,z}, 3 = {x,y,z}, 4 = {x,z} (x is the marched axis)
  constexpr bool HX = AXES == 3 || AXES == 4, HY = AXES == 2 || AXES == 3;
  constexpr int NT = AXES == 1 ? 1 : (AXES == 3 ? 3 : 2);
  constexpr bool FAST = MODE < 2, BOUNDED = MODE != 0;
  constexpr int VN = Vec<T>::N;
  constexpr int TE = VPT * VN;            // elements per thread per plane
  constexpr int NS = (R + VN - 1) / VN;   // neighbour vectors needed on each side along z
  constexpr int NP = 2 * R + 1;
  constexpr int NW = NP + 1;              // x window: NP planes in use + 1 prefetched
  constexpr int UNR = (HX && R <= 2) ? NW : 1;   // static rotation of the x window
  using PK = Pack<T, VN>;
  struct Row {
    T v[TE];
  };
  const DevFused<T>& F = P.F;
  const int n0 = F.n[0], n1 = F.n[1], n2 = F.n[2];
  const int nvec = n2 / VN;               // vectors per row
  // a block covers blockDim.x * VPT consecutive vectors of a row; vector v of a
  // thread sits blockDim.x vectors after vector v-1, so every load instruction
  // of a warp reads one contiguous 512-byte run
  const int vstride = blockDim.x;
  const int vz = F.box_lo[2] / VN + blockIdx.x * (blockDim.x * VPT) + threadIdx.x;   // first vector of this thread
  const int i1 = F.box_lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
  const bool active = (vz + (VPT - 1) * vstride < F.box_hi[2] / VN) && (i1 < F.box_hi[1]);
  const int i2 = vz * VN;
  const int estride = vstride * VN;       // elements between a thread's vectors

  const StarAxis<T>& az = P.ax[2];
  const StarAxis<T>& ay = P.ax[1];
  const StarAxis<T>& ax = P.ax[0];

  unsigned nan_cnt[2] = {0u, 0u};
  bool saw_inf = false, clash = false;

  const int x_lo = F.box_lo[0], x_hi = F.box_hi[0];
  const int x_begin = x_lo + blockIdx.z * P.x_chunk;
  const int x_end = min(x_begin + P.x_chunk, x_hi);
  const int plane = n1 * n2;               // elements; the host guarantees < 2^31

  auto load_row = [estride](const T* p) {
    Row r;
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      const PK t = load_vec<T>(p + v * estride);
#pragma unroll
      for (int e = 0; e < VN; ++e) r.v[v * VN + e] = t.v[e];
    }
    return r;
  };
  auto store_row = [estride](T* p, const Row& r) {
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      PK t;
#pragma unroll
      for (int e = 0; e < VN; ++e) t.v[e] = r.v[v * VN + e];
      store_vec<T>(p + v * estride, t);
    }
  };

  if (active) {
    const T* col = P.src + (int64_t)i1 * n2 + i2;   // (x = 0, i1, i2)
    Row win[HX ? NW : 1];
    if (HX) {
#pragma unroll
      for (int p = 0; p < NP; ++p)
        win[p] = load_row(col + (int64_t)wrap_or_clamp(x_begin - R + p, n0, ax.periodic) * plane);
    }
    // neighbour offsets within a plane, in elements (do not depend on x)
    int yoff[HY ? NP : 1];
    if (HY) {
#pragma unroll
      for (int p = 0; p < NP; ++p) yoff[p] = (wrap_or_clamp(i1 - R + p, n1, ay.periodic) - i1) * n2;
    }
    int zoff[VPT][2 * NS];                  // per vector: NS vectors to the left, NS to the right
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
#pragma unroll
      for (int q = 0; q < NS; ++q) {
        zoff[v][q] = (wrap_or_clamp(vz + v * vstride - NS + q, nvec, az.periodic) - vz) * VN;
        zoff[v][NS + q] = (wrap_or_clamp(vz + v * vstride + 1 + q, nvec, az.periodic) - vz) * VN;
      }
    }
    // fixed axes: outputs outside [lo, hi) are unknown (NaN)
    bool zbad[TE];
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const int j = i2 + v * estride + e;
        zbad[v * VN + e] = BOUNDED && !az.periodic && (j < az.lo || j >= az.hi);
      }
    }
    const bool ybad = BOUNDED && HY && !ay.periodic && (i1 < ay.lo || i1 >= ay.hi);
    const int fm = BOUNDED ? F.any_face : 0;
    const bool rim_y = (fm & 2) && (i1 == F.rim_lo[1] || i1 == F.rim_hi[1]);
    const bool rim_z = (fm & 4) && (i2 <= F.rim_lo[2] || i2 + (VPT - 1) * estride + VN > F.rim_hi[2]);

    int64_t idx0 = ((int64_t)x_begin * n1 + i1) * n2 + i2;
    // running pointers: one 64-bit add per iteration instead of re-deriving
    // every address from the indices
    const T* here = P.src + idx0;
    T* out0 = F.out[0] + idx0;
    int xnew = wrap_or_clamp(x_begin + R + 1, n0, ax.periodic);   // plane being prefetched
    const T* pnew = col + (int64_t)xnew * plane;

    auto body = [&](const int u, const int i0) {
      Row center;
      if (HX) {
        // prefetch the plane the NEXT iteration needs; this one is already here
        win[(NP + u) % NW] = load_row(pnew);
        center = win[(R + u) % NW];
        if (++xnew == n0) {
          if (ax.periodic) {
            xnew = 0;
            pnew = col;
          } else {
            xnew = n0 - 1;   // clamped: those outputs are outside [lo, hi) anyway
          }
        } else {
          pnew += plane;
        }
      } else {
        center = load_row(here);
      }
      // z window of each vector: NS vectors | the vector | NS vectors
      constexpr int ZW = (2 * NS + 1) * VN;
      T zw[VPT][ZW];
#pragma unroll
      for (int v = 0; v < VPT; ++v) {
#pragma unroll
        for (int q = 0; q < NS; ++q) {
          const PK l = load_vec<T>(here + zoff[v][q]);
          const PK r = load_vec<T>(here + zoff[v][NS + q]);
#pragma unroll
          for (int e = 0; e < VN; ++e) {
            zw[v][q * VN + e] = l.v[e];
            zw[v][(NS + 1 + q) * VN + e] = r.v[e];
          }
        }
#pragma unroll
        for (int e = 0; e < VN; ++e) zw[v][NS * VN + e] = center.v[v * VN + e];
      }
      Row yv[HY ? NP : 1];
      if (HY) {
#pragma unroll
        for (int p = 0; p < NP; ++p)
          if (p != R) yv[p] = load_row(here + yoff[p]);
      }
      const bool xbad = BOUNDED && HX && !ax.periodic && (i0 < ax.lo || i0 >= ax.hi);

      Row rhs;
#pragma unroll
      for (int e = 0; e < TE; ++e) {
        T d[3] = {(T)0, (T)0, (T)0};
        {
          T acc = (T)0;
#pragma unroll
          for (int p = 0; p < NP; ++p) acc = add_rn(acc, mul_rn(az.w[p], zw[e / VN][NS * VN - R + e % VN + p]));
          if (!FAST) acc = apply_scale(acc, az.div_mode, az.scale, az.inv);
          if (BOUNDED && zbad[e]) acc = quiet_nan<T>();
          saw_inf |= is_inf(acc);
          d[2] = acc;
        }
        if (HY) {
          T acc = (T)0;
#pragma unroll
          for (int p = 0; p < NP; ++p) acc = add_rn(acc, mul_rn(ay.w[p], p == R ? center.v[e] : yv[p].v[e]));
          if (!FAST) acc = apply_scale(acc, ay.div_mode, ay.scale, ay.inv);
          if (BOUNDED && ybad) acc = quiet_nan<T>();
          saw_inf |= is_inf(acc);
          d[1] = acc;
        }
        if (HX) {
          T acc = (T)0;
#pragma unroll
          for (int p = 0; p < NP; ++p) acc = add_rn(acc, mul_rn(ax.w[p], win[(p + u) % NW].v[e]));
          if (!FAST) acc = apply_scale(acc, ax.div_mode, ax.scale, ax.inv);
          if (BOUNDED && xbad) acc = quiet_nan<T>();
          saw_inf |= is_inf(acc);
          d[0] = acc;
        }
        T total;
        if (FAST) {
          // canonical order: terms follow the axes, all added
          total = (HX && HY) ? add_rn(add_rn(d[0], d[1]), d[2])
                             : (HX ? add_rn(d[0], d[2]) : (HY ? add_rn(d[1], d[2]) : d[2]));
          total = mul_rn(total, P.mul);   // 1.0 when the expression has no factor: exact
        } else {
          total = (T)0;
#pragma unroll
          for (int k = 0; k < NT; ++k) {
            T v = P.order[k] == 2 ? d[2] : (P.order[k] == 1 ? d[1] : d[0]);
            if (P.negate[k]) v = -v;
            if (P.has_pre[k]) v = mul_rn(v, P.pre[k]);
            total = (k == 0) ? v : (P.subtract[k] ? sub_rn(total, v) : add_rn(total, v));
          }
          total = apply_post(F, total, idx0 + (e / VN) * estride + e % VN);
        }
        rhs.v[e] = total;
      }

      // epilogue on vectors: optional .dot store, Euler chain, boundary merge
      if (F.dot_out) store_row(F.dot_out + idx0, rhs);
      const bool rim_row = rim_y || ((fm & 1) && (i0 == F.rim_lo[0] || i0 == F.rim_hi[0]));
      Row rate = rhs;
#pragma unroll
      for (int l = 0; l < 2; ++l) {
        if (l < F.depth) {
          Row nv;
          if (P.prev_is_src[l]) {
#pragma unroll
            for (int e = 0; e < TE; ++e) nv.v[e] = add_rn(center.v[e], mul_rn(rate.v[e], F.dt));
          } else {
            const Row pv = load_row(F.prev[l] + idx0);
#pragma unroll
            for (int e = 0; e < TE; ++e) nv.v[e] = add_rn(pv.v[e], mul_rn(rate.v[e], F.dt));
          }
#pragma unroll
          for (int e = 0; e < TE; ++e) {
            T v = nv.v[e];
            const int j2 = i2 + (e / VN) * estride + e % VN;
            if (BOUNDED && (rim_row || (rim_z && (j2 == F.rim_lo[2] || j2 == F.rim_hi[2])))) {
              const T stored = stored_face_value(F, l, i0, i1, j2);
              if (is_nan(v))
                v = stored;
              else
                clash |= !is_nan(stored);
            }
            nan_cnt[l] += is_nan(v) ? 1u : 0u;
            nv.v[e] = v;
          }
          store_row(l == 0 ? out0 : F.out[l] + idx0, nv);
          rate = nv;
        }
      }
      idx0 += plane;
      here += plane;
      out0 += plane;
    };

    if (UNR > 1) {
      for (int i0 = x_begin; i0 < x_end; i0 += UNR) {
#pragma unroll
        for (int u = 0; u < UNR; ++u)
          if (i0 + u < x_end) body(u, i0 + u);
      }
    } else {
      for (int i0 = x_begin; i0 < x_end; ++i0) {
        body(0, i0);
        if (HX) {
#pragma unroll
          for (int p = 0; p < NW - 1; ++p) win[p] = win[p + 1];
        }
      }
    }
  }

  if (saw_inf) raise_flag(F.flags, FD_FLAG_INF);
  if (clash) raise_flag(F.flags, FD_FLAG_OVERRIDE);
  warp_add_count(F.nan_count[0], nan_cnt[0]);
  if (F.depth > 1) warp_add_count(F.nan_count[1], nan_cnt[1]);
}

// FD_B200_NO_LEAN=1 keeps the general star kernel for every statement (tests).
inline bool lean_enabled() {
  static const bool on = [] {
    const char* e = getenv("FD_B200_NO_LEAN");
    return !(e && e[0] == '1');
  }();
  return on;
}

// FD_B200_STAR_CHUNK=<planes> overrides the marching length (tuning runs only).
inline int tuning_chunk() {
  static const int v = [] {
    const char* e = getenv("FD_B200_STAR_CHUNK");
    return e ? atoi(e) : 0;
  }();
  return v;
}

template <typename T, int R, int AXES, int MODE, int VPT>
static void launch_star_vpt(StarParams<T>& P, int bx, int by, cudaStream_t stream) {
  constexpr int VN = Vec<T>::N;
  // extents of the box being processed
  const int n0 = P.F.box_hi[0] - P.F.box_lo[0], n1 = P.F.box_hi[1] - P.F.box_lo[1];
  const int nvec = (P.F.box_hi[2] - P.F.box_lo[2]) / VN;
  if (n0 <= 0 || n1 <= 0 || nvec <= 0) return;
  const int gx = (nvec + bx * VPT - 1) / (bx * VPT), gy = (n1 + by - 1) / by;
  // planes marched by one block: long enough to amortise the 2R+1 preloaded
  // planes, short enough to give every SM several blocks
  int chunk = R <= 2 ? 8 : 16;
  while (chunk > 1 && (int64_t)gx * gy * ((n0 + chunk - 1) / chunk) < 148 * 16) chunk >>= 1;
  if (const int forced = tuning_chunk()) chunk = forced;
  while ((n0 + chunk - 1) / chunk > 65535) chunk <<= 1;
  P.x_chunk = chunk;
  dim3 grid(gx, gy, (n0 + chunk - 1) / chunk), block(bx, by, 1);
  note_kernel("star f%d R=%d axes=%d mode=%d vpt=%d chunk=%d grid=%ux%ux%u block=%dx%d", (int)sizeof(T) * 8, R, AXES, MODE,
              VPT, chunk, grid.x, grid.y, grid.z, bx, by);
  star_kernel<T, R, AXES, MODE, VPT><<<grid, block, 0, stream>>>(P);
}

template <typename T, int R, int AXES, int MODE>
static void launch_star(StarParams<T>& P, cudaStream_t stream) {
  constexpr int VN = Vec<T>::N;
  const int n1 = P.F.box_hi[1] - P.F.box_lo[1], nvec = (P.F.box_hi[2] - P.F.box_lo[2]) / VN;
  int by = 1;                            // rows per block
  while (by < 8 && by < n1) by <<= 1;
  int bx = 256 / by;                     // threads along the contiguous axis
  // two vectors per thread (R <= 2) when whole blocks tile the row exactly
  if (R <= 2) {
    int b2 = bx;
    while (b2 > 32 && nvec % (2 * b2) != 0) b2 >>= 1;
    if (nvec % (2 * b2) == 0) {
      launch_star_vpt<T, R, AXES, MODE, (R <= 2 ? 2 : 1)>(P, b2, by, stream);
      return;
    }
  }
  while (bx > 1 && bx / 2 >= nvec) bx >>= 1;
  if (bx * by < 32) by = 32 / bx;        // whole warps: the count reductions use full-warp shuffles
  launch_star_vpt<T, R, AXES, MODE, 1>(P, bx, by, stream);
}

}  // namespace fdb

#include "fd_star_lean.cuh"
#include "fd_star_row.cuh"

namespace fdb {

template <typename T, int R>
void dispatch_axes(StarParams<T>& P, int axes, int mode, cudaStream_t stream) {
  if (lean_enabled() && try_star_lean<T, R>(P, axes, mode, stream)) return;
  if (try_star_row<T, R>(P, axes, stream)) return;
#define FD_STAR(AX)                                    \
  if (mode == 0)                                       \
    launch_star<T, R, AX, 0>(P, stream);               \
  else if (mode == 1)                                  \
    launch_star<T, R, AX, 1>(P, stream);               \
  else                                                 \
    launch_star<T, R, AX, 2>(P, stream);
  if (axes == 3) {
    FD_STAR(3)
  } else if (axes == 4) {
    FD_STAR(4)
  } else if (axes == 2) {
    FD_STAR(2)
  } else {
    FD_STAR(1)
  }
#undef FD_STAR
}

}  // namespace fdb
