// Lean instantiation of the star kernel for the commonest statement
//     F.dot = c * (d_x + [d_y +] d_z)(F.prev);  F.new = F.prev + F.dot*dt
// in a "clean" box (no output outside a fixed axis' [lo, hi), no boundary
// face, unit grid spacing), depth 1, no .dot store: the 512^3 benchmark case
// and the interior box of fixed-boundary problems.  Same arithmetic and the
// same register marching as star_kernel, minus every run-time option; each
// thread owns VY consecutive rows x VZ interleaved 128-bit vectors, which
// shares the y neighbours between a thread's rows.  Measured on 512^3 fp64:
// 0.362-0.368 ms against 0.41 ms for the general star kernel (profiles/lab/
// map_lab.cu).
#pragma once

#include "fd_star_params.cuh"

namespace fdb {

// SCALE: 0 = every axis divides by 1 (skipped), 1 = exact reciprocal multiply
// (power-of-two spacing), 3 = FMA-corrected division by the invariant dx**order.
template <typename T, int SCALE>
__device__ __forceinline__ T lean_scale(T v, const StarAxis<T>& a) {
  if (SCALE == 0) return v;
  if (SCALE == 1) return mul_rn(v, a.inv);
  return div_invariant(v, a.scale, a.inv);
}

// Exact NaN count and infinite-der test (FD.py:1112, 528-529) for the points of
// one thread at plane x, recomputed from the (unchanged) source level.  Rarely
// executed: out of line, so that the marching loop holds no registers for it.
// Returns NaN count | inf << 16.
template <typename T, int R, bool HY, int VY, int VZ, int SCALE, bool PRE>
__device__ __noinline__ unsigned lean_exact_checks(const StarParams<T>& P, int x, int y0, int vz0) {
  constexpr int VN = Vec<T>::N;
  constexpr int NP = 2 * R + 1;
  const DevFused<T>& F = P.F;
  const int n[3] = {F.n[0], F.n[1], F.n[2]};
  unsigned cnt = 0;
  bool inf = false;
  for (int r = 0; r < VY; ++r)
    for (int v = 0; v < VZ; ++v)
      for (int e = 0; e < VN; ++e) {
        const int i[3] = {x, y0 + r, (vz0 + 32 * v) * VN + e};
        const int64_t idx = ((int64_t)i[0] * n[1] + i[1]) * n[2] + i[2];
        T d[3] = {(T)0, (T)0, (T)0};
        for (int a = 0; a < 3; ++a) {
          if (a == 1 && !HY) continue;
          const StarAxis<T>& A = P.ax[a];
          const int64_t stride = a == 0 ? (int64_t)n[1] * n[2] : (a == 1 ? n[2] : 1);
          T acc = (T)0;
          for (int p = 0; p < NP; ++p) {
            const int j = wrap_or_clamp(i[a] + p - R, n[a], A.periodic != 0);
            acc = add_rn(acc, mul_rn(A.w[p], P.src[idx + (int64_t)(j - i[a]) * stride]));
          }
          d[a] = lean_scale<T, SCALE>(acc, A);
          inf |= is_inf(d[a]);
        }
        T total;
        if (HY) {
          if (PRE) {
            const T a = mul_rn(d[0], P.pre[0]), b = mul_rn(d[1], P.pre[1]), cz = mul_rn(d[2], P.pre[2]);
            const T ab = P.subtract[1] ? sub_rn(a, b) : add_rn(a, b);
            total = P.subtract[2] ? sub_rn(ab, cz) : add_rn(ab, cz);
          } else {
            total = add_rn(add_rn(d[0], d[1]), d[2]);
          }
        } else if (PRE) {
          const T a = mul_rn(d[0], P.pre[0]), cz = mul_rn(d[2], P.pre[1]);
          total = P.subtract[1] ? sub_rn(a, cz) : add_rn(a, cz);
        } else {
          total = add_rn(d[0], d[2]);
        }
        const T nv = add_rn(P.src[idx], mul_rn(mul_rn(total, P.mul), F.dt));
        cnt += is_nan(nv) ? 1u : 0u;
      }
  return cnt | (inf ? 0x10000u : 0u);
}

__device__ __forceinline__ float lean_abs(float v) { return fabsf(v); }
__device__ __forceinline__ double lean_abs(double v) { return fabs(v); }
__device__ __forceinline__ bool lean_finite(float v) { return v < __int_as_float(0x7f800000); }
__device__ __forceinline__ bool lean_finite(double v) { return v < __longlong_as_double(0x7ff0000000000000LL); }

// Compile-time experiment, off by default (-DFD_LEAN_TWO_STAGE=1): without BZ
// the NaN / infinity bookkeeping in two stages, as in the marching pair kernel
// (fd_pair_march.cuh): every iteration sums the magnitudes of the values it
// stores; a NaN or infinite der result always leaves a non-finite stored value
// behind, so a finite sum proves there is nothing to report, and only otherwise
// are the exact tests run after the march (lean_exact_checks).  It removes 13 of
// ~112 instructions per iteration, yet 512^3 fp64 runs at 0.401 ms against
// 0.363 ms with the per-value tests (profiles/lab/ab_two_stage.sh): the
// dependent magnitude sum and the state kept for the second stage cost more
// than the DSETPs they replace.  The per-value tests stay.
//
// BZ: the contiguous axis is a fixed axis inside the box: outputs outside its
// [lo, hi) are NaN and its two boundary faces are merged here (a z rim cannot be
// peeled off as a thin slab without giving up whole 128-bit vector groups).
// PRE: every term carries its own factor and sign:  c*((k0*d_x +/- k1*d_y) +/- k2*d_z)
template <typename T, int R, bool HY, int VY, int VZ, int SCALE, bool BZ, bool PRE>
__global__ void __launch_bounds__(128, R > 1 ? 4 : ((SCALE == 3 || BZ || PRE) ? 6 : 8))
star_lean_kernel(const __grid_constant__ StarParams<T> P) {
  constexpr int VN = Vec<T>::N;
  constexpr int NP = 2 * R + 1;
#ifndef FD_LEAN_TWO_STAGE
#define FD_LEAN_TWO_STAGE 0
#endif
  constexpr bool EXACT = BZ || !FD_LEAN_TWO_STAGE;   // per-value NaN / inf tests inside the loop
  static_assert(R <= VN, "z neighbours come from one vector on each side");
  using PK = Pack<T, VN>;
  const DevFused<T>& F = P.F;
  const int n0 = F.n[0], n1 = F.n[1], n2 = F.n[2];
  const int nvec = n2 / VN;
  // 3-D: a block is 32 lanes along z by blockDim.y row groups; 2-D (one row per
  // x plane): blockDim.y indexes further z groups instead
  const int zgroup = HY ? blockIdx.x : blockIdx.x * blockDim.y + threadIdx.y;
  const int vz0 = F.box_lo[2] / VN + zgroup * (32 * VZ) + threadIdx.x;   // vector v sits 32 vectors further
  const int y0 = F.box_lo[1] + (HY ? (blockIdx.y * blockDim.y + threadIdx.y) * VY : blockIdx.y);
  const int x_begin = F.box_lo[0] + blockIdx.z * P.x_chunk;
  const int x_end = min(x_begin + P.x_chunk, F.box_hi[0]);
  const int plane = n1 * n2;
  const StarAxis<T>& ax = P.ax[0];
  const StarAxis<T>& ay = P.ax[1];
  const StarAxis<T>& az = P.ax[2];
  unsigned nan_cnt = 0;
  bool saw_inf = false, clash = false, suspect = false;
  // the launch tiles the box exactly except for the last vector group of a row
  // (VZ == 1 only) and the last row group
  const bool active = (vz0 + 32 * (VZ - 1) < F.box_hi[2] / VN) && (y0 + VY <= F.box_hi[1]);

  if (active) {
    // neighbour offsets (elements) relative to (x, y0, vz0*VN)
    int zl[VZ], zr[VZ];
#pragma unroll
    for (int v = 0; v < VZ; ++v) {
      zl[v] = (wrap_or_clamp(vz0 + 32 * v - 1, nvec, az.periodic) - vz0) * VN;
      zr[v] = (wrap_or_clamp(vz0 + 32 * v + 1, nvec, az.periodic) - vz0) * VN;
    }
    int oym[HY ? R : 1], oyp[HY ? R : 1];
    if (HY) {
#pragma unroll
      for (int k = 0; k < R; ++k) {
        oym[k] = (wrap_or_clamp(y0 - 1 - k, n1, ay.periodic) - y0) * n2;
        oyp[k] = (wrap_or_clamp(y0 + VY + k, n1, ay.periodic) - y0) * n2;
      }
    }
    const T* col = P.src + (int64_t)y0 * n2 + (int64_t)vz0 * VN;
    // x window: planes x-R .. x+R in use, x+R+1 prefetched
    PK win[NP + 1][VY][VZ];
    auto load_plane = [&](PK (&w)[VY][VZ], int x) {
      const T* p = col + (int64_t)wrap_or_clamp(x, n0, ax.periodic) * plane;
#pragma unroll
      for (int r = 0; r < VY; ++r)
#pragma unroll
        for (int v = 0; v < VZ; ++v) w[r][v] = load_vec<T>(p + r * n2 + 32 * VN * v);
    };
#pragma unroll
    for (int p = 0; p < NP; ++p) load_plane(win[p], x_begin - R + p);
    const T* here = col + (int64_t)x_begin * plane;
    T* outp = F.out[0] + (int64_t)x_begin * plane + (int64_t)y0 * n2 + (int64_t)vz0 * VN;

    for (int x = x_begin; x < x_end; ++x) {
      load_plane(win[NP], x + R + 1);
      T mag = (T)0;                                  // sum of |stored values| (two-stage test)
      PK hm[HY ? R : 1][VZ], hp[HY ? R : 1][VZ];     // rows below / above the thread's rows
      if (HY) {
#pragma unroll
        for (int k = 0; k < R; ++k)
#pragma unroll
          for (int v = 0; v < VZ; ++v) {
            hm[k][v] = load_vec<T>(here + oym[k] + 32 * VN * v);
            hp[k][v] = load_vec<T>(here + oyp[k] + 32 * VN * v);
          }
      }
#pragma unroll
      for (int r = 0; r < VY; ++r) {
#pragma unroll
        for (int v = 0; v < VZ; ++v) {
          const PK c = win[R][r][v];
          // z window: R elements | the vector | R elements
          T zw[VN + 2 * R];
          if (R == 1) {
            zw[0] = __ldg(here + r * n2 + zl[v] + VN - 1);
            zw[VN + 1] = __ldg(here + r * n2 + zr[v]);
          } else {
            const PK l = load_vec<T>(here + r * n2 + zl[v]), rr = load_vec<T>(here + r * n2 + zr[v]);
#pragma unroll
            for (int k = 0; k < R; ++k) {
              zw[k] = l.v[VN - R + k];
              zw[VN + R + k] = rr.v[k];
            }
          }
#pragma unroll
          for (int e = 0; e < VN; ++e) zw[R + e] = c.v[e];
          PK res;
#pragma unroll
          for (int e = 0; e < VN; ++e) {
            const int j2 = (vz0 + 32 * v) * VN + e;          // global index along z
            T dz = (T)0;
#pragma unroll
            for (int p = 0; p < NP; ++p) dz = add_rn(dz, mul_rn(az.w[p], zw[e + p]));
            dz = lean_scale<T, SCALE>(dz, az);
            if (BZ && (j2 < az.lo || j2 >= az.hi)) dz = quiet_nan<T>();
            if (EXACT) saw_inf |= is_inf(dz);
            T dx = (T)0;
#pragma unroll
            for (int p = 0; p < NP; ++p) dx = add_rn(dx, mul_rn(ax.w[p], win[p][r][v].v[e]));
            dx = lean_scale<T, SCALE>(dx, ax);
            if (EXACT) saw_inf |= is_inf(dx);
            T total;
            if (HY) {
              T dy = (T)0;
#pragma unroll
              for (int p = 0; p < NP; ++p) {
                const int rr = r + p - R;          // row offset within the thread's rows
                const T f = (rr < 0) ? hm[rr < 0 ? -rr - 1 : 0][v].v[e]
                                     : (rr >= VY ? hp[rr >= VY ? rr - VY : 0][v].v[e] : win[R][rr < 0 ? 0 : (rr >= VY ? 0 : rr)][v].v[e]);
                dy = add_rn(dy, mul_rn(ay.w[p], f));
              }
              dy = lean_scale<T, SCALE>(dy, ay);
              if (EXACT) saw_inf |= is_inf(dy);
              if (PRE) {
                const T a = mul_rn(dx, P.pre[0]), b = mul_rn(dy, P.pre[1]), cz = mul_rn(dz, P.pre[2]);
                const T ab = P.subtract[1] ? sub_rn(a, b) : add_rn(a, b);
                total = P.subtract[2] ? sub_rn(ab, cz) : add_rn(ab, cz);
              } else {
                total = add_rn(add_rn(dx, dy), dz);
              }
            } else if (PRE) {
              const T a = mul_rn(dx, P.pre[0]), cz = mul_rn(dz, P.pre[1]);
              total = P.subtract[1] ? sub_rn(a, cz) : add_rn(a, cz);
            } else {
              total = add_rn(dx, dz);
            }
            total = mul_rn(total, P.mul);
            T nv = add_rn(c.v[e], mul_rn(total, F.dt));
            if (BZ && (j2 == F.rim_lo[2] || j2 == F.rim_hi[2])) {
              // boundary face of the contiguous axis (FD.py:541-576)
              const T* face = F.face[0][2][j2 == F.rim_lo[2] ? 0 : 1];
              if (face) {
                const T stored = face[(int64_t)x * n1 + y0 + r];
                if (is_nan(nv))
                  nv = stored;
                else
                  clash |= !is_nan(stored);
              }
            }
            if (EXACT)
              nan_cnt += is_nan(nv) ? 1u : 0u;
            else
              mag += lean_abs(nv);
            res.v[e] = nv;
          }
          store_vec<T>(outp + r * n2 + 32 * VN * v, res);
        }
      }
      if (!EXACT) suspect |= !lean_finite(mag);
      here += plane;
      outp += plane;
#pragma unroll
      for (int p = 0; p < NP; ++p)
#pragma unroll
        for (int r = 0; r < VY; ++r)
#pragma unroll
          for (int v = 0; v < VZ; ++v) win[p][r][v] = win[p + 1][r][v];
    }
  }
  if (!EXACT && suspect) {
    // after the march (nothing is live any more): exact tests over the thread's planes
    for (int x = x_begin; x < x_end; ++x) {
      const unsigned c = lean_exact_checks<T, R, HY, VY, VZ, SCALE, PRE>(P, x, y0, vz0);
      nan_cnt += c & 0xffffu;
      saw_inf |= (c >> 16) != 0;
    }
  }
  if (saw_inf) raise_flag(F.flags, FD_FLAG_INF);
  if (BZ && clash) raise_flag(F.flags, FD_FLAG_OVERRIDE);
  warp_add_count(F.nan_count[0], nan_cnt);
}

template <typename T, int R, bool HY, int VY, int VZ>
static void launch_lean(StarParams<T>& P, dim3 grid, dim3 block, bool bz, cudaStream_t stream) {
  note_kernel("star_lean f%d R=%d axes=%d vy=%d vz=%d scale=%d bz=%d pre=%d chunk=%d grid=%ux%ux%u block=%ux%u",
              (int)sizeof(T) * 8, R, HY ? 3 : 4, VY, VZ, P.scale_class, bz ? 1 : 0, P.lean_expr == 2 ? 1 : 0, P.x_chunk,
              grid.x, grid.y, grid.z, block.x, block.y);
#define FD_LEAN(SC)                                                                        \
  if (P.lean_expr == 2) {                                                                  \
    if (bz)                                                                                \
      star_lean_kernel<T, R, HY, VY, VZ, SC, true, true><<<grid, block, 0, stream>>>(P);   \
    else                                                                                   \
      star_lean_kernel<T, R, HY, VY, VZ, SC, false, true><<<grid, block, 0, stream>>>(P);  \
  } else if (bz) {                                                                         \
    star_lean_kernel<T, R, HY, VY, VZ, SC, true, false><<<grid, block, 0, stream>>>(P);    \
  } else {                                                                                 \
    star_lean_kernel<T, R, HY, VY, VZ, SC, false, false><<<grid, block, 0, stream>>>(P);   \
  }
  if (P.scale_class == 0) {
    FD_LEAN(0)
  } else if (P.scale_class == 1) {
    FD_LEAN(1)
  } else {
    FD_LEAN(3)
  }
#undef FD_LEAN
}

// Launch the lean kernel if the statement qualifies; returns false otherwise.
template <typename T, int R>
bool try_star_lean(StarParams<T>& P, int axes, int mode, cudaStream_t stream) {
  constexpr int VN = Vec<T>::N;
  if constexpr (R > VN || R > 2) {   // z neighbours come from one vector per side; radius 3, 4: general kernel
    return false;
  } else {
    const DevFused<T>& F = P.F;
    (void)mode;
    const bool bz = !P.clean && P.only_z_bounded;           // clean in x and y, fixed along z
    if (!P.lean_expr || !(P.clean || bz) || F.depth != 1 || !P.prev_is_src[0] || F.dot_out || (axes != 3 && axes != 4))
      return false;
    if (P.scale_class == 2) return false;                      // IEEE division somewhere: general kernel
    const int rows = F.box_hi[1] - F.box_lo[1], vecs = (F.box_hi[2] - F.box_lo[2]) / VN, nx = F.box_hi[0] - F.box_lo[0];
    if (rows <= 0 || vecs <= 0 || nx <= 0) return false;
    if (F.box_lo[2] % VN != 0 || F.box_hi[2] % VN != 0) return false;
    int chunk = 8;
    {
      // small grids: shorter marches so that every SM still gets several blocks
      const int64_t tiles = (int64_t)((vecs + 31) / 32) * ((rows + 7) / 8);
      while (chunk > 1 && tiles * ((nx + chunk - 1) / chunk) < 148 * 8) chunk >>= 1;
    }
    if (const int forced = tuning_chunk()) chunk = forced;
    P.x_chunk = chunk;
    const unsigned gz = (nx + chunk - 1) / chunk;
    if (gz > 65535 || vecs < 3) return false;
    if (axes == 3) {
      if (F.n[1] < 2 + 2 * R) return false;
      const int vy = rows % 2 == 0 ? 2 : 1;                   // rows per thread
      const int groups = rows / vy;
      int by = 4;                                  // row groups per block; the last block may be partial
      while (by > 1 && by / 2 >= groups) by >>= 1;
      dim3 grid((vecs + 31) / 32, (groups + by - 1) / by, gz), block(32, by, 1);
      if (grid.y > 65535) return false;
      if (vy == 2)
        launch_lean<T, R, true, 2, 1>(P, grid, block, bz, stream);
      else
        launch_lean<T, R, true, 1, 1>(P, grid, block, bz, stream);
    } else {
      if (rows > 65535) return false;
      if (vecs % 64 == 0) {
        const int groups = vecs / 64;
        int by = 4;                                  // z groups per block; the last block may be partial
        while (by > 1 && by / 2 >= groups) by >>= 1;
        dim3 grid((groups + by - 1) / by, rows, gz), block(32, by, 1);
        launch_lean<T, R, false, 1, 2>(P, grid, block, bz, stream);
      } else {
        const int groups = (vecs + 31) / 32;
        int by = 4;
        while (by > 1 && by / 2 >= groups) by >>= 1;
        dim3 grid((groups + by - 1) / by, rows, gz), block(32, by, 1);
        launch_lean<T, R, false, 1, 1>(P, grid, block, bz, stream);
      }
    }
    return true;
  }
}

}  // namespace fdb
