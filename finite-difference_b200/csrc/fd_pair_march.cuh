// Marching variant of the fused staggered pair (included by fd_pair.cu after
// PairAxis / PairParams / pair_scale).
//
// Same statement group and the same arithmetic as pair_kernel, for the
// canonical 3-D case (one axis entry per dimension, in x, y, z order, no .dot
// stores).  Each thread owns VY consecutive rows x one 128-bit vector and
// marches along x:
//   * p(x-1), p(x), p(x+1) stay in registers, p(x+2) and v_x(x+2) are prefetched;
//   * v_x.new at x+1, which the pressure update needs, is the value stored at
//     the next iteration: it is carried, not recomputed;
//   * v_y.new is needed at VY+1 rows, v_z.new at VN+1 elements: the neighbours
//     inside the thread's rows / vector are shared, only the outer one is extra;
//   * z neighbours are single-element loads.
// That brings the loads per row from 13 vectors to 7 vectors + 3 elements
// (VY = 1) and removes a third of the arithmetic.

#pragma once

namespace fdb {

// scaled up-stencil of p between one point and the one before it along the axis (FD.py:1054-1055, 1107-1110)
template <typename T, int SCALE>
__device__ __forceinline__ T pair_grad(const PairAxis<T>& A, T p_minus, T p_here) {
  const T g = add_rn(add_rn((T)0, mul_rn(A.u0, p_minus)), mul_rn(A.u1, p_here));
  return pair_scale<T, SCALE>(g, A.u_mode, A.us, A.ui);
}

// v.new at one point: v + (grad * c_up) * dt   (FD.py:845-863, 515-516)
template <typename T, int SCALE>
__device__ __forceinline__ T pair_vnew(const PairAxis<T>& A, T p_minus, T p_here, T v, T dt) {
  return add_rn(v, mul_rn(mul_rn(pair_grad<T, SCALE>(A, p_minus, p_here), A.c_up), dt));
}

// scaled down-stencil of v.new between one point and the next along the axis
template <typename T, int SCALE>
__device__ __forceinline__ T pair_div(const PairAxis<T>& A, T v_here, T v_next) {
  const T d = add_rn(add_rn((T)0, mul_rn(A.w0, v_here)), mul_rn(A.w1, v_next));
  return pair_scale<T, SCALE>(d, A.d_mode, A.ds, A.di);
}

__device__ __forceinline__ float abs_dev(float v) { return fabsf(v); }
__device__ __forceinline__ double abs_dev(double v) { return fabs(v); }
template <typename T>
__device__ __forceinline__ T infinity();
template <>
__device__ __forceinline__ float infinity<float>() { return __int_as_float(0x7f800000); }
template <>
__device__ __forceinline__ double infinity<double>() { return __longlong_as_double(0x7ff0000000000000LL); }

// Exact NaN counts and infinite-der test for the VY x VN points of one thread
// at plane x, recomputed from the (unchanged) input levels.  Rarely executed:
// kept out of line so that the marching loop does not hold registers for it.
// Returns (NaN count of p | inf << 16, NaN counts of v_x, v_y, v_z).
template <typename T, int SCALE, int VY>
__device__ __noinline__ uint4 pair_exact_checks(const PairParams<T>& P, int x, int y0, int z0) {
  constexpr int VN = Vec<T>::N;
  const int n[3] = {P.n[0], P.n[1], P.n[2]};
  unsigned cnt[4] = {0u, 0u, 0u, 0u};
  bool inf = false;
  for (int r = 0; r < VY; ++r)
    for (int e = 0; e < VN; ++e) {
      const int i[3] = {x, y0 + r, z0 + e};
      const int64_t idx = ((int64_t)i[0] * n[1] + i[1]) * n[2] + i[2];
      const T p0 = P.p_prev[idx];
      T total = (T)0;
      for (int a = 0; a < 3; ++a) {
        const PairAxis<T>& A = P.ax[a];
        const int64_t stride = a == 0 ? (int64_t)n[1] * n[2] : (a == 1 ? n[2] : 1);
        const int64_t im = idx + (int64_t)(wrap_or_clamp(i[a] - 1, n[a], A.periodic != 0) - i[a]) * stride;
        const int64_t ip = idx + (int64_t)(wrap_or_clamp(i[a] + 1, n[a], A.periodic != 0) - i[a]) * stride;
        const T g = pair_grad<T, SCALE>(A, P.p_prev[im], p0), h = pair_grad<T, SCALE>(A, p0, P.p_prev[ip]);
        const T v0 = add_rn(A.v_prev[idx], mul_rn(mul_rn(g, A.c_up), P.dt));
        const T v1 = add_rn(A.v_prev[ip], mul_rn(mul_rn(h, A.c_up), P.dt));
        const T d = pair_div<T, SCALE>(A, v0, v1);
        inf |= is_inf(g) || is_inf(d);
        cnt[1 + a] += is_nan(v0) ? 1u : 0u;
        total = a == 0 ? d : (A.subtract ? sub_rn(total, d) : add_rn(total, d));
      }
      const T pn = add_rn(p0, mul_rn(mul_rn(total, P.c_down), P.dt));
      cnt[0] += is_nan(pn) ? 1u : 0u;
    }
  return make_uint4(cnt[0] | (inf ? 0x10000u : 0u), cnt[1], cnt[2], cnt[3]);
}

// Infinite der results and NaN outputs (FD.py:1112, 528-529) are looked for in
// two stages.  Every iteration sums the magnitudes of the values it stores; an
// infinite or NaN intermediate (grad, div) always leaves a non-finite stored
// value behind (inf*c and inf+x are never finite), so a finite sum proves that
// there is nothing to report.  Only when the sum is not finite are the exact
// per-value tests run, after the march (pair_exact_checks).
#ifndef FD_PAIR_MARCH_BLOCKS
#define FD_PAIR_MARCH_BLOCKS(T, VY) (sizeof(T) == 8 ? ((VY) == 2 ? 3 : 5) : ((VY) == 2 ? 4 : 5))
#endif
template <typename T, int SCALE, int VY>
__global__ void __launch_bounds__(128, FD_PAIR_MARCH_BLOCKS(T, VY))
pair_march_kernel(const __grid_constant__ PairParams<T> P) {
  constexpr int VN = Vec<T>::N;
  using PK = Pack<T, VN>;
  const int n0 = P.n[0], n1 = P.n[1], n2 = P.n[2];
  const int nvec = n2 / VN;
  const int vz0 = P.box_lo[2] / VN + blockIdx.x * 32 + threadIdx.x;
  const int y0 = P.box_lo[1] + (blockIdx.y * blockDim.y + threadIdx.y) * VY;
  const int x_begin = P.box_lo[0] + blockIdx.z * P.x_chunk;
  const int x_end = min(x_begin + P.x_chunk, P.box_hi[0]);
  const int64_t plane = (int64_t)n1 * n2;
  const PairAxis<T>& AX = P.ax[0];
  const PairAxis<T>& AY = P.ax[1];
  const PairAxis<T>& AZ = P.ax[2];
  unsigned nan_p = 0, nan_x = 0, nan_y = 0, nan_z = 0;
  bool saw_inf = false, suspect = false;
  const bool active = vz0 < P.box_hi[2] / VN && y0 + VY <= P.box_hi[1];

  if (active) {
    // neighbour offsets (elements) relative to (x, y0, vz0*VN)
    const int zl = (wrap_or_clamp(vz0 - 1, nvec, AZ.periodic != 0) - vz0) * VN + VN - 1;
    const int zr = (wrap_or_clamp(vz0 + 1, nvec, AZ.periodic != 0) - vz0) * VN;
    const int oym = (wrap_or_clamp(y0 - 1, n1, AY.periodic != 0) - y0) * n2;
    const int oyp = (wrap_or_clamp(y0 + VY, n1, AY.periodic != 0) - y0) * n2;
    const bool perx = AX.periodic != 0;
    // running pointers at (x, y0, vz0*VN)
    const int64_t start = (int64_t)x_begin * plane + (int64_t)y0 * n2 + (int64_t)vz0 * VN;
    const T* ip = P.p_prev + start;
    const T* ix = AX.v_prev + start;
    const T* iy = AY.v_prev + start;
    const T* iz = AZ.v_prev + start;
    T* op = P.p_out + start;
    T* ox = AX.v_out + start;
    T* oy = AY.v_out + start;
    T* oz = AZ.v_out + start;

    PK pw[4][VY];         // p at x-1, x, x+1 and (prefetched) x+2
    PK vxn[VY];           // v_x.new at x (carried from the previous iteration)
    PK vx1[VY], vx2[VY];  // v_x.prev at x+1 and (prefetched) x+2
    {
      const int64_t dm = (int64_t)(wrap_or_clamp(x_begin - 1, n0, perx) - x_begin) * plane;
      const int64_t dp = (int64_t)(wrap_or_clamp(x_begin + 1, n0, perx) - x_begin) * plane;
#pragma unroll
      for (int r = 0; r < VY; ++r) {
        pw[0][r] = load_vec<T>(ip + dm + r * n2);
        pw[1][r] = load_vec<T>(ip + r * n2);
        pw[2][r] = load_vec<T>(ip + dp + r * n2);
        const PK v = load_vec<T>(ix + r * n2);
        vx1[r] = load_vec<T>(ix + dp + r * n2);
#pragma unroll
        for (int e = 0; e < VN; ++e) vxn[r].v[e] = pair_vnew<T, SCALE>(AX, pw[0][r].v[e], pw[1][r].v[e], v.v[e], P.dt);
      }
    }

    for (int x = x_begin; x < x_end; ++x) {
      {
        const int64_t d2 = (int64_t)(wrap_or_clamp(x + 2, n0, perx) - x) * plane;
#pragma unroll
        for (int r = 0; r < VY; ++r) {
          pw[3][r] = load_vec<T>(ip + d2 + r * n2);
          vx2[r] = load_vec<T>(ix + d2 + r * n2);
        }
      }
      // this plane: rows around the thread's rows, v_y, v_z, z neighbours
      const PK p_below = load_vec<T>(ip + oym), p_above = load_vec<T>(ip + oyp);
      PK vy[VY + 1], vz[VY];
      T p_zl[VY], p_zr[VY], vz_zr[VY];
#pragma unroll
      for (int r = 0; r < VY; ++r) {
        vy[r] = load_vec<T>(iy + r * n2);
        vz[r] = load_vec<T>(iz + r * n2);
        p_zl[r] = __ldg(ip + r * n2 + zl);
        p_zr[r] = __ldg(ip + r * n2 + zr);
        vz_zr[r] = __ldg(iz + r * n2 + zr);
      }
      vy[VY] = load_vec<T>(iy + oyp);

      // v_y.new at rows y0 .. y0+VY
      PK vyn[VY + 1];
#pragma unroll
      for (int k = 0; k <= VY; ++k)
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          const T pm = k == 0 ? p_below.v[e] : pw[1][k == 0 ? 0 : k - 1].v[e];
          const T pc = k < VY ? pw[1][k < VY ? k : 0].v[e] : p_above.v[e];
          vyn[k].v[e] = pair_vnew<T, SCALE>(AY, pm, pc, vy[k].v[e], P.dt);
        }

      T mag = (T)0;
#pragma unroll
      for (int r = 0; r < VY; ++r) {
        const PK pc = pw[1][r];
        T vzn[VN + 1];     // v_z.new at elements 0 .. VN of the row
#pragma unroll
        for (int e = 0; e < VN; ++e)
          vzn[e] = pair_vnew<T, SCALE>(AZ, e == 0 ? p_zl[r] : pc.v[e == 0 ? 0 : e - 1], pc.v[e], vz[r].v[e], P.dt);
        vzn[VN] = pair_vnew<T, SCALE>(AZ, pc.v[VN - 1], p_zr[r], vz_zr[r], P.dt);
        PK out_p, out_z, next_x;
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          const T vx_next = pair_vnew<T, SCALE>(AX, pc.v[e], pw[2][r].v[e], vx1[r].v[e], P.dt);
          const T dx = pair_div<T, SCALE>(AX, vxn[r].v[e], vx_next);
          const T dy = pair_div<T, SCALE>(AY, vyn[r].v[e], vyn[r + 1].v[e]);
          const T dz = pair_div<T, SCALE>(AZ, vzn[e], vzn[e + 1]);
          T total = AY.subtract ? sub_rn(dx, dy) : add_rn(dx, dy);
          total = AZ.subtract ? sub_rn(total, dz) : add_rn(total, dz);
          const T pn = add_rn(pc.v[e], mul_rn(mul_rn(total, P.c_down), P.dt));
          mag += abs_dev(pn);
          mag += abs_dev(vxn[r].v[e]);
          mag += abs_dev(vyn[r].v[e]);
          mag += abs_dev(vzn[e]);
          out_p.v[e] = pn;
          out_z.v[e] = vzn[e];
          next_x.v[e] = vx_next;
        }
        store_vec<T>(op + r * n2, out_p);
        store_vec<T>(ox + r * n2, vxn[r]);
        store_vec<T>(oy + r * n2, vyn[r]);
        store_vec<T>(oz + r * n2, out_z);
        vxn[r] = next_x;
      }
      suspect |= !(mag < infinity<T>());
      ip += plane; ix += plane; iy += plane; iz += plane;
      op += plane; ox += plane; oy += plane; oz += plane;
#pragma unroll
      for (int r = 0; r < VY; ++r) {
        pw[0][r] = pw[1][r];
        pw[1][r] = pw[2][r];
        pw[2][r] = pw[3][r];
        vx1[r] = vx2[r];
      }
    }
  }
  if (suspect) {
    // something stored was NaN or infinite (or the magnitudes overflowed): exact
    // tests over the thread's planes, after the march (nothing is live any more)
    for (int x = x_begin; x < x_end; ++x) {
      const uint4 c = pair_exact_checks<T, SCALE, VY>(P, x, y0, vz0 * VN);
      nan_p += c.x & 0xffffu;
      nan_x += c.y;
      nan_y += c.z;
      nan_z += c.w;
      saw_inf |= (c.x >> 16) != 0;
    }
  }
  if (saw_inf) raise_flag(P.flags, FD_FLAG_INF);
  warp_add_count(P.p_nan, nan_p);
  warp_add_count(AX.v_nan, nan_x);
  warp_add_count(AY.v_nan, nan_y);
  warp_add_count(AZ.v_nan, nan_z);
}

// FD_B200_NO_MARCH=1 keeps pair_kernel for every pair (tests, comparisons);
// FD_B200_PAIR_CHUNK=<planes> / FD_B200_PAIR_VY=<1|2> override the marching
// length and the rows per thread (tuning runs only).
inline bool march_enabled() {
  const char* e = getenv("FD_B200_NO_MARCH");
  return !(e && e[0] == '1');
}
inline int pair_tuning(const char* name) {
  const char* e = getenv(name);
  return e ? atoi(e) : 0;
}

// Launch the marching kernel if the pair qualifies; returns false otherwise.
template <typename T>
bool try_pair_march(PairParams<T>& P, int naxes, bool any_dot, int scale_class, cudaStream_t stream) {
  constexpr int VN = Vec<T>::N;
  if (!march_enabled() || naxes != 3 || any_dot) return false;
  for (int a = 0; a < 3; ++a)
    if (P.ax[a].axis != a) return false;
  const int nv = (P.box_hi[2] - P.box_lo[2]) / VN, rows = P.box_hi[1] - P.box_lo[1], nx = P.box_hi[0] - P.box_lo[0];
  if (P.n[2] / VN < 3 || P.n[1] < 4) return false;
  // rows per thread: sharing the y neighbours pays in fp64 (6.1 TB/s against
  // 4.4 with one row); in fp32 (4-element vectors) the second row costs more
  // registers than it saves loads (5.5 against 5.8 TB/s)
  int vy = sizeof(T) == 8 ? 2 : 1;
  if (const int f = pair_tuning("FD_B200_PAIR_VY")) vy = f == 2 ? 2 : 1;
  if (rows % vy != 0) {
    if (sizeof(T) == 8 && !pair_tuning("FD_B200_PAIR_VY")) return false;
    vy = 1;
  }
  const int groups = rows / vy;
  int by = 4;
  while (by > 1 && by / 2 >= groups) by >>= 1;
  const int64_t tiles = (int64_t)((nv + 31) / 32) * ((groups + by - 1) / by);
  if ((groups + by - 1) / by > 65535) return false;
  // long marches on large grids (the planes around a chunk are read twice),
  // shorter ones when the grid would not fill the SMs otherwise; thin boxes
  // (rim slabs) and small grids stay with pair_kernel
  // (fp64 runs 3 blocks of 168 registers per SM and needs more blocks to pay off)
  int chunk = sizeof(T) == 8 ? 32 : 64;
  const int min_chunk = sizeof(T) == 8 ? 16 : 8;
  while (chunk > min_chunk && tiles * ((nx + chunk - 1) / chunk) < 148 * 8) chunk >>= 1;
  if (const int f = pair_tuning("FD_B200_PAIR_CHUNK"))
    chunk = f;
  else if (nx < 4 || tiles * ((nx + chunk - 1) / chunk) < 148 * (sizeof(T) == 8 ? 12 : 4))
    return false;
  while ((nx + chunk - 1) / chunk > 65535) chunk <<= 1;
  P.x_chunk = chunk;
  dim3 grid((nv + 31) / 32, (groups + by - 1) / by, (nx + chunk - 1) / chunk), block(32, by, 1);
  note_kernel("pair_march f%d vy=%d scale=%d chunk=%d grid=%ux%ux%u block=32x%d", (int)sizeof(T) * 8, vy, scale_class, chunk,
              grid.x, grid.y, grid.z, by);
#define FD_MARCH(SC)                                             \
  if (vy == 2)                                                   \
    pair_march_kernel<T, SC, 2><<<grid, block, 0, stream>>>(P);  \
  else                                                           \
    pair_march_kernel<T, SC, 1><<<grid, block, 0, stream>>>(P);
  if (scale_class == 0) {
    FD_MARCH(0)
  } else if (scale_class == 1) {
    FD_MARCH(1)
  } else {
    FD_MARCH(2)
  }
#undef FD_MARCH
  return true;
}

}  // namespace fdb
