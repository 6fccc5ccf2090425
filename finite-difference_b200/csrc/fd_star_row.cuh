// Wide star stencils (radius 3 and 4) on 2-D grids: the wave-equation family
//     F.dotdot = c * (d_x + d_z)(F.prev);  F.dot += dt*F.dotdot;  F += dt*F.dot
// (depth-2 Euler chain) and its depth-1 sibling, in a "clean" box (periodic
// axes, or the interior box of a fixed-boundary problem), canonical expression.
//
// The general star kernel fetches the 2R neighbours along the contiguous axis
// with extra vector loads through L1 (four 128-bit loads per thread and row at
// radius 4, a third of them L1 hits) and runs latency-bound at 0.71 of the HBM
// peak (profiles/r01_star_c4_ncu_summary.txt).  Here a warp owns one run of 32
// consecutive 128-bit vectors of a row and marches along the slow axis:
//   * the 2R+1 rows a point needs along the marched axis rotate through
//     registers, the row R+1 ahead is prefetched (as before);
//   * the neighbours along the contiguous axis are the *centre row registers of
//     the neighbouring lanes*: warp shuffles, no load at all.  The HL outermost
//     lanes on either side of the run only feed their neighbours (their own
//     results are not stored), so consecutive warps overlap by 2*HL vectors;
//   * the row entering the window and the row of the other integrated level are
//     copied to shared memory asynchronously several rows ahead (cp.async, one
//     private 16-byte slot per lane and row: no barrier).
// Arithmetic is the reference's, operation for operation (FD.py:1054-1055,
// 1107-1110, 515-516), as in star_kernel.
#pragma once

#include "fd_star_params.cuh"

namespace fdb {

template <typename T>
__device__ __forceinline__ Pack<T, Vec<T>::N> shfl_pack(const Pack<T, Vec<T>::N>& v, int src_lane) {
  Pack<T, Vec<T>::N> r;
#pragma unroll
  for (int e = 0; e < Vec<T>::N; ++e) r.v[e] = __shfl_sync(0xffffffffu, v.v[e], src_lane);
  return r;
}

template <typename T, int SCALE>
__device__ __forceinline__ T row_scale(T v, const StarAxis<T>& a) {
  if (SCALE == 0) return v;
  if (SCALE == 1) return mul_rn(v, a.inv);
  return div_invariant(v, a.scale, a.inv);
}

// ---- asynchronous staging: every lane copies the 16 bytes it will need kRowAhead
// rows later straight into its own slot of shared memory (cp.async), so that the
// bytes in flight per SM do not depend on the register budget (one row ahead in
// registers kept 16 KB per SM in flight and ran at 0.66 of the HBM peak with long-
// scoreboard stalls; the B200 needs about three times that).  A lane only ever
// reads back what it copied itself: no barrier, cp.async.wait_group alone.
constexpr int kRowAhead = 4;

__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// 4 warps per block, each an independent run; blockIdx.y = chunk of marched rows
template <typename T, int R, int DEPTH, int SCALE>
__global__ void __launch_bounds__(128, sizeof(T) == 8 ? 5 : 6)
star_row_kernel(const __grid_constant__ StarParams<T> P) {
  constexpr int VN = Vec<T>::N;
  constexpr int NP = 2 * R + 1;                // rows in the register window
  constexpr int HL = (R + VN - 1) / VN;        // feeding lanes on either side of a run
  constexpr int OUTL = 32 - 2 * HL;            // lanes whose results are stored
  constexpr int D = kRowAhead, SLOTS = D + 1;  // D copies in flight, one slot being read
  using PK = Pack<T, VN>;
  __shared__ __align__(16) PK stage[SLOTS][1 + DEPTH][128];   // [slot][0: differentiated level, 1+l: prev[l]][thread]
  const DevFused<T>& F = P.F;
  const int n0 = F.n[0], n2 = F.n[2];          // marched rows, contiguous extent (F.n[1] == 1)
  const int nvec = n2 / VN;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int seg = blockIdx.x * 4 + warp;
  const int v_out0 = F.box_lo[2] / VN + seg * OUTL;    // first vector this warp stores
  if (v_out0 >= F.box_hi[2] / VN) return;               // (whole warp)
  const int v = v_out0 - HL + lane;                     // this lane's vector
  const StarAxis<T>& ax = P.ax[0];
  const StarAxis<T>& az = P.ax[2];
  const int vw = wrap_or_clamp(v, nvec, az.periodic);
  const bool is_out = lane >= HL && lane < 32 - HL && v < F.box_hi[2] / VN;
  const int x_begin = F.box_lo[0] + blockIdx.y * P.x_chunk;
  const int x_end = min(x_begin + P.x_chunk, F.box_hi[0]);
  const int64_t col_off = (int64_t)vw * VN;
  const T* col = P.src + col_off;

  unsigned nan_cnt[2] = {0u, 0u};
  bool saw_inf = false;

  // what iteration j (row x_begin + j) needs beyond the rows already in the
  // window: the row R ahead of it, and its own row of the other integrated levels
  auto stage_rows = [&](const int j) {
    const int x = x_begin + j;
    if (x < x_end) {
      const int slot = j % SLOTS;
      cp_async_16(&stage[slot][0][tid], col + (int64_t)wrap_or_clamp(x + R, n0, ax.periodic) * n2);
#pragma unroll
      for (int l = 0; l < DEPTH; ++l)
        if (!P.prev_is_src[l] && is_out) cp_async_16(&stage[slot][1 + l][tid], F.prev[l] + (int64_t)x * n2 + col_off);
    }
    cp_async_commit();      // (possibly empty: the group count stays in step with the iterations)
  };
#pragma unroll
  for (int j = 0; j < D; ++j) stage_rows(j);

  PK win[NP];               // rows x-R .. x+R; the last one arrives through shared memory
#pragma unroll
  for (int p = 0; p < NP - 1; ++p) win[p] = load_vec<T>(col + (int64_t)wrap_or_clamp(x_begin - R + p, n0, ax.periodic) * n2);

  int64_t row_off = (int64_t)x_begin * n2 + col_off;
  int slot = 0;
  // one marched row; `u` is the position of the iteration within a group of NP:
  // the window rotates by index arithmetic that is static once the group is
  // unrolled, instead of by moving 2R+1 vectors from register to register
  auto body = [&](const int u, const int x) {
    cp_async_wait<D - 1>();                     // the copies issued for this iteration have landed
    win[(NP - 1 + u) % NP] = stage[slot][0][tid];
    PK cur[DEPTH];
#pragma unroll
    for (int l = 0; l < DEPTH; ++l)
      if (!P.prev_is_src[l]) cur[l] = stage[slot][1 + l][tid];
    stage_rows(x - x_begin + D);
    slot = slot + 1 == SLOTS ? 0 : slot + 1;
    const PK c = win[(R + u) % NP];
    // window along the contiguous axis: HL vectors | the lane's vector | HL vectors
    T zw[(2 * HL + 1) * VN];
#pragma unroll
    for (int k = 1; k <= HL; ++k) {
      const PK l = shfl_pack<T>(c, lane - k), r = shfl_pack<T>(c, lane + k);
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        zw[(HL - k) * VN + e] = l.v[e];
        zw[(HL + k) * VN + e] = r.v[e];
      }
    }
#pragma unroll
    for (int e = 0; e < VN; ++e) zw[HL * VN + e] = c.v[e];

    PK rhs;
#pragma unroll
    for (int e = 0; e < VN; ++e) {
      T dz = (T)0;
#pragma unroll
      for (int p = 0; p < NP; ++p) dz = add_rn(dz, mul_rn(az.w[p], zw[HL * VN - R + e + p]));
      dz = row_scale<T, SCALE>(dz, az);
      T dx = (T)0;
#pragma unroll
      for (int p = 0; p < NP; ++p) dx = add_rn(dx, mul_rn(ax.w[p], win[(p + u) % NP].v[e]));
      dx = row_scale<T, SCALE>(dx, ax);
      if (is_out) saw_inf |= is_inf(dz) | is_inf(dx);
      rhs.v[e] = mul_rn(add_rn(dx, dz), P.mul);      // canonical order: marched axis, then the contiguous one
    }
    if (is_out) {
      if (F.dot_out) store_vec<T>(F.dot_out + row_off, rhs);
      PK rate = rhs;
#pragma unroll
      for (int l = 0; l < DEPTH; ++l) {
        const PK pv = P.prev_is_src[l] ? c : cur[l];
        PK nv;
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          nv.v[e] = add_rn(pv.v[e], mul_rn(rate.v[e], F.dt));
          nan_cnt[l] += is_nan(nv.v[e]) ? 1u : 0u;
        }
        store_vec<T>(F.out[l] + row_off, nv);
        rate = nv;
      }
    }
    row_off += n2;
  };
  for (int x = x_begin; x < x_end; x += NP) {
#pragma unroll
    for (int u = 0; u < NP; ++u)
      if (x + u < x_end) body(u, x + u);
  }
  cp_async_wait<0>();
  if (saw_inf) raise_flag(F.flags, FD_FLAG_INF);
  warp_add_count(F.nan_count[0], nan_cnt[0]);
  if (DEPTH > 1) warp_add_count(F.nan_count[1], nan_cnt[1]);
}

// FD_B200_NO_ROW=1 keeps the general star kernel for wide 2-D stencils (tests, A/B runs).
inline bool row_enabled() {
  static const bool on = [] {
    const char* e = getenv("FD_B200_NO_ROW");
    return !(e && e[0] == '1');
  }();
  return on;
}

template <typename T, int R, int DEPTH>
static void launch_row_depth(StarParams<T>& P, dim3 grid, cudaStream_t stream) {
  if (P.scale_class == 0)
    star_row_kernel<T, R, DEPTH, 0><<<grid, 128, 0, stream>>>(P);
  else if (P.scale_class == 1)
    star_row_kernel<T, R, DEPTH, 1><<<grid, 128, 0, stream>>>(P);
  else
    star_row_kernel<T, R, DEPTH, 3><<<grid, 128, 0, stream>>>(P);
}

// Launch the row kernel if the statement qualifies; returns false otherwise.
template <typename T, int R>
bool try_star_row(StarParams<T>& P, int axes, cudaStream_t stream) {
  if constexpr (R < 3) {
    return false;
  } else {
    constexpr int VN = Vec<T>::N;
    constexpr int HL = (R + VN - 1) / VN, OUTL = 32 - 2 * HL;
    const DevFused<T>& F = P.F;
    if (!row_enabled() || axes != 4 || F.n[1] != 1 || !P.clean || P.lean_expr != 1 || P.scale_class == 2) return false;
    const int vecs = (F.box_hi[2] - F.box_lo[2]) / VN, nx = F.box_hi[0] - F.box_lo[0];
    if (vecs <= 0 || nx <= 0) return false;
    if (F.box_lo[2] % VN != 0 || F.box_hi[2] % VN != 0 || F.n[2] / VN < 32) return false;
    const int segs = (vecs + OUTL - 1) / OUTL;
    const unsigned gx = (segs + 3) / 4;
    // the 2R rows around a chunk are read twice (2R/chunk of one level), longer
    // marches leave fewer blocks for the last wave: 8192^2 radius 4 measured
    // 0.352 / 0.351 / 0.358 / 0.365 / 0.393 ms with 16 / 32 / 64 / 128 / 256 rows
    // (profiles/r02_c4_row_chunks.txt); shorter while SMs would be left without a block
    int chunk = 32;
    while (chunk > 8 && (int64_t)gx * ((nx + chunk - 1) / chunk) < 148 * 4) chunk >>= 1;
    if (const int forced = tuning_chunk()) chunk = forced;
    while ((nx + chunk - 1) / chunk > 65535) chunk <<= 1;
    P.x_chunk = chunk;
    dim3 grid(gx, (nx + chunk - 1) / chunk, 1);
    note_kernel("star_row f%d R=%d depth=%d scale=%d chunk=%d grid=%ux%u block=128", (int)sizeof(T) * 8, R, F.depth,
                P.scale_class, chunk, grid.x, grid.y);
    if (F.depth == 2)
      launch_row_depth<T, R, 2>(P, grid, stream);
    else
      launch_row_depth<T, R, 1>(P, grid, stream);
    return true;
  }
}

}  // namespace fdb
