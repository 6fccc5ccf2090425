// Fused staggered pair: all velocity updates and the pressure update of one
// cell/edge leapfrog step in a single pass (config 5 of BASELINE.json).
//
// Per point (i) and axis a with unit vector e_a, in the reference's order of
// operations (FD.py:1054-1055, 1107-1110, 845-863, 515-516):
//     g(j)      = ((0 + u0*p(j-e_a)) + u1*p(j)) / su          up.der(p.prev, a)
//     vnew_a(j) = v_a(j) + (g(j)*c_up) * dt                    assign + integrate
//     d_a       = ((0 + w0*vnew_a(i)) + w1*vnew_a(i+e_a)) / sd down.der(v_a.new, a)
//     pnew      = p + ((d_0 +/- d_1 +/- d_2) * c_down) * dt
// vnew_a(i+e_a) is recomputed from p and v_a instead of being read back, so
// p, v_x, v_y, v_z are each read once and written once: 8*sizeof(T) bytes per
// point instead of 14*sizeof(T) for the four separate statements.
//
// One 128-bit vector (2 doubles / 4 floats) per thread along z; neighbours
// along x and y are vector loads of the adjacent plane / row (L1/L2),
// neighbours along z come from the adjacent vectors.

#include "fd_fused.cuh"

namespace fdb {

template <typename T>
struct PairAxis {
  const T* v_prev;
  T* v_out;
  T* vdot_out;
  long long* v_nan;
  T c_up;
  int32_t axis, periodic, subtract;
  T u0, u1, w0, w1;
  T us, ui, ds, di;          // scale / exact inverse of the two stencils
  int32_t u_mode, d_mode;
};

template <typename T>
struct PairParams {
  int32_t n[3];
  int32_t box_lo[3], box_hi[3];
  int32_t x_chunk;
  PairAxis<T> ax[3];
  const T* p_prev;
  T* p_out;
  T* pdot_out;
  long long* p_nan;
  long long* flags;
  T c_down, dt;
};

// SCALE 0: every stencil divides by 1 (skipped); 1: by powers of two (exact
// multiplication by the inverse); 2: general (IEEE division where needed).
template <typename T, int SCALE>
__device__ __forceinline__ T pair_scale(T v, int mode, T scale, T inv) {
  if (SCALE == 0) return v;
  if (SCALE == 1) return mul_rn(v, inv);
  return apply_scale(v, mode, scale, inv);
}

template <typename T, int NA, int SCALE>
__global__ void __launch_bounds__(256, sizeof(T) == 4 ? 3 : 2)
pair_kernel(const __grid_constant__ PairParams<T> P) {
  constexpr int VN = Vec<T>::N;
  using PK = Pack<T, VN>;
  const int n0 = P.n[0], n1 = P.n[1], n2 = P.n[2];
  const int nvec = n2 / VN;
  const int vz = P.box_lo[2] / VN + blockIdx.x * blockDim.x + threadIdx.x;
  const int i2 = vz * VN;
  const int i1 = P.box_lo[1] + blockIdx.y * blockDim.y + threadIdx.y;
  const bool active = i2 < P.box_hi[2] && i1 < P.box_hi[1];
  unsigned nan_p = 0, nan_v[NA];
#pragma unroll
  for (int a = 0; a < NA; ++a) nan_v[a] = 0;
  bool saw_inf = false;
  const int64_t row = n2, plane = (int64_t)n1 * n2;

  if (active) {
    // neighbour offsets that do not depend on x (elements)
    bool per0 = true, per1 = true, per2 = true;
#pragma unroll
    for (int a = 0; a < NA; ++a) {
      const bool pa = P.ax[a].periodic != 0;
      if (P.ax[a].axis == 0) per0 = pa;
      if (P.ax[a].axis == 1) per1 = pa;
      if (P.ax[a].axis == 2) per2 = pa;
    }
    const int64_t ym = (int64_t)(wrap_or_clamp(i1 - 1, n1, per1) - i1) * row;
    const int64_t yp = (int64_t)(wrap_or_clamp(i1 + 1, n1, per1) - i1) * row;
    const int zm = (wrap_or_clamp(vz - 1, nvec, per2) - vz) * VN;
    const int zp = (wrap_or_clamp(vz + 1, nvec, per2) - vz) * VN;

    for (int i0 = P.box_lo[0] + blockIdx.z * P.x_chunk, i0_end = min(i0 + P.x_chunk, P.box_hi[0]); i0 < i0_end; ++i0) {
      const int64_t idx = (int64_t)i0 * plane + (int64_t)i1 * row + i2;
      const int64_t xm = (int64_t)(wrap_or_clamp(i0 - 1, n0, per0) - i0) * plane;
      const int64_t xp = (int64_t)(wrap_or_clamp(i0 + 1, n0, per0) - i0) * plane;
      const T* pc = P.p_prev + idx;
      const PK p0 = load_vec<T>(pc);
      T total[VN];
#pragma unroll
      for (int a = 0; a < NA; ++a) {
        const PairAxis<T>& A = P.ax[a];
        // p and v one step back / forward along this axis
        T pm[VN], pp[VN], vc[VN], vp[VN];
        {
          const PK t = load_vec<T>(A.v_prev + idx);
#pragma unroll
          for (int e = 0; e < VN; ++e) vc[e] = t.v[e];
        }
        if (A.axis == 2) {
          const PK l = load_vec<T>(pc + zm), r = load_vec<T>(pc + zp), vr = load_vec<T>(A.v_prev + idx + zp);
#pragma unroll
          for (int e = 0; e < VN; ++e) {
            pm[e] = e == 0 ? l.v[VN - 1] : p0.v[e - 1];
            pp[e] = e == VN - 1 ? r.v[0] : p0.v[e + 1];
            vp[e] = e == VN - 1 ? vr.v[0] : vc[e + 1];
          }
        } else {
          const int64_t om = A.axis == 0 ? xm : ym, op = A.axis == 0 ? xp : yp;
          const PK l = load_vec<T>(pc + om), r = load_vec<T>(pc + op), vr = load_vec<T>(A.v_prev + idx + op);
#pragma unroll
          for (int e = 0; e < VN; ++e) {
            pm[e] = l.v[e];
            pp[e] = r.v[e];
            vp[e] = vr.v[e];
          }
        }
        PK vnew, vdot;
#pragma unroll
        for (int e = 0; e < VN; ++e) {
          // v update at this point ...
          T g = add_rn(add_rn((T)0, mul_rn(A.u0, pm[e])), mul_rn(A.u1, p0.v[e]));
          g = pair_scale<T, SCALE>(g, A.u_mode, A.us, A.ui);
          saw_inf |= is_inf(g);
          const T rhs0 = mul_rn(g, A.c_up);
          const T v0 = add_rn(vc[e], mul_rn(rhs0, P.dt));
          // ... and at the next point along the axis (recomputed, not re-read)
          T h = add_rn(add_rn((T)0, mul_rn(A.u0, p0.v[e])), mul_rn(A.u1, pp[e]));
          h = pair_scale<T, SCALE>(h, A.u_mode, A.us, A.ui);
          const T v1 = add_rn(vp[e], mul_rn(mul_rn(h, A.c_up), P.dt));
          T d = add_rn(add_rn((T)0, mul_rn(A.w0, v0)), mul_rn(A.w1, v1));
          d = pair_scale<T, SCALE>(d, A.d_mode, A.ds, A.di);
          saw_inf |= is_inf(d);
          total[e] = (a == 0) ? d : (A.subtract ? sub_rn(total[e], d) : add_rn(total[e], d));
          nan_v[a] += is_nan(v0) ? 1u : 0u;
          vnew.v[e] = v0;
          vdot.v[e] = rhs0;
        }
        store_vec<T>(A.v_out + idx, vnew);
        if (A.vdot_out) store_vec<T>(A.vdot_out + idx, vdot);
      }
      PK pnew, pdot;
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const T rhs = mul_rn(total[e], P.c_down);
        const T v = add_rn(p0.v[e], mul_rn(rhs, P.dt));
        nan_p += is_nan(v) ? 1u : 0u;
        pnew.v[e] = v;
        pdot.v[e] = rhs;
      }
      store_vec<T>(P.p_out + idx, pnew);
      if (P.pdot_out) store_vec<T>(P.pdot_out + idx, pdot);
    }
  }
  if (saw_inf) raise_flag(P.flags, FD_FLAG_INF);
  warp_add_count(P.p_nan, nan_p);
#pragma unroll
  for (int a = 0; a < NA; ++a) warp_add_count(P.ax[a].v_nan, nan_v[a]);
}

}  // namespace fdb

#include "fd_pair_march.cuh"

namespace fdb {

template <typename T>
static int pair_impl(const fd_pair_t* d, void* stream) {
  constexpr int VN = Vec<T>::N;
  if (!d) return fail("fd_fused_pair: descriptor is NULL");
  if (d->naxes < 1 || d->naxes > 3) return fail("fd_fused_pair: %d axes, supported 1..3", d->naxes);
  auto aligned = [](const void* p) { return ((uintptr_t)p & 15u) == 0; };
  PairParams<T> P;
  bool whole0 = d->box_hi[0] == 0 && d->box_hi[1] == 0 && d->box_hi[2] == 0;
  for (int a = 0; a < 3; ++a) {
    P.n[a] = d->n[a];
    P.box_lo[a] = whole0 ? 0 : d->box_lo[a];
    P.box_hi[a] = whole0 ? d->n[a] : d->box_hi[a];
    if (P.box_lo[a] < 0 || P.box_hi[a] > d->n[a] || P.box_lo[a] > P.box_hi[a])
      return fail("fd_fused_pair: box outside dimension %d", a);
  }
  if (d->n[2] % VN != 0 || d->n[2] < 3 * VN) return fail("fd_fused_pair: contiguous extent %d must be a multiple of %d", d->n[2], VN);
  if (P.box_lo[2] % VN != 0 || P.box_hi[2] % VN != 0) return fail("fd_fused_pair: box must fall on vector boundaries along z");
  if (!d->p_prev || !d->p_out || !aligned(d->p_prev) || !aligned(d->p_out) || (d->pdot_out && !aligned(d->pdot_out)))
    return fail("fd_fused_pair: p pointers must be non-NULL and 16-byte aligned");
  int seen[3] = {0, 0, 0};
  for (int a = 0; a < d->naxes; ++a) {
    const fd_pair_axis_t& s = d->ax[a];
    PairAxis<T>& A = P.ax[a];
    if (s.axis < 0 || s.axis > 2 || seen[s.axis]++) return fail("fd_fused_pair: bad or repeated axis");
    if (!s.v_prev || !s.v_out || !aligned(s.v_prev) || !aligned(s.v_out) || (s.vdot_out && !aligned(s.vdot_out)))
      return fail("fd_fused_pair: v pointers must be non-NULL and 16-byte aligned");
    if (s.up.npts != 2 || s.up.shifts[0] != -1 || s.up.shifts[1] != 0)
      return fail("fd_fused_pair: the cell->edge stencil must sample shifts (-1, 0)");
    if (s.down.npts != 2 || s.down.shifts[0] != 0 || s.down.shifts[1] != 1)
      return fail("fd_fused_pair: the edge->cell stencil must sample shifts (0, +1)");
    if (d->n[s.axis] < 3) return fail("fd_fused_pair: axis %d too short", s.axis);
    if (!s.periodic && (P.box_lo[s.axis] < 1 || P.box_hi[s.axis] > d->n[s.axis] - 1))
      return fail("fd_fused_pair: a non-periodic axis needs ghost planes around the box");
    DevStencil<T> up, down;
    if (make_dev_stencil<T>(&s.up, &up) || make_dev_stencil<T>(&s.down, &down)) return 1;
    A.v_prev = (const T*)s.v_prev;
    A.v_out = (T*)s.v_out;
    A.vdot_out = (T*)s.vdot_out;
    A.v_nan = (long long*)s.v_nan_count;
    A.c_up = (T)s.c_up;
    A.axis = s.axis;
    A.periodic = s.periodic;
    A.subtract = a > 0 ? s.subtract : 0;
    A.u0 = up.weights[0]; A.u1 = up.weights[1];
    A.w0 = down.weights[0]; A.w1 = down.weights[1];
    A.us = up.scale; A.ui = up.inv; A.u_mode = up.div_mode;
    A.ds = down.scale; A.di = down.inv; A.d_mode = down.div_mode;
  }
  for (int a = d->naxes; a < 3; ++a) P.ax[a] = P.ax[0];
  P.p_prev = (const T*)d->p_prev;
  P.p_out = (T*)d->p_out;
  P.pdot_out = (T*)d->pdot_out;
  P.p_nan = (long long*)d->p_nan_count;
  P.flags = (long long*)d->flags;
  P.c_down = (T)d->c_down;
  P.dt = (T)d->dt;

  const int nv = (P.box_hi[2] - P.box_lo[2]) / VN, rows = P.box_hi[1] - P.box_lo[1], nx = P.box_hi[0] - P.box_lo[0];
  if (nv <= 0 || rows <= 0 || nx <= 0) return 0;
  int bx = 32;
  while (bx < 64 && bx < nv) bx <<= 1;
  int by = 256 / bx;
  while (by > 1 && by / 2 >= rows) by >>= 1;
  if (bx * by < 32) by = 32 / bx;
  const int64_t gy = (rows + by - 1) / by;
  if (gy > 65535) return fail("fd_fused_pair: middle dimension too large");
  int chunk = 8;
  while (chunk > 1 && (int64_t)((nv + bx - 1) / bx) * gy * ((nx + chunk - 1) / chunk) < 148 * 16) chunk >>= 1;
  while ((nx + chunk - 1) / chunk > 65535) chunk <<= 1;
  P.x_chunk = chunk;
  dim3 grid((nv + bx - 1) / bx, (unsigned)gy, (nx + chunk - 1) / chunk), block(bx, by, 1);
  int scale_class = 0;
  for (int a = 0; a < d->naxes; ++a) {
    const int m = P.ax[a].u_mode > P.ax[a].d_mode ? P.ax[a].u_mode : P.ax[a].d_mode;
    if (m > scale_class) scale_class = m;
    // mode 0 under class 1: multiplying by inv == 1 is exact
  }
  cudaStream_t st = (cudaStream_t)stream;
  {
    bool any_dot = d->pdot_out != nullptr;
    for (int a = 0; a < d->naxes; ++a) any_dot |= d->ax[a].vdot_out != nullptr;
    if (try_pair_march<T>(P, d->naxes, any_dot, scale_class, st)) return check_launch("fd_fused_pair(march)");
  }
  note_kernel("pair f%d axes=%d scale=%d chunk=%d grid=%ux%ux%u block=%dx%d", (int)sizeof(T) * 8, d->naxes, scale_class, chunk,
              grid.x, grid.y, grid.z, bx, by);
#define FD_PAIR(NA)                                                        \
  if (scale_class == 0)                                                    \
    pair_kernel<T, NA, 0><<<grid, block, 0, st>>>(P);                      \
  else if (scale_class == 1)                                               \
    pair_kernel<T, NA, 1><<<grid, block, 0, st>>>(P);                      \
  else                                                                     \
    pair_kernel<T, NA, 2><<<grid, block, 0, st>>>(P);
  switch (d->naxes) {
    case 1: FD_PAIR(1) break;
    case 2: FD_PAIR(2) break;
    default: FD_PAIR(3) break;
  }
#undef FD_PAIR
  return check_launch("fd_fused_pair");
}

}  // namespace fdb

extern "C" {
int fd_fused_pair_f64(const fd_pair_t* desc, void* stream) { return fdb::pair_impl<double>(desc, stream); }
int fd_fused_pair_f32(const fd_pair_t* desc, void* stream) { return fdb::pair_impl<float>(desc, stream); }
}
