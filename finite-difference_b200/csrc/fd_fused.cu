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
      const int j = tm.axis == 0 ? i0 : (tm.axis == 1 ? i1 : i2);
      T d;
      if (!tm.periodic && (j < tm.lo || j >= tm.hi)) {
        d = quiet_nan<T>();
      } else {
        // offset of the source element with the axis coordinate taken out
        const int64_t base = (int64_t)i0 * tm.stride[0] + (int64_t)i1 * tm.stride[1] + i2 -
                             (int64_t)j * tm.stride[tm.axis];
        const int64_t sa = tm.stride[tm.axis];
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
  fused_kernel<T><<<grid, block, 0, (cudaStream_t)stream>>>(P);
  return check_launch("fd_fused_step");
}

}  // namespace fdb

extern "C" {
int fd_fused_step_f64(const fd_fused_t* desc, void* stream) { return fdb::fused_impl<double>(desc, stream); }
int fd_fused_step_f32(const fd_fused_t* desc, void* stream) { return fdb::fused_impl<float>(desc, stream); }
}
