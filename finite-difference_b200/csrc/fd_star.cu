// Host side of the star-stencil fused kernel: recognise the star form, fill
// StarParams and hand over to the per-(dtype, radius) instantiations, which are
// compiled in separate translation units (fd_star_<dtype>_r<R>.cu) so that the
// library builds in parallel.

#include "fd_star_params.cuh"

namespace fdb {

template <typename T, int R>
void dispatch_axes(StarParams<T>& P, int axes, int mode, cudaStream_t stream);

// Returns 1 if the star kernel was launched, 0 if the statement does not have
// the star form (the caller then uses the general fused kernel), -1 on error.
template <typename T>
int try_star(const DevFused<T>& F, cudaStream_t stream) {
  StarParams<T> P;
  int R = 0, axes = 0;
  int mode = 2;
  if (!F.prev[0]) return 0;     // assign_update without an Euler chain: general kernels
  if (!star_applicable<T>(F, &P, &R, &axes, &mode)) return 0;
  if ((int64_t)F.n[1] > 65535LL * 8) return 0;
  switch (R) {
    case 1: dispatch_axes<T, 1>(P, axes, mode, stream); break;
    case 2: dispatch_axes<T, 2>(P, axes, mode, stream); break;
    case 3: dispatch_axes<T, 3>(P, axes, mode, stream); break;
    case 4: dispatch_axes<T, 4>(P, axes, mode, stream); break;
    default: return 0;
  }
  return check_launch("fd_fused_step(star)") ? -1 : 1;
}


template int try_star<double>(const DevFused<double>&, cudaStream_t);
template int try_star<float>(const DevFused<float>&, cudaStream_t);

}  // namespace fdb
