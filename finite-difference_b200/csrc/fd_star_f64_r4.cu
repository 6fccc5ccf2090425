// f64, radius 4 instantiations of the star-stencil fused kernel.
#include "fd_star.cuh"

namespace fdb {
template void dispatch_axes<double, 4>(StarParams<double>&, int, int, cudaStream_t);
}
