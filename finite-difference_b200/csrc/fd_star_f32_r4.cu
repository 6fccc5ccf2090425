// f32, radius 4 instantiations of the star-stencil fused kernel.
#include "fd_star.cuh"

namespace fdb {
template void dispatch_axes<float, 4>(StarParams<float>&, int, int, cudaStream_t);
}
