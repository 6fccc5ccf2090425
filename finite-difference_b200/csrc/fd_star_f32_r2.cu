// f32, radius 2 instantiations of the star-stencil fused kernel.
#include "fd_star.cuh"

namespace fdb {
template void dispatch_axes<float, 2>(StarParams<float>&, int, int, cudaStream_t);
}
