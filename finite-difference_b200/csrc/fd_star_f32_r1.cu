// f32, radius 1 instantiations of the star-stencil fused kernel.
#include "fd_star.cuh"

namespace fdb {
template void dispatch_axes<float, 1>(StarParams<float>&, int, int, cudaStream_t);
}
