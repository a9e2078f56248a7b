// Translation unit of the tiled Kzz kernels (tens_tens_tile_kernels.cuh); abi.cu links the launchers.
#include "tens_tens_tile_kernels.cuh"

namespace gpsig {
template bool launch_tens_tens_tile<float>(const TensTensParams&, cudaStream_t);
template bool launch_tens_tens_tile<double>(const TensTensParams&, cudaStream_t);
template bool launch_tens_tens_tile_vjp<float>(const TensTensVjpParams&, cudaStream_t);
template bool launch_tens_tens_tile_vjp<double>(const TensTensVjpParams&, cudaStream_t);
}  // namespace gpsig
