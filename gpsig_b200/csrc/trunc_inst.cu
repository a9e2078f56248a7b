// One object per (T, OP, DP): nvcc -DGPSIG_T=double -DGPSIG_OP=1 -DGPSIG_DP=8
#include "tens_kernels.cuh"
#include "trunc_kernel.cuh"
#include "trunc_skew_kernel.cuh"
namespace gpsig {
template int launch_trunc<GPSIG_T, GPSIG_OP, GPSIG_DP>(const TruncParams&, cudaStream_t);
template int launch_trunc_skew<GPSIG_T, GPSIG_OP, GPSIG_DP>(const TruncParams&, cudaStream_t);
template int launch_tens_vs_seq<GPSIG_T, GPSIG_OP, GPSIG_DP>(const TensSeqParams&, int, cudaStream_t);
}
