// TEST / BENCH INFRASTRUCTURE ONLY (see oracle/__init__.py): C-ABI doorway to the reference's own
// CUDA launcher, so that bench.py can time the UNMODIFIED reference kernel beside pde_block_kernel.
//
// The reference kernel source is NOT in this repository: oracle/build_ref.py compiles
// /root/reference/gpsig/covariance_op/untrunc_cov_op_gpu.cu where it lies (nvcc -arch=sm_100a) together
// with this file into oracle/_ref/libuntrunc_cov_ref.so.  The launcher declared below is the plain C++
// function of untrunc_cov_op_gpu.cu:71-81 (declared the same way by untrunc_cov_op_gpu.cc:43-50); it
// launches on the legacy default stream and checks nothing, so the check is done here.
#include <cuda_runtime.h>

void UntruncCovKernelLauncher(const double* incr, const int batch_samples, const int nb_incr, const int paths_length,
                              const int paths_length_, const int nb_diagonals, double* pdes_sol);

// incr: [A, P, P] increments (device), sol: [A, P+2, P+2] initialised to ones (device, overwritten in place as
// UntruncCovOp::Compute does, untrunc_cov_op_gpu.cc:61-102).  Same argument derivation as the op:
// paths_length = P+1 (= threads per block, so P+1 <= 1024), paths_length_ = P+2, nb_diagonals = 2(P+1)-1.
// Returns the cudaError_t of the launch (0 = ok).
extern "C" int gpsig_ref_untrunc_cov(const double* incr, int A, int P, double* sol) {
  UntruncCovKernelLauncher(incr, A, P, P + 1, P + 2, 2 * (P + 1) - 1, sol);
  return (int)cudaGetLastError();
}
