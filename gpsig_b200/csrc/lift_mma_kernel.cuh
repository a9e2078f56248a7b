// float32 lift on the 5th-generation tensor cores (tcgen05.mma kind::tf32) -- see lifted_kernels.cuh.
#pragma once
#include "lifted_kernels.cuh"

namespace gpsig {

struct LiftMmaOperands {
  const void* xop; const void* yop;      // pre-split operands in the canonical K-major core-matrix layout, or NULL
  size_t xop_blk, yop_blk;               // elements per sequence
  int KP;                                // padded K of one split term
};

template <int COLS>
int launch_lift_mma(const LiftParams& P, const LiftMmaOperands& op, cudaStream_t s) {
  GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "tensor-core lift is not built");
}

}  // namespace gpsig
