// Host side of the large-state-dimension path (lifted_kernels.cuh): block loop, instantiations.
#include <stdlib.h>

#include "lifted_kernels.cuh"
#include "lift_mma_kernel.cuh"

namespace gpsig {

template <typename T, int COLS, int MLEV>
static int lift_block(const LiftParams& P, const LiftMmaOperands* mma, cudaStream_t s) {
  const int npair_slots = P.nstreams * P.SP;
  bool done = false;
  if constexpr (sizeof(T) == 4) {
    if (mma && mma->xop) { if (int rc = launch_lift_mma<COLS>(P, *mma, s)) return rc; done = true; }
  }
  if (!done) {
    dim3 grid((unsigned)npair_slots, (unsigned)((P.L + 63) / 64));
    lift_tile_kernel<T, COLS><<<grid, 256, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("lift_tile_kernel");
  }
  lift_recur_kernel<T, COLS, MLEV><<<(unsigned)((P.nstreams + kLiftWarps - 1) / kLiftWarps), kLiftWarps * 32, 0, s>>>(P);
  GPSIG_LAUNCH_CHECK("lift_recur_kernel");
  return GPSIG_OK;
}

template <typename T, int COLS>
static int lift_block_mlev(const LiftParams& P, const LiftMmaOperands* mma, cudaStream_t s) {
  if (P.m <= 3) return lift_block<T, COLS, 3>(P, mma, s);
  if (P.m <= 4) return lift_block<T, COLS, 4>(P, mma, s);
  if (P.m <= 5) return lift_block<T, COLS, 5>(P, mma, s);
  if (P.m <= 6) return lift_block<T, COLS, 6>(P, mma, s);
  if (P.m <= 8) return lift_block<T, COLS, 8>(P, mma, s);
  GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported (1..8)", P.m);
}

// P: everything but the block fields (ib, ie, jb, je, SP, nci, nstreams, stream_elems, scratch)
template <typename T>
int launch_lift(LiftParams P, void* scratch, size_t scratch_bytes, const LiftMmaOperands* mma, cudaStream_t s) {
  // the plan is sized by the WHOLE matrix (not this call's row range): it must agree with gpsig_*_workspace_bytes
  const size_t npairs = P.diag ? (size_t)P.N : (size_t)P.N * (size_t)P.N2;
  if (npairs == 0 || (!P.diag && P.row_begin == P.row_end)) return GPSIG_OK;
  const LiftPlan pl = lift_plan(npairs, P.L, P.L2, sizeof(T));
  if (scratch_bytes < pl.scratch_bytes)
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small for the lifted tiles: need %zu bytes, got %zu", pl.scratch_bytes, scratch_bytes);
  P.SP = pl.SP; P.stream_elems = pl.stream_elems; P.scratch = scratch;
  const int r0 = P.diag ? 0 : P.row_begin, r1 = P.diag ? P.N : P.row_end;
  // rows per block: as many streams of SP rows as fit; then as many columns as fit next to them
  const long long max_rows = (long long)pl.max_streams * pl.SP;
  const int NI = (int)((long long)(r1 - r0) < max_rows ? (r1 - r0) : max_rows);
  prof_begin(s);
  for (int ib = r0; ib < r1; ib += NI) {
    P.ib = ib; P.ie = ib + NI < r1 ? ib + NI : r1;
    P.nci = (P.ie - P.ib + P.SP - 1) / P.SP;
    int NJ = P.diag ? 1 : pl.max_streams / P.nci;
    if (NJ < 1) NJ = 1;
    if (!P.diag && NJ > P.N2) NJ = P.N2;
    const int jend = P.diag ? 1 : P.N2;
    for (int jb = 0; jb < jend; jb += NJ) {
      P.jb = jb; P.je = jb + NJ < jend ? jb + NJ : jend;
      P.nstreams = P.nci * (P.diag ? 1 : (P.je - P.jb));
      int rc;
      switch (pl.COLS) {
        case 2: rc = lift_block_mlev<T, 2>(P, mma, s); break;
        case 4: rc = lift_block_mlev<T, 4>(P, mma, s); break;
        default: rc = lift_block_mlev<T, 8>(P, mma, s); break;
      }
      if (rc) return rc;
    }
  }
  prof_end(s);
  return GPSIG_OK;
}

template int launch_lift<float>(LiftParams, void*, size_t, const LiftMmaOperands*, cudaStream_t);
template int launch_lift<double>(LiftParams, void*, size_t, const LiftMmaOperands*, cudaStream_t);

}  // namespace gpsig
