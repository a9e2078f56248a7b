// Host side of the large-state-dimension path (lifted_kernels.cuh): block loop, instantiations.
#include <stdlib.h>

#include "lifted_kernels.cuh"
#include "lift_mma_kernel.cuh"
#include "tens_lifted_kernels.cuh"

namespace gpsig {

// lift on stream s, recursion on stream sr (== s, or the side stream when consecutive blocks are overlapped)
template <typename T, int COLS, int MLEV>
static int lift_block(const LiftParams& P, const LiftMmaOperands* mma, cudaStream_t s, cudaStream_t sr, cudaEvent_t lifted) {
  const int npair_slots = P.nstreams * P.SP;
  bool done = false;
  if constexpr (sizeof(T) == 4) {
    if (mma && mma->xop) { if (int rc = launch_lift_mma<COLS>(P, *mma, s)) return rc; done = true; }
  }
  if constexpr (sizeof(T) == 8) {
    if (!done && getenv("GPSIG_LIFT_TILE_OLD") == nullptr) {
      if (int rc = lift_tile2_wide(P.L, P.L2) ? launch_lift_tile2<T, COLS, 7, 7>(P, npair_slots, s) : launch_lift_tile2<T, COLS, 4, 4>(P, npair_slots, s))
        return rc;
      done = true;
    }
  }
  if (!done) {
    dim3 grid((unsigned)npair_slots, (unsigned)((P.L + 63) / 64));
    lift_tile_kernel<T, COLS><<<grid, 256, 0, s>>>(P);
    GPSIG_LAUNCH_CHECK("lift_tile_kernel");
  }
  if (sr != s) { GPSIG_CUDA_CHECK(cudaEventRecord(lifted, s)); GPSIG_CUDA_CHECK(cudaStreamWaitEvent(sr, lifted, 0)); }
  lift_recur_kernel<T, COLS, MLEV><<<(unsigned)((P.nstreams + kLiftWarps - 1) / kLiftWarps), kLiftWarps * 32, 0, sr>>>(P);
  GPSIG_LAUNCH_CHECK("lift_recur_kernel");
  return GPSIG_OK;
}

template <typename T, int COLS>
static int lift_block_mlev(const LiftParams& P, const LiftMmaOperands* mma, cudaStream_t s, cudaStream_t sr, cudaEvent_t ev) {
  if (P.m <= 3) return lift_block<T, COLS, 3>(P, mma, s, sr, ev);
  if (P.m <= 4) return lift_block<T, COLS, 4>(P, mma, s, sr, ev);
  if (P.m <= 5) return lift_block<T, COLS, 5>(P, mma, s, sr, ev);
  if (P.m <= 6) return lift_block<T, COLS, 6>(P, mma, s, sr, ev);
  if (P.m <= 8) return lift_block<T, COLS, 8>(P, mma, s, sr, ev);
  GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported (1..8)", P.m);
}

// Side stream and events of the block pipeline, per host thread and device (created on first use; streams and events
// are driver objects, not device memory).
constexpr int kLiftMaxDev = 64;
struct LiftSide { cudaStream_t stream = nullptr; cudaEvent_t lifted[2] = {nullptr, nullptr}, consumed[2] = {nullptr, nullptr}; };
static thread_local LiftSide g_side[kLiftMaxDev];
static LiftSide* lift_side() {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kLiftMaxDev) return nullptr;
  LiftSide& S = g_side[dev];
  if (!S.stream) {
    if (cudaStreamCreateWithFlags(&S.stream, cudaStreamNonBlocking) != cudaSuccess) { S.stream = nullptr; return nullptr; }
    for (int k = 0; k < 2; ++k) {
      if (cudaEventCreateWithFlags(&S.lifted[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
      if (cudaEventCreateWithFlags(&S.consumed[k], cudaEventDisableTiming) != cudaSuccess) return nullptr;
    }
  }
  return &S;
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
  P.SP = pl.SP; P.stream_elems = pl.stream_elems;
  const int r0 = P.diag ? 0 : P.row_begin, r1 = P.diag ? P.N : P.row_end;
  const size_t call_pairs = P.diag ? (size_t)(r1 - r0) : (size_t)(r1 - r0) * (size_t)P.N2;
  // More pairs than one block of scratch holds: the scratch is used as two halves and the recursion of block b (side
  // stream) overlaps the lift of block b + 1 (caller's stream): the lift is bound by the tensor pipe / its stores, the
  // recursion by its loads, and both fit on an SM together.  (GPSIG_LIFT_NO_OVERLAP: one stream, whole scratch.)
  int max_streams = pl.max_streams;
  LiftSide* side = nullptr;
  if (call_pairs > (size_t)pl.max_streams * pl.SP && pl.max_streams >= 2 && getenv("GPSIG_LIFT_NO_OVERLAP") == nullptr) side = lift_side();
  if (side) max_streams = pl.max_streams / 2;
  // rows per block: as many streams of SP rows as fit; then as many columns as fit next to them
  const long long max_rows = (long long)max_streams * pl.SP;
  const int NI = (int)((long long)(r1 - r0) < max_rows ? (r1 - r0) : max_rows);
  prof_begin(s);
  int blk = 0;
  for (int ib = r0; ib < r1; ib += NI) {
    P.ib = ib; P.ie = ib + NI < r1 ? ib + NI : r1;
    P.nci = (P.ie - P.ib + P.SP - 1) / P.SP;
    int NJ = P.diag ? 1 : max_streams / P.nci;
    if (NJ < 1) NJ = 1;
    if (!P.diag && NJ > P.N2) NJ = P.N2;
    const int jend = P.diag ? 1 : P.N2;
    for (int jb = 0; jb < jend; jb += NJ, ++blk) {
      P.jb = jb; P.je = jb + NJ < jend ? jb + NJ : jend;
      P.nstreams = P.nci * (P.diag ? 1 : (P.je - P.jb));
      const int h = blk & 1;
      P.scratch = static_cast<char*>(scratch) + (side ? (size_t)h * max_streams * pl.stream_elems * sizeof(T) : 0);
      // this half was last read by the recursion of block blk - 2
      if (side && blk >= 2) GPSIG_CUDA_CHECK(cudaStreamWaitEvent(s, side->consumed[h], 0));
      int rc;
      cudaStream_t sr = side ? side->stream : s;
      cudaEvent_t ev = side ? side->lifted[h] : nullptr;
      switch (pl.COLS) {
        case 2: rc = lift_block_mlev<T, 2>(P, mma, s, sr, ev); break;
        case 4: rc = lift_block_mlev<T, 4>(P, mma, s, sr, ev); break;
        default: rc = lift_block_mlev<T, 8>(P, mma, s, sr, ev); break;
      }
      if (rc) return rc;
      if (side) GPSIG_CUDA_CHECK(cudaEventRecord(side->consumed[h], sr));
    }
  }
  // join: the caller's stream continues after the last recursions
  if (side) {
    if (blk >= 1) GPSIG_CUDA_CHECK(cudaStreamWaitEvent(s, side->consumed[(blk - 1) & 1], 0));
    if (blk >= 2) GPSIG_CUDA_CHECK(cudaStreamWaitEvent(s, side->consumed[(blk - 2) & 1], 0));
  }
  prof_end(s);
  return GPSIG_OK;
}

template int launch_lift<float>(LiftParams, void*, size_t, const LiftMmaOperands*, cudaStream_t);
template int launch_lift<double>(LiftParams, void*, size_t, const LiftMmaOperands*, cudaStream_t);

// ---- Kzx at d > 16 (tens_lifted_kernels.cuh): blocks of sequences, GEMM + recursion per block ----
template <typename T, int TM, int TN>
static int tens_lift_gemm(const TensSeqParams& P, int DP, int n0, int nb, T* C, cudaStream_t s) {
  constexpr int BM = 16 * TM, BN = 16 * TN, BK = 16;
  const int R = P.m * (P.m + 1) / 2 * P.zs * P.Mz;
  const long E = (long)P.L * nb;
  const int nrb = (R + BM - 1) / BM, ncb = (int)((E + BN - 1) / BN);
  const size_t smem = ((size_t)2 * BK * ((BM + 1) + (BN + 1)) + BM + BN) * sizeof(T);
  auto kern = tens_lift_gemm_kernel<T, TM, TN>;
  if (smem > 48 * 1024) GPSIG_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int occ = 1;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, 256, smem) != cudaSuccess || occ < 1) occ = 1;
  const long nitems = (long)nrb * ncb;
  if (nitems > 0x7fffffffL || E > 0x7fffffffL) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "Kzx block too large (%ld tiles)", nitems);
  const long want = (long)occ * device_sm_count();
  kern<<<(unsigned)(nitems < want ? nitems : want), 256, smem, s>>>(P, DP, n0, nb, C, nrb, ncb);
  GPSIG_LAUNCH_CHECK("tens_lift_gemm_kernel");
  return GPSIG_OK;
}

template <typename T>
int launch_tens_seq_lifted(const TensSeqParams& P, int DP, void* scratch, size_t scratch_bytes, cudaStream_t s) {
  const int ncols = P.col_end - P.col_begin;
  if (ncols <= 0 || P.Mz <= 0) return GPSIG_OK;
  if (P.m < 1 || P.m > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported (1..8)", P.m);
  // the plan is sized by N (like the workspace query), the blocks cover this call's columns
  const int nbmax = tens_lift_block(P.Mz, P.zs, P.m, P.L, P.N, sizeof(T));
  if (scratch_bytes < tens_lift_bytes(P.Mz, P.zs, P.m, P.L, P.N, sizeof(T)))
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE, "workspace too small for the lifted Kzx values");
  T* C = static_cast<T*>(scratch);
  prof_begin(s);
  for (int n0 = 0; n0 < ncols; n0 += nbmax) {
    const int nb = ncols - n0 < nbmax ? ncols - n0 : nbmax;
    if (int rc = tens_lift_gemm<T, 7, 7>(P, DP, n0, nb, C, s)) return rc;
    dim3 grid((unsigned)((nb + 127) / 128), (unsigned)P.Mz);
    if (P.return_levels && (long)nb * P.Mz < 148L * 2048) {      // raw levels of a small block: one thread per level as well
      grid.z = (unsigned)P.m;
      if (P.m <= 2) tens_lift_recur_level_kernel<T, 2><<<grid, 128, 0, s>>>(P, n0, nb, C);
      else if (P.m <= 4) tens_lift_recur_level_kernel<T, 4><<<grid, 128, 0, s>>>(P, n0, nb, C);
      else if (P.m <= 6) tens_lift_recur_level_kernel<T, 6><<<grid, 128, 0, s>>>(P, n0, nb, C);
      else tens_lift_recur_level_kernel<T, 8><<<grid, 128, 0, s>>>(P, n0, nb, C);
      GPSIG_LAUNCH_CHECK("tens_lift_recur_level_kernel");
      continue;
    }
    if (P.m <= 2) tens_lift_recur_kernel<T, 2><<<grid, 128, 0, s>>>(P, n0, nb, C);
    else if (P.m <= 4) tens_lift_recur_kernel<T, 4><<<grid, 128, 0, s>>>(P, n0, nb, C);
    else if (P.m <= 6) tens_lift_recur_kernel<T, 6><<<grid, 128, 0, s>>>(P, n0, nb, C);
    else tens_lift_recur_kernel<T, 8><<<grid, 128, 0, s>>>(P, n0, nb, C);
    GPSIG_LAUNCH_CHECK("tens_lift_recur_kernel");
  }
  prof_end(s);
  return GPSIG_OK;
}
template <typename T>
int launch_tens_seq_lifted_gd(const TensSeqParams& P, int DP, T* C, const T* g, cudaStream_t s) {
  const int ncols = P.col_end - P.col_begin;
  if (ncols <= 0 || P.Mz <= 0) return GPSIG_OK;
  if (P.m < 1 || P.m > 8) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "num_levels=%d is not supported (1..8)", P.m);
  prof_begin(s);
  if (int rc = tens_lift_gemm<T, 7, 7>(P, DP, 0, ncols, C, s)) return rc;
  if (g) {
    const int by_value = P.base == GPSIG_BASE_RBF ? 1 : 0;
    dim3 grid((unsigned)((ncols + 127) / 128), (unsigned)P.Mz, (unsigned)P.m);      // one thread per (sequence, tensor, level)
    if (P.m <= 2) tens_lift_reverse_kernel<T, 2><<<grid, 128, 0, s>>>(P, ncols, C, g, by_value);
    else if (P.m <= 4) tens_lift_reverse_kernel<T, 4><<<grid, 128, 0, s>>>(P, ncols, C, g, by_value);
    else if (P.m <= 6) tens_lift_reverse_kernel<T, 6><<<grid, 128, 0, s>>>(P, ncols, C, g, by_value);
    else tens_lift_reverse_kernel<T, 8><<<grid, 128, 0, s>>>(P, ncols, C, g, by_value);
    GPSIG_LAUNCH_CHECK("tens_lift_reverse_kernel");
  }
  prof_end(s);
  return GPSIG_OK;
}
template int launch_tens_seq_lifted_gd<float>(const TensSeqParams&, int, float*, const float*, cudaStream_t);
template int launch_tens_seq_lifted_gd<double>(const TensSeqParams&, int, double*, const double*, cudaStream_t);
template int launch_tens_seq_lifted<float>(const TensSeqParams&, int, void*, size_t, cudaStream_t);
template int launch_tens_seq_lifted<double>(const TensSeqParams&, int, void*, size_t, cudaStream_t);

}  // namespace gpsig
