// gpsig_trunc_diag_static_f64: the diagonal kernel for callers that cannot allocate.
//
// NO COUNTERPART IN THE REFERENCE -- added for callers that
// cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid
// it).  Same computation as gpsig_trunc_diag_f64 (raw per-level diagonals k_l(x_i, x_i), [m+1, N]);
// the only difference is where the packed-sequence scratch lives: in the file-scope device array
// below instead of a caller-provided workspace.
//
// * The array costs GPSIG_STATIC_SCRATCH_BYTES of device memory on EVERY GPU that loads the library,
//   from module load until the process exits.
// * In HOST code the name `g_static_scratch` is a host-side stand-in, not a device address: the
//   device address is obtained here, once per device and with that device current, with
//   cudaGetSymbolAddress, and only that pointer is handed to the kernels.
// * Every call on a device shares the array and finds the previous call's contents in it.  Nothing
//   in it has to start at zero: the pack kernel overwrites every element the diagonal kernel reads
//   (coordinates and squared norms of the N sequences); the alignment padding it leaves untouched is
//   copied around but never used in arithmetic.
// * Calls that OVERLAP on the GPU (different streams, or graph branches that may run concurrently)
//   race on the array.  Serialise them: one stream, or an event between them.
// * The entry point calls no cudaMalloc / cudaFree / allocating library routine (and no torch);
//   shapes whose scratch exceeds the array are refused with GPSIG_ERR_WORKSPACE.
#include <mutex>

#include "common.cuh"

namespace gpsig {

static __device__ __align__(256) unsigned char g_static_scratch[GPSIG_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
static void* g_scratch_addr[kMaxDevices] = {};   // device address of g_static_scratch, per device ordinal
static std::mutex g_scratch_mu;

// Device address of the array on the CURRENT device (looked up on first use per device).
static int static_scratch_ptr(void** out) {
  int dev = -1;
  GPSIG_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) GPSIG_FAIL(GPSIG_ERR_UNSUPPORTED, "device ordinal %d outside 0..%d", dev, kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(g_scratch_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    GPSIG_CUDA_CHECK(cudaGetSymbolAddress(&p, g_static_scratch));
    if (!p) GPSIG_FAIL(GPSIG_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch array");
    g_scratch_addr[dev] = p;
  }
  *out = g_scratch_addr[dev];
  return GPSIG_OK;
}

}  // namespace gpsig

using namespace gpsig;

extern "C" int gpsig_trunc_diag_static_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L, int d,
                                            const double* lengthscales, double* out, void* stream) {
  if (N < 0 || L < 1 || d < 1) GPSIG_FAIL(GPSIG_ERR_INVALID_ARGUMENT, "bad sizes N=%d L=%d d=%d", N, L, d);
  const size_t need = gpsig_trunc_diag_workspace_bytes(N, L, d, 8);
  if (need > (size_t)GPSIG_STATIC_SCRATCH_BYTES)
    GPSIG_FAIL(GPSIG_ERR_WORKSPACE,
               "gpsig_trunc_diag_static_f64: N=%d, L=%d, d=%d needs %zu bytes of scratch, the static array holds %zu; "
               "use gpsig_trunc_diag_f64 with a caller-provided workspace",
               N, L, d, need, (size_t)GPSIG_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int rc = static_scratch_ptr(&scratch)) return rc;
  // same validation, pack kernel and diagonal kernel as the workspace variant; `scratch` is a device address
  return gpsig_trunc_diag_f64(desc, X, N, L, d, lengthscales, out, scratch, (size_t)GPSIG_STATIC_SCRATCH_BYTES, stream);
}
