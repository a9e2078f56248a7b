/* gpsig_b200 -- C ABI of the B200-native signature-kernel Gram library.
 *
 * This header is the drop-in boundary for the hot path of maudl3116/GPSig:
 * everything below `gpsig.kernels` / `gpsig.kernels_pde` that touches numbers.
 * Plain C: raw device pointers, sizes, a cudaStream_t passed as void*.  No
 * torch / TensorFlow / C++ types cross it.  All functions
 *   - return 0 (GPSIG_OK) or a negative GPSIG_ERR_* code; the message is
 *     available (thread-local) from gpsig_last_error_string();
 *   - launch asynchronously on the caller's stream and never synchronise;
 *   - never allocate, free or retain device memory: the caller owns inputs,
 *     outputs and the workspace (size from the matching *_workspace_bytes);
 *   - treat every input as const.
 * Sequences are dense row-major [N, L, d] (== the reference's 2-D GPflow
 * layout [N, L*d], gpsig/kernels.py:418-420).  `lengthscales` is a DEVICE
 * pointer to d values or NULL (gpsig/kernels.py:358-359: X /= lengthscales).
 *
 * Reference interface each entry point replaces (paths relative to the
 * reference repository root):
 *   gpsig_trunc_gram_*   gpsig/kernels.py:209-238  SignatureKernel._K_seq
 *                        + gpsig/signature_algs.py:8-35 / :37-74, with the
 *                        normalisation / variance epilogue of kernels.py:431-477
 *   gpsig_trunc_diag_*   gpsig/kernels.py:190-207  SignatureKernel._K_seq_diag
 *   gpsig_tens_vs_seq_*  gpsig/kernels.py:314-341  _K_tens_vs_seq
 *                        + gpsig/signature_algs.py:101-127 / :129-160, epilogue kernels.py:604-620
 *   gpsig_tens_tens_*    gpsig/kernels.py:264-284  _K_tens + signature_algs.py:76-99
 *   gpsig_pde_*          gpsig/sigKer_fast.pyx:15-62 sig_kern_diag and the TF op
 *                        gpsig/covariance_op/untrunc_cov_op_gpu.cc:43-50,61-102
 *                        (UntruncCovKernelLauncher / UntruncCovOp::Compute)
 */
#ifndef GPSIG_B200_H_
#define GPSIG_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GPSIG_VERSION 100

/* status codes */
#define GPSIG_OK 0
#define GPSIG_ERR_INVALID_ARGUMENT (-1)
#define GPSIG_ERR_UNSUPPORTED (-2)
#define GPSIG_ERR_WORKSPACE (-3)
#define GPSIG_ERR_CUDA (-4)

/* static (base) kernels on R^d, gpsig/kernels.py:966-1173 */
#define GPSIG_BASE_LINEAR 0   /* _lin      kernels.py:979  */
#define GPSIG_BASE_RBF 1      /* _rbf      kernels.py:1042 */
#define GPSIG_BASE_COSINE 2   /* _cos      kernels.py:1000 */
#define GPSIG_BASE_POLY 3     /* _poly     kernels.py:1024  params: gamma, degree */
#define GPSIG_BASE_MIX 4      /* _mix      kernels.py:1061  params: mixing */
#define GPSIG_BASE_MATERN12 5 /* _Matern12 kernels.py:1135 */
#define GPSIG_BASE_MATERN32 6 /* _Matern32 kernels.py:1154 */
#define GPSIG_BASE_MATERN52 7 /* _Matern52 kernels.py:1171 */
#define GPSIG_BASE_SPECTRAL 8 /* _spectral kernels.py:1101-1122: sum_q alpha_q exp(-|gamma_q*(x-y)|^2 / 2) cos(2 pi omega_q.(x-y))
                                 ('rbf' family) or exp(-|gamma_q*(x-y)| / 2) ('exp' family); forward only, general path */
#define GPSIG_SPECTRAL_RBF 0
#define GPSIG_SPECTRAL_EXP 1

/* What the truncated kernel computes (constructor state of
 * gpsig/kernels.py:19-89 that reaches the numerics). */
typedef struct gpsig_trunc_desc {
  int base;              /* GPSIG_BASE_* */
  int num_levels;        /* m >= 1 (kernels.py:56) */
  int order;             /* 1..m (kernels.py:58); 1 = first-order recursion */
  int difference;        /* kernels.py:64; 1 = difference the lifted Gram */
  double base_params[4]; /* POLY: {gamma, degree}; MIX: {mixing} */
  /* SPECTRAL only: device array of DOUBLES (whatever the data type) alpha[Q], omega[Q, d], gamma[Q, d]
   * (kernels.py:1092-1094), Q = spectral_q in 1..64, spectral_family = GPSIG_SPECTRAL_*. */
  const double* spectral;
  int spectral_q;
  int spectral_family;
} gpsig_trunc_desc;

/* How raw levels K_l become the returned matrix (kernels.py:431-434,456-477):
 *   v_l = K_l (+ jitter if X2 == NULL and i == j)
 *   v_l /= sqrt(norm_x[l, i]) * sqrt(norm_x2[l, j])      if norm_* != NULL
 *   v_l *= level_scale[l]                                 if level_scale != NULL
 *   out  = v_l per level (return_levels) or sum_l v_l
 * norm_* hold the (jittered) per-level diagonals [m+1, N] / [m+1, N2] computed
 * by gpsig_trunc_diag_*; level_scale = sigma * variances [m+1].  Device ptrs. */
typedef struct gpsig_epilogue {
  const void* norm_x;
  const void* norm_x2;
  const void* level_scale;
  double sym_jitter;
  int return_levels;
  /* K(X, X) only, row-sharded callers: entries with i > j MAY be left unwritten (the fused kernel skips the CTAs
   * that lie strictly below the diagonal); after assembling all rows complete the matrix with
   * gpsig_mirror_lower_*.  A single call that covers all rows of K(X, X) does both by itself. */
  int upper_only;
} gpsig_epilogue;

const char* gpsig_last_error_string(void);
int gpsig_version(void);

/* ---- truncated kernel: sequences vs sequences ------------------------------
 * out: rows [row_begin, row_end) of the [N, N2] Gram, i.e. device array
 *      [(m+1,)? row_end-row_begin, N2] (row-block sharding across GPUs).
 * X2 == NULL: K(X, X) (then N2 = N, L2 = L are ignored). */
size_t gpsig_trunc_gram_workspace_bytes(int N, int L, int N2, int L2, int d, int elem_size);
int gpsig_trunc_gram_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L,
                         const double* X2, int N2, int L2, int d, const double* lengthscales,
                         int row_begin, int row_end, const gpsig_epilogue* epi, double* out,
                         void* workspace, size_t workspace_bytes, void* stream);
int gpsig_trunc_gram_f32(const gpsig_trunc_desc* desc, const float* X, int N, int L,
                         const float* X2, int N2, int L2, int d, const float* lengthscales,
                         int row_begin, int row_end, const gpsig_epilogue* epi, float* out,
                         void* workspace, size_t workspace_bytes, void* stream);

/* out[p, i, j] = out[p, j, i] for i > j on `nplanes` [N, N] planes (see gpsig_epilogue.upper_only). */
int gpsig_mirror_lower_f64(double* out, int N, int nplanes, void* stream);
int gpsig_mirror_lower_f32(float* out, int N, int nplanes, void* stream);

/* ---- truncated kernel: diagonal k_l(x_i, x_i), raw levels out [m+1, N] ---- */
size_t gpsig_trunc_diag_workspace_bytes(int N, int L, int d, int elem_size);
int gpsig_trunc_diag_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L, int d,
                         const double* lengthscales, double* out,
                         void* workspace, size_t workspace_bytes, void* stream);
int gpsig_trunc_diag_f32(const gpsig_trunc_desc* desc, const float* X, int N, int L, int d,
                         const float* lengthscales, float* out,
                         void* workspace, size_t workspace_bytes, void* stream);

/* ---- diagonal without any allocation (gpsig_b200/csrc/static_scratch.cu) ----
 * No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same result as gpsig_trunc_diag_f64 (raw levels out [m+1, N]); the packed-sequence scratch lives
 * in a file-scope `static __device__` array of GPSIG_STATIC_SCRATCH_BYTES that every GPU loading the
 * library holds until exit.  The caller passes inputs and output only; the call allocates nothing.
 * Shapes with gpsig_trunc_diag_workspace_bytes(N, L, d, 8) > GPSIG_STATIC_SCRATCH_BYTES are refused
 * (GPSIG_ERR_WORKSPACE).  All calls on a device share the array: calls that overlap on the GPU
 * (different streams without an ordering event) race -- serialise them. */
#define GPSIG_STATIC_SCRATCH_BYTES (16u << 20)   /* 16 MiB, e.g. N=2000, L=100, d=8 (14.4 MB) */
int gpsig_trunc_diag_static_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L, int d,
                                const double* lengthscales, double* out, void* stream);

/* ---- inducing tensors vs sequences (Kzx) -----------------------------------
 * Z: [T, Mz, d] or, with increments, [T, Mz, 2, d]; T = m(m+1)/2.
 * out: columns [col_begin, col_end) of [(m+1,)? Mz, N]  (sharded over N).
 * epi->norm_x2 (if set) is the per-level diagonal of X [m+1, N]; norm_x unused. */
size_t gpsig_tens_vs_seq_workspace_bytes(int Mz, int increments, int N, int L, int d, int num_levels, int elem_size);
int gpsig_tens_vs_seq_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int increments,
                          const double* X, int N, int L, int d, const double* lengthscales,
                          int col_begin, int col_end, const gpsig_epilogue* epi, double* out,
                          void* workspace, size_t workspace_bytes, void* stream);
int gpsig_tens_vs_seq_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int increments,
                          const float* X, int N, int L, int d, const float* lengthscales,
                          int col_begin, int col_end, const gpsig_epilogue* epi, float* out,
                          void* workspace, size_t workspace_bytes, void* stream);

/* ---- inducing tensors vs inducing tensors (Kzz), out [(m+1,)? Mz, Mz] ------ */
int gpsig_tens_tens_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int increments, int d,
                        const double* lengthscales, const gpsig_epilogue* epi, double* out, void* stream);
int gpsig_tens_tens_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int increments, int d,
                        const float* lengthscales, const gpsig_epilogue* epi, float* out, void* stream);

/* ---- reverse mode (vector-Jacobian products) of the first-order truncated kernels -------------
 * The reference differentiates gpsig/signature_algs.py:8-35, 76-127 with TensorFlow autodiff (training.py
 * optimises lengthscales, variances and the inducing tensors through Kzz / Kzx / Kxx_diag,
 * inducing_variables.py:51-137).  These entry points are the hand-written adjoints of the RAW levels
 * (no normalisation, no level scaling: the host layer chains those).  Inputs are the dense coordinates the
 * forward kernels were called with, ALREADY divided by the lengthscales; gradients are w.r.t. those.
 *   grad_levels: dL/dK_l, same shape as the forward's return_levels output (level 0 is ignored).
 *   gradX / gradX2 / gradZ: same shapes as X / X2 / Z.  gradX, gradX2 are ACCUMULATED into (zero them first);
 *   X2 == NULL: symmetric case, both roles accumulate into gradX.  gpsig_tens_tens_vjp overwrites gradZ.
 * pair_mode: GPSIG_PDE_GRAM = rows [row_begin,row_end) x all N2 columns; GPSIG_PDE_DIAG = pairs (i, i).
 * num_levels <= 8.  order > 1 (signature_algs.py:37-74, :129-160): gpsig_trunc_vjp_* and gpsig_tens_vs_seq_vjp_* run the
 * general adjoint kernels (generic_vjp_kernels.cuh); gpsig_tens_tens_* does not depend on the order. */
size_t gpsig_trunc_vjp_workspace_bytes(int npairs, int L, int L2, int num_levels, int difference, int elem_size);
/* the same for any `order` (order > 1: the higher-order recursion of signature_algs.py:37-74 stores its per-level factor
 * tiles; order == 1 equals the function above) */
size_t gpsig_trunc_vjp_workspace_bytes_order(int npairs, int L, int L2, int num_levels, int order, int difference, int elem_size);
int gpsig_trunc_vjp_f64(const gpsig_trunc_desc* desc, const double* X, int N, int L, const double* X2, int N2, int L2,
                        int d, int pair_mode, int row_begin, int row_end, const double* grad_levels,
                        double* gradX, double* gradX2, void* workspace, size_t workspace_bytes, void* stream);
int gpsig_trunc_vjp_f32(const gpsig_trunc_desc* desc, const float* X, int N, int L, const float* X2, int N2, int L2,
                        int d, int pair_mode, int row_begin, int row_end, const float* grad_levels,
                        float* gradX, float* gradX2, void* workspace, size_t workspace_bytes, void* stream);
size_t gpsig_tens_vs_seq_vjp_workspace_bytes(int Mz, int ncols, int L, int num_levels, int elem_size);
int gpsig_tens_vs_seq_vjp_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int increments, const double* X,
                              int N, int L, int d, int col_begin, int col_end, const double* grad_levels,
                              double* gradZ, double* gradX, void* workspace, size_t workspace_bytes, void* stream);
int gpsig_tens_vs_seq_vjp_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int increments, const float* X,
                              int N, int L, int d, int col_begin, int col_end, const float* grad_levels,
                              float* gradZ, float* gradX, void* workspace, size_t workspace_bytes, void* stream);
/* Reverse mode of gpsig_tens_vs_seq_* at state dimensions above 8, lifted form (csrc/tens_lifted_kernels.cuh; the reference gets
 * this gradient from TensorFlow autodiff of gpsig/signature_algs.py:101-127 + kernels.py:322-341).  First-order linear / RBF.
 * Gd [m(m+1)/2 * zs * Mz, L * C], C = col_end - col_begin: row r = (c zs + s) Mz + z (component c, increment half s, tensor z),
 * column e = l C + n; entry = dL/dkappa(z point r, x point e), times kappa for the RBF kernel.  The caller finishes with two
 * plain GEMMs: dL/dZ = Gd X - rowsum(Gd) Z, dL/dX = Gd^T Z - colsum(Gd) X (linear: without the sums).  Z, X as for
 * gpsig_tens_vs_seq_vjp_* (already scaled by the lengthscales); grad_levels [m+1, Mz, C] w.r.t. the raw levels. */
size_t gpsig_tens_vs_seq_lifted_gd_workspace_bytes(int Mz, int increments, int N, int L, int d, int num_levels, int elem);
int gpsig_tens_vs_seq_lifted_gd_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int increments, const double* X,
                                    int N, int L, int d, int col_begin, int col_end, const double* grad_levels,
                                    double* Gd, void* workspace, size_t workspace_bytes, void* stream);
int gpsig_tens_vs_seq_lifted_gd_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int increments, const float* X,
                                    int N, int L, int d, int col_begin, int col_end, const float* grad_levels,
                                    float* Gd, void* workspace, size_t workspace_bytes, void* stream);
int gpsig_tens_tens_vjp_f64(const gpsig_trunc_desc* desc, const double* Z, int Mz, int increments, int d,
                            const double* grad_levels, double* gradZ, void* stream);
int gpsig_tens_tens_vjp_f32(const gpsig_trunc_desc* desc, const float* Z, int Mz, int increments, int d,
                            const float* grad_levels, float* gradZ, void* stream);

/* ---- untruncated (Goursat PDE) kernel ---------------------------------------
 * Explicit 2nd-order stencil of sigKer_fast.pyx:48 on the 2^dyadic_order
 * refined grid, linear base kernel (what the reference's Cython path solves).
 * pair_mode GPSIG_PDE_DIAG: out[a] = sigma * k(x_a, y_a), a in [row_begin,row_end)
 *           (Y == NULL -> y = x: the reference's Kdiag, kernels_pde.py:160-185);
 * pair_mode GPSIG_PDE_GRAM: out[i - row_begin, j] = sigma * k(x_i, y_j). */
#define GPSIG_PDE_DIAG 0
#define GPSIG_PDE_GRAM 1
size_t gpsig_pde_workspace_bytes(int N, int L, int N2, int L2, int d, int dyadic_order, int pair_mode);
int gpsig_pde_f64(const double* X, int N, int L, const double* Y, int N2, int L2, int d,
                  const double* lengthscales, int dyadic_order, int pair_mode,
                  int row_begin, int row_end, double sigma, double* out,
                  void* workspace, size_t workspace_bytes, void* stream);

/* ---- reverse mode of the PDE kernel diagonal ----------------------------------
 * The reference's hand-written gradient of Kdiag w.r.t. the sequences: `_KdiagGrad`
 * (gpsig/kernels_pde.py:465-509, Cython path) / `_untrunc_cov_grad`
 * (gpsig/covariance_op/_untrunc_cov_grad.py:25-77, TF op): the continuous adjoint
 * sum_q K[p,q] K_rev[P-1-p,P-1-q] dx_q, NOT the derivative of the stencil.
 *   X          [N, L, d] coordinates ALREADY divided by the lengthscales; gradients are w.r.t. those
 *   rev_solver stencil of the reversed-path solve: 0 = 1st order (sigKer_fast.pyx:50,61, implementation
 *              'cython'), 1 = 2nd order (the TF op re-runs `untrunc_cov`, implementation 'gpu_op')
 *   grad_out   [row_end - row_begin] dL/dKdiag (fold sigma in)
 *   gradX      [N, L, d]; rows [row_begin, row_end) are OVERWRITTEN
 *   workspace  any multiple of gpsig_pde_kdiag_vjp_workspace_bytes(1, L, d, order) (8 P^2 bytes of grid scratch
 *              per sequence in flight); the call runs the rows in chunks of what fits. */
size_t gpsig_pde_kdiag_vjp_workspace_bytes(int nseq, int L, int d, int dyadic_order);
int gpsig_pde_kdiag_vjp_f64(const double* X, int N, int L, int d, int dyadic_order, int rev_solver,
                            int row_begin, int row_end, const double* grad_out, double* gradX,
                            void* workspace, size_t workspace_bytes, void* stream);

/* ---- diagnostics (no reference counterpart; used by bench.py) ----------------
 * gpsig_launch_count: kernels this library has launched since it was loaded.
 * gpsig_profile_enable(1): from now on (this host thread) the launcher of each
 *   MAIN kernel (fused Gram / tensors-vs-sequences / PDE wavefront) brackets it
 *   with CUDA events on the caller's stream; gpsig_profile_last_kernel_ms()
 *   waits for the most recent one and returns its duration (-1 if none).
 * gpsig_fma_peak: FMA-chain microbenchmark (`blocks` CTAs x 256 threads x 8
 *   chains x `iters` FMAs) -- the measured CUDA-core FP32/FP64 peak the
 *   compute roofline is quoted against. */
unsigned long long gpsig_launch_count(void);
void gpsig_profile_enable(int on);
double gpsig_profile_last_kernel_ms(void);
int gpsig_fma_peak(int elem_size, int blocks, int iters, void* scratch, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* GPSIG_B200_H_ */
