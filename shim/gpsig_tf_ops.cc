// TensorFlow custom-op shim over the C ABI of libgpsig_b200.so (include/gpsig_b200.h).
//
// OPTIONAL: built only where TensorFlow headers exist (shim/build_shim.py); nothing in gpsig_b200 depends on it.
// It follows the reference's own native-op pattern file by file:
//   REGISTER_OP + shape function      gpsig/covariance_op/untrunc_cov_op_gpu.cc:13-40
//   OpKernel::Compute                 gpsig/covariance_op/untrunc_cov_op_gpu.cc:61-102
//   REGISTER_KERNEL_BUILDER (GPU)     gpsig/covariance_op/untrunc_cov_op_gpu.cc:148-149
//   gradient registered in Python     gpsig/covariance_op/_untrunc_cov_grad.py:25  (shim/gpsig_tf_grad.py)
// with three differences the reference's op gets wrong: outputs and scratch are allocated through the
// OpKernelContext instead of aliasing an input (cc:70,85), kernels run on TensorFlow's stream instead of the legacy
// default stream (cu:80), and every call's status is checked.
//
// Ops (T in {float, double}; sequences [N, L, d] already reshaped as kernels.py:419 does):
//   GpsigTruncGram     x, x2, lengthscales -> raw levels [m+1, N, N2]       replaces _K_seq + signature_kern_first_order /
//                                                                            higher_order (kernels.py:209-238, signature_algs.py:8-74)
//   GpsigTruncGramGrad x, x2, grad_levels  -> grad_x, grad_x2               (TF autodiff of the same in the reference)
//   GpsigTensVsSeq     z, x, lengthscales  -> raw levels [m+1, Mz, N]       replaces _K_tens_vs_seq (kernels.py:314-341)
//   GpsigPdeDiag       x, lengthscales     -> kdiag [N]                      replaces UntruncCov + sol[:, -1, -1] (kernels_pde.py:176-185)
//   GpsigPdeDiagGrad   x, grad_kdiag       -> grad_x                         replaces _untrunc_cov_grad.py:25-77 / _KdiagGrad
#include <cuda_runtime.h>
#include <string.h>

#include "tensorflow/core/framework/op.h"
#include "tensorflow/core/framework/op_kernel.h"
#include "tensorflow/core/framework/shape_inference.h"

#include "gpsig_b200.h"

using namespace tensorflow;

namespace {

// TensorFlow's CUDA stream of this op (the reference launches on stream 0, untrunc_cov_op_gpu.cu:80)
void* TfStream(OpKernelContext* ctx) { return (void*)ctx->eigen_device<Eigen::GpuDevice>().stream(); }

Status Check(int rc) { return rc == GPSIG_OK ? Status::OK() : errors::Internal("gpsig_b200: ", gpsig_last_error_string()); }

struct TruncAttrs {
  gpsig_trunc_desc desc;
  explicit TruncAttrs(OpKernelConstruction* c) {
    memset(&desc, 0, sizeof(desc));
    float p0 = 0.f, p1 = 0.f;
    bool diff = true;
    OP_REQUIRES_OK(c, c->GetAttr("base", &desc.base));
    OP_REQUIRES_OK(c, c->GetAttr("num_levels", &desc.num_levels));
    OP_REQUIRES_OK(c, c->GetAttr("order", &desc.order));
    OP_REQUIRES_OK(c, c->GetAttr("difference", &diff));
    OP_REQUIRES_OK(c, c->GetAttr("base_p0", &p0));
    OP_REQUIRES_OK(c, c->GetAttr("base_p1", &p1));
    desc.difference = diff ? 1 : 0;
    desc.base_params[0] = p0; desc.base_params[1] = p1;
  }
};

template <typename T> struct Abi;
template <> struct Abi<double> {
  static constexpr auto gram = gpsig_trunc_gram_f64;
  static constexpr auto gram_vjp = gpsig_trunc_vjp_f64;
  static constexpr auto tens_vs_seq = gpsig_tens_vs_seq_f64;
};
template <> struct Abi<float> {
  static constexpr auto gram = gpsig_trunc_gram_f32;
  static constexpr auto gram_vjp = gpsig_trunc_vjp_f32;
  static constexpr auto tens_vs_seq = gpsig_tens_vs_seq_f32;
};

#define GPSIG_TRUNC_ATTRS \
  .Attr("T: {float, double}").Attr("base: int = 0").Attr("num_levels: int").Attr("order: int = 1") \
  .Attr("difference: bool = true").Attr("base_p0: float = 0.0").Attr("base_p1: float = 0.0")

}  // namespace

// ---------------------------------------------------------------------------------------------------------------
REGISTER_OP("GpsigTruncGram")
    .Input("x: T").Input("x2: T").Input("lengthscales: T").Output("levels: T") GPSIG_TRUNC_ATTRS
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      shape_inference::ShapeHandle x, x2;
      TF_RETURN_IF_ERROR(c->WithRank(c->input(0), 3, &x));           // like cc:22: rank checks only
      TF_RETURN_IF_ERROR(c->WithRank(c->input(1), 3, &x2));
      int m = 0;
      TF_RETURN_IF_ERROR(c->GetAttr("num_levels", &m));
      c->set_output(0, c->MakeShape({c->MakeDim(m + 1), c->Dim(x, 0), c->Dim(x2, 0)}));
      return Status::OK();
    });

template <typename T>
class GpsigTruncGramOp : public OpKernel {
 public:
  explicit GpsigTruncGramOp(OpKernelConstruction* c) : OpKernel(c), a_(c) {}
  void Compute(OpKernelContext* ctx) override {
    const Tensor& x = ctx->input(0);
    const Tensor& x2 = ctx->input(1);
    OP_REQUIRES(ctx, x.dims() == 3 && x2.dims() == 3 && x.dim_size(2) == x2.dim_size(2),
                errors::InvalidArgument("x, x2 must be [N, L, d] with the same d"));
    const int N = x.dim_size(0), L = x.dim_size(1), d = x.dim_size(2), N2 = x2.dim_size(0), L2 = x2.dim_size(1);
    OP_REQUIRES(ctx, ctx->input(2).NumElements() == d, errors::InvalidArgument("lengthscales must have d entries"));
    Tensor* out = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({a_.desc.num_levels + 1, N, N2}), &out));
    const size_t wsb = gpsig_trunc_gram_workspace_bytes(N, L, N2, L2, d, sizeof(T));
    Tensor ws;
    OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, TensorShape({(int64)wsb + 16}), &ws));
    OP_REQUIRES_OK(ctx, Check(Abi<T>::gram(&a_.desc, x.flat<T>().data(), N, L, x2.flat<T>().data(), N2, L2, d,
                                           ctx->input(2).flat<T>().data(), 0, N, nullptr /* raw levels */,
                                           out->flat<T>().data(), ws.flat<uint8>().data(), wsb, TfStream(ctx))));
  }
 private:
  TruncAttrs a_;
};
REGISTER_KERNEL_BUILDER(Name("GpsigTruncGram").Device(DEVICE_GPU).TypeConstraint<double>("T"), GpsigTruncGramOp<double>);
REGISTER_KERNEL_BUILDER(Name("GpsigTruncGram").Device(DEVICE_GPU).TypeConstraint<float>("T"), GpsigTruncGramOp<float>);

// x, x2: the coordinates ALREADY divided by the lengthscales (the Python gradient chains the division, gpsig_tf_grad.py)
REGISTER_OP("GpsigTruncGramGrad")
    .Input("x: T").Input("x2: T").Input("grad_levels: T").Output("grad_x: T").Output("grad_x2: T") GPSIG_TRUNC_ATTRS
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      c->set_output(0, c->input(0));
      c->set_output(1, c->input(1));
      return Status::OK();
    });

template <typename T>
class GpsigTruncGramGradOp : public OpKernel {
 public:
  explicit GpsigTruncGramGradOp(OpKernelConstruction* c) : OpKernel(c), a_(c) {}
  void Compute(OpKernelContext* ctx) override {
    const Tensor& x = ctx->input(0);
    const Tensor& x2 = ctx->input(1);
    const Tensor& g = ctx->input(2);
    const int N = x.dim_size(0), L = x.dim_size(1), d = x.dim_size(2), N2 = x2.dim_size(0), L2 = x2.dim_size(1);
    OP_REQUIRES(ctx, g.dims() == 3 && g.dim_size(0) == a_.desc.num_levels + 1 && g.dim_size(1) == N && g.dim_size(2) == N2,
                errors::InvalidArgument("grad_levels must be [m+1, N, N2]"));
    Tensor *gx = nullptr, *gx2 = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, x.shape(), &gx));
    OP_REQUIRES_OK(ctx, ctx->allocate_output(1, x2.shape(), &gx2));
    // the adjoint ACCUMULATES into its outputs (include/gpsig_b200.h): zero them on the op's stream first
    cudaStream_t st = (cudaStream_t)TfStream(ctx);
    OP_REQUIRES(ctx, cudaMemsetAsync(gx->flat<T>().data(), 0, sizeof(T) * (size_t)x.NumElements(), st) == cudaSuccess &&
                         cudaMemsetAsync(gx2->flat<T>().data(), 0, sizeof(T) * (size_t)x2.NumElements(), st) == cudaSuccess,
                errors::Internal("cudaMemsetAsync failed"));
    const size_t wsb = gpsig_trunc_vjp_workspace_bytes(N * N2, L, L2, a_.desc.num_levels, a_.desc.difference, sizeof(T));
    Tensor ws;
    OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, TensorShape({(int64)wsb + 16}), &ws));
    OP_REQUIRES_OK(ctx, Check(Abi<T>::gram_vjp(&a_.desc, x.flat<T>().data(), N, L, x2.flat<T>().data(), N2, L2, d,
                                               GPSIG_PDE_GRAM, 0, N, g.flat<T>().data(), gx->flat<T>().data(),
                                               gx2->flat<T>().data(), ws.flat<uint8>().data(), wsb, TfStream(ctx))));
  }
 private:
  TruncAttrs a_;
};
REGISTER_KERNEL_BUILDER(Name("GpsigTruncGramGrad").Device(DEVICE_GPU).TypeConstraint<double>("T"), GpsigTruncGramGradOp<double>);
REGISTER_KERNEL_BUILDER(Name("GpsigTruncGramGrad").Device(DEVICE_GPU).TypeConstraint<float>("T"), GpsigTruncGramGradOp<float>);

// ---------------------------------------------------------------------------------------------------------------
REGISTER_OP("GpsigTensVsSeq")
    .Input("z: T").Input("x: T").Input("lengthscales: T").Output("levels: T") GPSIG_TRUNC_ATTRS
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      shape_inference::ShapeHandle x;
      TF_RETURN_IF_ERROR(c->WithRank(c->input(1), 3, &x));
      int m = 0;
      TF_RETURN_IF_ERROR(c->GetAttr("num_levels", &m));
      c->set_output(0, c->MakeShape({c->MakeDim(m + 1), c->Dim(c->input(0), 1), c->Dim(x, 0)}));
      return Status::OK();
    });

template <typename T>
class GpsigTensVsSeqOp : public OpKernel {
 public:
  explicit GpsigTensVsSeqOp(OpKernelConstruction* c) : OpKernel(c), a_(c) {}
  void Compute(OpKernelContext* ctx) override {
    const Tensor& z = ctx->input(0);                                  // [T, Mz, d] or [T, Mz, 2, d] (increments, kernels.py:329-331)
    const Tensor& x = ctx->input(1);
    OP_REQUIRES(ctx, (z.dims() == 3 || z.dims() == 4) && x.dims() == 3, errors::InvalidArgument("z must be rank 3 or 4, x rank 3"));
    const int increments = z.dims() == 4 ? 1 : 0;
    const int Mz = z.dim_size(1), N = x.dim_size(0), L = x.dim_size(1), d = x.dim_size(2), m = a_.desc.num_levels;
    OP_REQUIRES(ctx, z.dim_size(0) == m * (m + 1) / 2 && z.dim_size(z.dims() - 1) == d,
                errors::InvalidArgument("z must hold m(m+1)/2 components of dimension d"));
    Tensor* out = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({m + 1, Mz, N}), &out));
    const size_t wsb = gpsig_tens_vs_seq_workspace_bytes(Mz, increments, N, L, d, m, sizeof(T));
    Tensor ws;
    OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, TensorShape({(int64)wsb + 16}), &ws));
    OP_REQUIRES_OK(ctx, Check(Abi<T>::tens_vs_seq(&a_.desc, z.flat<T>().data(), Mz, increments, x.flat<T>().data(), N, L, d,
                                                  ctx->input(2).flat<T>().data(), 0, N, nullptr, out->flat<T>().data(),
                                                  ws.flat<uint8>().data(), wsb, TfStream(ctx))));
  }
 private:
  TruncAttrs a_;
};
REGISTER_KERNEL_BUILDER(Name("GpsigTensVsSeq").Device(DEVICE_GPU).TypeConstraint<double>("T"), GpsigTensVsSeqOp<double>);
REGISTER_KERNEL_BUILDER(Name("GpsigTensVsSeq").Device(DEVICE_GPU).TypeConstraint<float>("T"), GpsigTensVsSeqOp<float>);

// ---------------------------------------------------------------------------------------------------------------
// float64 only, like the reference op (`Input("paths: double")`, cc:14)
REGISTER_OP("GpsigPdeDiag")
    .Input("paths: double").Input("lengthscales: double").Output("kdiag: double").Attr("order: int = 0")
    .SetShapeFn([](shape_inference::InferenceContext* c) {
      shape_inference::ShapeHandle x;
      TF_RETURN_IF_ERROR(c->WithRank(c->input(0), 3, &x));
      c->set_output(0, c->MakeShape({c->Dim(x, 0)}));
      return Status::OK();
    });

class GpsigPdeDiagOp : public OpKernel {
 public:
  explicit GpsigPdeDiagOp(OpKernelConstruction* c) : OpKernel(c) { OP_REQUIRES_OK(c, c->GetAttr("order", &order_)); }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& x = ctx->input(0);
    OP_REQUIRES(ctx, x.dims() == 3, errors::InvalidArgument("paths must be [N, L, d]"));
    const int N = x.dim_size(0), L = x.dim_size(1), d = x.dim_size(2);
    Tensor* out = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, TensorShape({N}), &out));
    const size_t wsb = gpsig_pde_workspace_bytes(N, L, 0, 0, d, order_, GPSIG_PDE_DIAG);
    Tensor ws;
    OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, TensorShape({(int64)wsb + 16}), &ws));
    OP_REQUIRES_OK(ctx, Check(gpsig_pde_f64(x.flat<double>().data(), N, L, nullptr, 0, 0, d, ctx->input(1).flat<double>().data(),
                                            order_, GPSIG_PDE_DIAG, 0, N, 1.0, out->flat<double>().data(),
                                            ws.flat<uint8>().data(), wsb, TfStream(ctx))));
  }
 private:
  int order_ = 0;
};
REGISTER_KERNEL_BUILDER(Name("GpsigPdeDiag").Device(DEVICE_GPU), GpsigPdeDiagOp);

// paths: ALREADY divided by the lengthscales; rev_solver 0 = the Cython path's 1st-order reversed solve, 1 = the TF op's
REGISTER_OP("GpsigPdeDiagGrad")
    .Input("paths: double").Input("grad_kdiag: double").Output("grad_paths: double").Attr("order: int = 0").Attr("rev_solver: int = 1")
    .SetShapeFn([](shape_inference::InferenceContext* c) { c->set_output(0, c->input(0)); return Status::OK(); });

class GpsigPdeDiagGradOp : public OpKernel {
 public:
  explicit GpsigPdeDiagGradOp(OpKernelConstruction* c) : OpKernel(c) {
    OP_REQUIRES_OK(c, c->GetAttr("order", &order_));
    OP_REQUIRES_OK(c, c->GetAttr("rev_solver", &rev_));
  }
  void Compute(OpKernelContext* ctx) override {
    const Tensor& x = ctx->input(0);
    const Tensor& g = ctx->input(1);
    OP_REQUIRES(ctx, x.dims() == 3 && g.NumElements() == x.dim_size(0), errors::InvalidArgument("paths [N, L, d], grad_kdiag [N]"));
    const int N = x.dim_size(0), L = x.dim_size(1), d = x.dim_size(2);
    Tensor* gx = nullptr;
    OP_REQUIRES_OK(ctx, ctx->allocate_output(0, x.shape(), &gx));
    // scratch for up to 64 sequences in flight (the call chunks the rows to what it is given)
    const size_t wsb = gpsig_pde_kdiag_vjp_workspace_bytes(N < 64 ? N : 64, L, d, order_);
    Tensor ws;
    OP_REQUIRES_OK(ctx, ctx->allocate_temp(DT_UINT8, TensorShape({(int64)wsb + 16}), &ws));
    OP_REQUIRES_OK(ctx, Check(gpsig_pde_kdiag_vjp_f64(x.flat<double>().data(), N, L, d, order_, rev_, 0, N, g.flat<double>().data(),
                                                      gx->flat<double>().data(), ws.flat<uint8>().data(), wsb, TfStream(ctx))));
  }
 private:
  int order_ = 0, rev_ = 1;
};
REGISTER_KERNEL_BUILDER(Name("GpsigPdeDiagGrad").Device(DEVICE_GPU), GpsigPdeDiagGradOp);
