// TEST INFRASTRUCTURE: a declaration-only stand-in for the handful of TensorFlow 1.15 types shim/gpsig_tf_ops.cc uses,
// so that the shim can be SYNTAX- and TYPE-checked against include/gpsig_b200.h in an image without TensorFlow
// (tests/test_abi_cpu.py::test_tf_shim_compiles_against_the_c_abi).  Nothing here is TensorFlow code and nothing links.
#pragma once
#include <initializer_list>
#include <string>
#include <vector>

namespace Eigen { struct GpuDevice { void* stream() const; }; }

namespace tensorflow {
typedef long long int64;
typedef unsigned char uint8;
enum DataType { DT_UINT8 = 4 };
extern const char* const DEVICE_GPU;

class Status {
 public:
  static Status OK();
  bool ok() const;
};
namespace errors {
template <typename... A> Status Internal(A...);
template <typename... A> Status InvalidArgument(A...);
}  // namespace errors

class TensorShape {
 public:
  TensorShape(std::initializer_list<int64> dims);
};
template <typename T> struct Flat { T* data() const; };
class Tensor {
 public:
  int dims() const;
  int64 dim_size(int i) const;
  int64 NumElements() const;
  const TensorShape& shape() const;
  template <typename T> Flat<T> flat() const;
};

class OpKernelConstruction {
 public:
  template <typename T> Status GetAttr(const char* name, T* value) const;
  void CtxFailure(const Status& s);
};
class OpKernelContext {
 public:
  const Tensor& input(int i);
  Status allocate_output(int i, const TensorShape& shape, Tensor** out);
  Status allocate_temp(DataType dt, const TensorShape& shape, Tensor* out);
  template <typename D> const D& eigen_device() const;
  void CtxFailure(const Status& s);
};
class OpKernel {
 public:
  explicit OpKernel(OpKernelConstruction*);
  virtual ~OpKernel();
  virtual void Compute(OpKernelContext* ctx) = 0;
};

#define OP_REQUIRES(CTX, EXP, STATUS) do { if (!(EXP)) { (CTX)->CtxFailure((STATUS)); return; } } while (0)
#define OP_REQUIRES_OK(CTX, ...) do { ::tensorflow::Status s_((__VA_ARGS__)); if (!s_.ok()) { (CTX)->CtxFailure(s_); return; } } while (0)
#define TF_RETURN_IF_ERROR(...) do { ::tensorflow::Status s_((__VA_ARGS__)); if (!s_.ok()) return s_; } while (0)

struct KernelDefBuilder {
  KernelDefBuilder& Device(const char*);
  template <typename T> KernelDefBuilder& TypeConstraint(const char*);
};
KernelDefBuilder Name(const char*);
#define TF_STUB_CAT2(a, b) a##b
#define TF_STUB_CAT(a, b) TF_STUB_CAT2(a, b)
#define REGISTER_KERNEL_BUILDER(BUILDER, ...) \
  static ::tensorflow::OpKernel* TF_STUB_CAT(tf_stub_make_, __LINE__)(::tensorflow::OpKernelConstruction* c) { (void)(BUILDER); return new __VA_ARGS__(c); }
}  // namespace tensorflow
