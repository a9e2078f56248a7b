// TEST INFRASTRUCTURE: see op_kernel.h in this directory.
#pragma once
#include "tensorflow/core/framework/op_kernel.h"
namespace tensorflow {
namespace shape_inference {
struct ShapeHandle {};
struct DimensionHandle {};
class InferenceContext {
 public:
  ShapeHandle input(int i);
  Status WithRank(ShapeHandle s, int rank, ShapeHandle* out);
  DimensionHandle Dim(ShapeHandle s, int i);
  DimensionHandle MakeDim(int64 v);
  ShapeHandle MakeShape(std::initializer_list<DimensionHandle> dims);
  void set_output(int i, ShapeHandle s);
  template <typename T> Status GetAttr(const char* name, T* value) const;
};
}  // namespace shape_inference
}  // namespace tensorflow
