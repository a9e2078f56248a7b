// TEST INFRASTRUCTURE: see op_kernel.h in this directory.
#pragma once
#include "tensorflow/core/framework/shape_inference.h"
namespace tensorflow {
struct OpDefBuilderStub {
  OpDefBuilderStub& Input(const char*);
  OpDefBuilderStub& Output(const char*);
  OpDefBuilderStub& Attr(const char*);
  OpDefBuilderStub& SetShapeFn(Status (*fn)(shape_inference::InferenceContext*));
};
OpDefBuilderStub RegisterOpStub(const char*);
#define REGISTER_OP(NAME) static ::tensorflow::OpDefBuilderStub& TF_STUB_CAT(tf_stub_op_, __LINE__) = ::tensorflow::RegisterOpStub(NAME)
}  // namespace tensorflow
