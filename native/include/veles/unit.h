// veles_b200 native runtime — Unit: node of the inference graph (reference: libVeles/inc/veles/unit.h:105-200,
// libVeles/src/unit.cc:43-85).  Units consume ONE device tensor (bf16, [batch, ...]) produced by their first
// parent and produce one; execution order and buffer reuse are resolved once by the Workflow.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "veles/numpy_array_loader.h"

namespace veles {

struct Property {
  enum Kind { None, Bool, Number, String, Array, List } kind = None;
  bool b = false;
  double num = 0.0;
  std::string str;
  std::shared_ptr<NumpyArray> array;
  std::vector<double> list;
};

using Shape = std::vector<int64_t>;  // sample shape, WITHOUT the batch dimension

inline int64_t shape_size(const Shape& s) {
  int64_t n = 1;
  for (auto d : s) n *= d;
  return n;
}

class Unit {
 public:
  virtual ~Unit() = default;
  virtual const char* Uuid() const = 0;
  virtual const char* Name() const = 0;
  virtual void SetParameter(const std::string& name, const Property& value) = 0;
  // Shape of one output sample given the shape of one input sample.
  virtual Shape OutputShape(const Shape& input) const = 0;
  // Bytes of private device scratch needed for `batch` samples (0 if none).
  virtual size_t ScratchBytes(int /*batch*/, const Shape& /*input*/) const { return 0; }
  // Upload parameters; called once after every SetParameter, with the device current.
  virtual void Initialize(int /*batch*/, const Shape& /*input*/) {}
  // Enqueue the computation: in/out are bf16 device buffers with `batch` samples.
  virtual void Execute(const void* in, void* out, void* scratch, int batch, const Shape& input,
                       cudaStream_t stream) = 0;

  void LinkFrom(Unit* parent) {
    parents_.push_back(parent);
    parent->children_.push_back(this);
  }
  const std::vector<Unit*>& parents() const { return parents_; }
  const std::vector<Unit*>& children() const { return children_; }

 private:
  std::vector<Unit*> parents_;
  std::vector<Unit*> children_;
};

}  // namespace veles
