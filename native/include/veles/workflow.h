// veles_b200 native runtime — Workflow: units in execution order + arena plan + stream engine
// (reference: libVeles/inc/veles/workflow.h:72-127, libVeles/src/workflow.cc:73-158, engine.cc:61-75).
// The reference ran units on a 2-thread pool; here the whole forward pass is enqueued on one CUDA stream and,
// after the first run, replayed as a CUDA graph.
#pragma once
#include <cuda_runtime.h>

#include <memory>
#include <string>
#include <vector>

#include "veles/memory_optimizer.h"
#include "veles/unit.h"

namespace veles {

class Workflow {
 public:
  Workflow() = default;
  ~Workflow();
  Workflow(const Workflow&) = delete;
  Workflow& operator=(const Workflow&) = delete;

  const std::string& name() const { return name_; }
  const std::string& checksum() const { return checksum_; }
  void set_name(const std::string& n) { name_ = n; }
  void set_checksum(const std::string& c) { checksum_ = c; }
  void AddUnit(std::unique_ptr<Unit> unit) { units_.push_back(std::move(unit)); }
  size_t Size() const { return units_.size(); }
  Unit* Get(size_t i) const { return units_[i].get(); }

  // Host-only planning (testable without a GPU): topological order, per-unit shapes, arena offsets.
  void Plan(int batch, const Shape& input_shape);
  const std::vector<int>& order() const { return order_; }
  const std::vector<Shape>& shapes() const { return out_shapes_; }
  const std::vector<MemoryNode>& memory_plan() const { return nodes_; }
  size_t arena_bytes() const { return arena_bytes_; }
  Shape output_shape() const { return out_shapes_.empty() ? Shape() : out_shapes_[order_.back()]; }

  // Device part: allocate the arena, upload parameters.
  void Initialize(int batch, const Shape& input_shape, int device = 0);
  // input: batch * prod(input_shape) floats on the host; output: batch * prod(output_shape()) floats.
  void Run(const float* input, float* output, int batch = -1);
  bool uses_graph() const { return graph_exec_ != nullptr; }
  void set_use_graph(bool v) { use_graph_ = v; }

 private:
  void Enqueue(int batch);
  std::string name_, checksum_;
  std::vector<std::unique_ptr<Unit>> units_;
  std::vector<int> order_;
  std::vector<Shape> in_shapes_, out_shapes_;
  std::vector<MemoryNode> nodes_;   // [0] = network input, [1 + i] = output of unit order_[i], then scratch buffers
  std::vector<int> scratch_node_;   // per ordered unit: node index or -1
  size_t arena_bytes_ = 0;
  int batch_ = 0;
  Shape input_shape_;
  char* arena_ = nullptr;
  float* staging_in_ = nullptr;
  float* staging_out_ = nullptr;
  cudaStream_t stream_ = nullptr;
  cudaGraphExec_t graph_exec_ = nullptr;
  bool use_graph_ = true;
  int runs_ = 0;
};

}  // namespace veles
