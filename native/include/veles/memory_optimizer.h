// veles_b200 native runtime — arena planner: every unit output is a buffer alive during
// [time_start, time_finish); buffers whose lifetimes do not overlap may share bytes.  Greedy
// first-fit-decreasing placement into one arena (idea of libVeles/src/memory_optimizer.cc:38-98), so the whole
// inference needs ONE cudaMalloc.
#pragma once
#include <cstddef>
#include <vector>

namespace veles {

struct MemoryNode {
  size_t size = 0;        // bytes (aligned up by the optimizer)
  int time_start = 0;     // first step that writes the buffer
  int time_finish = 0;    // first step at which it is dead (exclusive)
  size_t position = 0;    // out: byte offset in the arena
};

class MemoryOptimizer {
 public:
  explicit MemoryOptimizer(size_t alignment = 256) : alignment_(alignment) {}
  // Assigns positions; returns the arena size in bytes.
  size_t Optimize(std::vector<MemoryNode>* nodes) const;

 private:
  size_t alignment_;
};

}  // namespace veles
