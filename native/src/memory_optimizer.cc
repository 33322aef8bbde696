#include "veles/memory_optimizer.h"

#include <algorithm>
#include <numeric>

namespace veles {

size_t MemoryOptimizer::Optimize(std::vector<MemoryNode>* nodes) const {
  std::vector<size_t> order(nodes->size());
  std::iota(order.begin(), order.end(), 0);
  auto up = [this](size_t v) { return (v + alignment_ - 1) / alignment_ * alignment_; };
  for (auto& n : *nodes) n.size = up(n.size);
  // biggest first; ties: earlier first
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) {
    if ((*nodes)[a].size != (*nodes)[b].size) return (*nodes)[a].size > (*nodes)[b].size;
    return (*nodes)[a].time_start < (*nodes)[b].time_start;
  });
  std::vector<size_t> placed;
  size_t arena = 0;
  for (size_t idx : order) {
    MemoryNode& n = (*nodes)[idx];
    // intervals occupied by already placed buffers that are alive at the same time
    std::vector<std::pair<size_t, size_t>> busy;
    for (size_t j : placed) {
      const MemoryNode& o = (*nodes)[j];
      if (o.time_start < n.time_finish && n.time_start < o.time_finish) busy.emplace_back(o.position, o.position + o.size);
    }
    std::sort(busy.begin(), busy.end());
    size_t pos = 0;
    for (const auto& iv : busy) {
      if (pos + n.size <= iv.first) break;  // fits into the gap before this interval
      pos = std::max(pos, iv.second);
    }
    n.position = pos;
    arena = std::max(arena, pos + n.size);
    placed.push_back(idx);
  }
  return arena;
}

}  // namespace veles
