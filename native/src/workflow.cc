#include "veles/workflow.h"

#include "veles/logger.h"

#include <algorithm>
#include <map>
#include <stdexcept>

#include "../../include/veles_b200.h"

namespace veles {

#define VN_CUDA(expr)                                                                          \
  do {                                                                                         \
    cudaError_t e_ = (expr);                                                                   \
    if (e_ != cudaSuccess) throw std::runtime_error(std::string(#expr) + ": " + cudaGetErrorString(e_)); \
  } while (0)

Workflow::~Workflow() {
  if (graph_exec_) cudaGraphExecDestroy(graph_exec_);
  if (arena_) cudaFree(arena_);
  if (staging_in_) cudaFree(staging_in_);
  if (staging_out_) cudaFree(staging_out_);
  if (stream_) cudaStreamDestroy(stream_);
}

void Workflow::Plan(int batch, const Shape& input_shape) {
  if (units_.empty()) throw std::runtime_error("workflow has no units");
  const int n = static_cast<int>(units_.size());
  std::map<const Unit*, int> index;
  for (int i = 0; i < n; ++i) index[units_[i].get()] = i;
  // Kahn topological sort (stable: file order among ready units)
  std::vector<int> indeg(n, 0);
  for (int i = 0; i < n; ++i) indeg[i] = static_cast<int>(units_[i]->parents().size());
  order_.clear();
  std::vector<bool> done(n, false);
  for (int step = 0; step < n; ++step) {
    int pick = -1;
    for (int i = 0; i < n; ++i)
      if (!done[i] && indeg[i] == 0) {
        pick = i;
        break;
      }
    if (pick < 0) throw std::runtime_error("workflow graph has a cycle");
    done[pick] = true;
    order_.push_back(pick);
    for (Unit* c : units_[pick]->children()) --indeg[index[c]];
  }
  // shapes
  in_shapes_.assign(n, Shape());
  out_shapes_.assign(n, Shape());
  for (int i : order_) {
    const auto& ps = units_[i]->parents();
    in_shapes_[i] = ps.empty() ? input_shape : out_shapes_[index[ps[0]]];
    out_shapes_[i] = units_[i]->OutputShape(in_shapes_[i]);
  }
  // lifetimes: node 0 = input (alive from step 0 until its last consumer), node 1+k = output of order_[k]
  std::vector<int> pos(n);
  for (int k = 0; k < n; ++k) pos[order_[k]] = k;
  nodes_.assign(1 + n, MemoryNode());
  nodes_[0].size = static_cast<size_t>(batch) * shape_size(input_shape) * 2;
  nodes_[0].time_start = 0;
  nodes_[0].time_finish = 1;
  for (int k = 0; k < n; ++k) {
    const int i = order_[k];
    MemoryNode& node = nodes_[1 + k];
    node.size = static_cast<size_t>(batch) * shape_size(out_shapes_[i]) * 2;
    node.time_start = k;
    int last = k + 1;  // at least alive while it is produced; the final output stays alive to the end
    for (Unit* c : units_[i]->children()) last = std::max(last, pos[index[c]] + 1);
    if (units_[i]->children().empty()) last = n + 1;
    node.time_finish = last;
    if (units_[i]->parents().empty()) nodes_[0].time_finish = std::max(nodes_[0].time_finish, k + 1);
  }
  scratch_node_.assign(n, -1);
  for (int k = 0; k < n; ++k) {
    const int i = order_[k];
    const size_t sb = units_[i]->ScratchBytes(batch, in_shapes_[i]);
    if (sb) {
      MemoryNode s;
      s.size = sb;
      s.time_start = k;
      s.time_finish = k + 1;
      scratch_node_[k] = static_cast<int>(nodes_.size());
      nodes_.push_back(s);
    }
  }
  arena_bytes_ = MemoryOptimizer(256).Optimize(&nodes_);
  batch_ = batch;
  input_shape_ = input_shape;
}

void Workflow::Initialize(int batch, const Shape& input_shape, int device) {
  static const Logger log("Workflow", kLogColorGreen);
  VN_CUDA(cudaSetDevice(device));
  Plan(batch, input_shape);
  size_t naive = 0;
  for (const auto& node : nodes_) naive += node.size;
  VELES_INF(log, "\"%s\": batch %d on device %d, activation arena %zu bytes (%zu without buffer reuse)", name_.c_str(),
            batch, device, arena_bytes_, naive);
  VN_CUDA(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
  VN_CUDA(cudaMalloc(reinterpret_cast<void**>(&arena_), arena_bytes_));
  VN_CUDA(cudaMemset(arena_, 0, arena_bytes_));
  VN_CUDA(cudaMalloc(reinterpret_cast<void**>(&staging_in_), sizeof(float) * batch * shape_size(input_shape)));
  VN_CUDA(cudaMalloc(reinterpret_cast<void**>(&staging_out_), sizeof(float) * batch * shape_size(output_shape())));
  for (int i : order_) units_[i]->Initialize(batch, in_shapes_[i]);
  VN_CUDA(cudaDeviceSynchronize());
}

void Workflow::Enqueue(int batch) {
  const int n = static_cast<int>(units_.size());
  std::map<const Unit*, int> pos;
  for (int k = 0; k < n; ++k) pos[units_[order_[k]].get()] = k;
  const long long in_elems = static_cast<long long>(batch) * shape_size(input_shape_);
  if (vb_cast(staging_in_, 0, arena_ + nodes_[0].position, 1, in_elems, stream_))
    throw std::runtime_error(std::string("vb_cast: ") + vb_last_error());
  for (int k = 0; k < n; ++k) {
    Unit* u = units_[order_[k]].get();
    const void* in = u->parents().empty() ? arena_ + nodes_[0].position
                                          : arena_ + nodes_[1 + pos[u->parents()[0]]].position;
    void* out = arena_ + nodes_[1 + k].position;
    void* scratch = scratch_node_[k] >= 0 ? arena_ + nodes_[scratch_node_[k]].position : nullptr;
    u->Execute(in, out, scratch, batch, in_shapes_[order_[k]], stream_);
  }
  const long long out_elems = static_cast<long long>(batch) * shape_size(output_shape());
  if (vb_cast(arena_ + nodes_[n].position, 1, staging_out_, 0, out_elems, stream_))
    throw std::runtime_error(std::string("vb_cast: ") + vb_last_error());
}

void Workflow::Run(const float* input, float* output, int batch) {
  if (!arena_) throw std::runtime_error("Workflow::Run before Initialize");
  if (batch < 0) batch = batch_;
  if (batch > batch_) throw std::runtime_error("batch exceeds the one the workflow was initialized for");
  const size_t in_bytes = sizeof(float) * batch * shape_size(input_shape_);
  const size_t out_bytes = sizeof(float) * batch * shape_size(output_shape());
  VN_CUDA(cudaMemcpyAsync(staging_in_, input, in_bytes, cudaMemcpyHostToDevice, stream_));
  const bool full = batch == batch_;
  if (use_graph_ && full && graph_exec_ == nullptr && runs_ >= 1) {
    // first run was eager (lazy module loading, attribute setup); capture the steady state now
    cudaGraph_t graph = nullptr;
    VN_CUDA(cudaStreamBeginCapture(stream_, cudaStreamCaptureModeThreadLocal));
    try {
      Enqueue(batch);
    } catch (...) {
      cudaStreamEndCapture(stream_, &graph);
      if (graph) cudaGraphDestroy(graph);
      throw;
    }
    VN_CUDA(cudaStreamEndCapture(stream_, &graph));
    VN_CUDA(cudaGraphInstantiate(&graph_exec_, graph, 0));
    cudaGraphDestroy(graph);
  }
  if (graph_exec_ && full) VN_CUDA(cudaGraphLaunch(graph_exec_, stream_));
  else Enqueue(batch);
  VN_CUDA(cudaMemcpyAsync(output, staging_out_, out_bytes, cudaMemcpyDeviceToHost, stream_));
  VN_CUDA(cudaStreamSynchronize(stream_));
  ++runs_;
}

}  // namespace veles
