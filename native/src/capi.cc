// veles_b200 native runtime — C API (used by the Python test through ctypes and by embedding applications).
#include <cstring>
#include <string>

#include "veles/logger.h"
#include "veles/unit_factory.h"
#include "veles/workflow_loader.h"

using namespace veles;

static thread_local std::string g_err;

extern "C" {

const char* veles_native_last_error() { return g_err.c_str(); }

void* veles_native_load(const char* package_path) {
  try {
    return WorkflowLoader::Load(package_path).release();
  } catch (const std::exception& e) {
    g_err = e.what();
    return nullptr;
  }
}

int veles_native_initialize(void* wf, int batch, const long long* input_shape, int ndim, int device, int use_graph) {
  try {
    Shape s(input_shape, input_shape + ndim);
    static_cast<Workflow*>(wf)->set_use_graph(use_graph != 0);
    static_cast<Workflow*>(wf)->Initialize(batch, s, device);
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

int veles_native_plan(void* wf, int batch, const long long* input_shape, int ndim, long long* arena_bytes,
                      long long* naive_bytes) {
  try {
    Workflow* w = static_cast<Workflow*>(wf);
    w->Plan(batch, Shape(input_shape, input_shape + ndim));
    *arena_bytes = static_cast<long long>(w->arena_bytes());
    long long total = 0;
    for (const auto& n : w->memory_plan()) total += static_cast<long long>(n.size);
    *naive_bytes = total;
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

long long veles_native_output_size(void* wf) { return shape_size(static_cast<Workflow*>(wf)->output_shape()); }
int veles_native_unit_count(void* wf) { return static_cast<int>(static_cast<Workflow*>(wf)->Size()); }
const char* veles_native_unit_name(void* wf, int i) { return static_cast<Workflow*>(wf)->Get(i)->Name(); }
const char* veles_native_name(void* wf) { return static_cast<Workflow*>(wf)->name().c_str(); }
int veles_native_uses_graph(void* wf) { return static_cast<Workflow*>(wf)->uses_graph() ? 1 : 0; }

int veles_native_run(void* wf, const float* input, float* output, int batch) {
  try {
    static_cast<Workflow*>(wf)->Run(input, output, batch);
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

void veles_native_free(void* wf) { delete static_cast<Workflow*>(wf); }

// -- logging (veles/logger.h) -------------------------------------------------------------------------------------
// level: 0 critical .. 4 debug; returns the previous global threshold, -1 for a bad argument
int veles_native_set_log_level(int level) {
  if (level < 0 || level > 4) return -1;
  const int prev = static_cast<int>(Logger::Level());
  Logger::SetLevel(static_cast<LogLevel>(level));
  return prev;
}
int veles_native_set_domain_log_level(const char* domain, int level) {
  if (domain == nullptr || level < 0 || level > 4) return -1;
  Logger::SetDomainLevel(domain, static_cast<LogLevel>(level));
  return 0;
}
// records go to sink(level, domain, message, user) instead of stderr; nullptr restores stderr
void veles_native_set_log_sink(LogSink sink, void* user) { Logger::SetSink(sink, user); }
// emit a record through the runtime's logger (embedding applications share its sink and thresholds)
void veles_native_log(int level, const char* domain, const char* message) {
  if (level < 0 || level > 4 || domain == nullptr || message == nullptr) return;
  Logger(domain).log(static_cast<LogLevel>(level), "%s", message);
}

}  // extern "C"
