// CPU-only checks of the native runtime (no GPU needed): JSON, .npy, archives, memory optimizer, unit factory,
// workflow loading + planning.  Mirrors libVeles/tests/{workflow_archive,main_file_loader,numpy_array_loader,
// memory_optimizer,unit_factory,workflow_loader,workflow}.cc.  Usage: test_native [package.zip [package.tgz]]
#include <cmath>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <string>

#include "veles/archive.h"
#include "veles/json.h"
#include "veles/memory_optimizer.h"
#include "veles/numpy_array_loader.h"
#include "veles/unit_factory.h"
#include <typeinfo>

#include "veles/logger.h"
#include "veles/workflow_loader.h"

using namespace veles;

static int g_failed = 0, g_checks = 0;
#define CHECK(cond)                                                                  \
  do {                                                                               \
    ++g_checks;                                                                      \
    if (!(cond)) {                                                                   \
      ++g_failed;                                                                    \
      std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #cond);                  \
    }                                                                                \
  } while (0)

static void test_json() {
  JsonValue v = parse_json(R"({"workflow": "W", "n": -1.5e2, "ok": true, "none": null,
                              "units": [{"data": {"weights": "@0000_2x3", "s": "a\"bA"}, "links": [1]}, {}]})");
  CHECK(v.is_object() && v.at("workflow").str == "W");
  CHECK(std::fabs(v.at("n").num + 150.0) < 1e-9 && v.at("ok").b && v.at("none").type == JsonValue::Null);
  CHECK(v.at("units").arr.size() == 2 && v.at("units").arr[0].at("links").arr[0].num == 1);
  CHECK(v.at("units").arr[0].at("data").at("s").str == "a\"bA");
  bool threw = false;
  try { parse_json("{\"a\": }"); } catch (const std::exception&) { threw = true; }
  CHECK(threw);
}

static std::string make_npy(const std::string& descr, bool fortran, const std::string& shape, const std::string& payload) {
  std::string header = "{'descr': '" + descr + "', 'fortran_order': " + (fortran ? "True" : "False") +
                       ", 'shape': " + shape + ", }";
  while ((10 + header.size() + 1) % 64) header += ' ';
  header += '\n';
  std::string out("\x93NUMPY\x01\x00", 8);
  out += static_cast<char>(header.size() & 0xFF);
  out += static_cast<char>(header.size() >> 8);
  return out + header + payload;
}

static void test_npy() {
  float vals[6] = {0, 1, 2, 3, 4, 5};
  NumpyArray a = load_npy(make_npy("<f4", false, "(2, 3)", std::string(reinterpret_cast<char*>(vals), 24)));
  CHECK(a.shape.size() == 2 && a.shape[0] == 2 && a.shape[1] == 3 && a.data[4] == 4.0f);
  // Fortran order: memory holds columns; element [1][2] is the last one
  NumpyArray f = load_npy(make_npy("<f4", true, "(2, 3)", std::string(reinterpret_cast<char*>(vals), 24)));
  CHECK(f.data[0] == 0 && f.data[1] == 2 && f.data[2] == 4 && f.data[3] == 1 && f.data[5] == 5);
  uint16_t halves[3] = {0x3C00 /*1.0*/, 0xC000 /*-2.0*/, 0x0001 /*5.96e-8*/};
  NumpyArray h = load_npy(make_npy("<f2", false, "(3,)", std::string(reinterpret_cast<char*>(halves), 6)));
  CHECK(h.data[0] == 1.0f && h.data[1] == -2.0f && std::fabs(h.data[2] - 5.9604645e-8f) < 1e-12f);
  int32_t ints[2] = {7, -9};
  NumpyArray i = load_npy(make_npy("<i4", false, "(2,)", std::string(reinterpret_cast<char*>(ints), 8)));
  CHECK(i.data[0] == 7.0f && i.data[1] == -9.0f);
  bool threw = false;
  try { load_npy("garbage"); } catch (const std::exception&) { threw = true; }
  CHECK(threw);
}

static void test_memory_optimizer() {
  // a chain: each buffer is alive for two consecutive steps -> two slots are enough
  std::vector<MemoryNode> nodes(6);
  size_t naive = 0;
  for (int i = 0; i < 6; ++i) {
    nodes[i].size = 1000 + 100 * i;
    nodes[i].time_start = i;
    nodes[i].time_finish = i + 2;
    naive += (nodes[i].size + 255) / 256 * 256;
  }
  const size_t arena = MemoryOptimizer(256).Optimize(&nodes);
  CHECK(arena < naive && arena <= 2 * 1536 + 256);
  for (size_t a = 0; a < nodes.size(); ++a)
    for (size_t b = a + 1; b < nodes.size(); ++b) {
      const bool overlap_time = nodes[a].time_start < nodes[b].time_finish && nodes[b].time_start < nodes[a].time_finish;
      const bool overlap_mem = nodes[a].position < nodes[b].position + nodes[b].size &&
                               nodes[b].position < nodes[a].position + nodes[a].size;
      CHECK(!(overlap_time && overlap_mem));
    }
  for (const auto& n : nodes) CHECK(n.position % 256 == 0 && n.position + n.size <= arena);
  // everything alive at once -> no sharing possible
  std::vector<MemoryNode> all(3);
  for (auto& n : all) { n.size = 512; n.time_start = 0; n.time_finish = 5; }
  CHECK(MemoryOptimizer(256).Optimize(&all) == 3 * 512);
}

static void test_factory() {
  auto& f = UnitFactory::Instance();
  CHECK(f.Has("b3a2bd5c-3c01-46ef-978a-fef22e008f31") && f.Has("420219fc-3e1a-45b1-87f8-aaa0c1540de4"));
  auto u = f.Create("b3a2bd5c-3c01-46ef-978a-fef22e008f31");
  CHECK(std::string(u->Name()) == "All2AllTanh");
  auto by_name = f.Create("00000000-0000-0000-0000-000000000000", "MaxPooling");
  CHECK(std::string(by_name->Name()) == "MaxPooling");
  bool threw = false;
  try { f.Create("nope", "Nope"); } catch (const std::exception&) { threw = true; }
  CHECK(threw);
  // shape inference without a device
  Property p;
  p.kind = Property::Number;
  p.num = 3;
  by_name->SetParameter("kx", p);
  by_name->SetParameter("ky", p);
  Property s;
  s.kind = Property::List;
  s.list = {2, 2};
  by_name->SetParameter("sliding", s);
  const Shape o = by_name->OutputShape(Shape{55, 55, 96});
  CHECK(o[0] == 27 && o[1] == 27 && o[2] == 96);
}

static void test_package(const char* path) {
  WorkflowArchive ar = WorkflowArchive::Load(path);
  CHECK(ar.has("contents.json"));
  auto wf = WorkflowLoader::Load(path);
  CHECK(wf->Size() >= 2);
  std::printf("package %s: workflow \"%s\", %zu units, members:", path, wf->name().c_str(), wf->Size());
  for (const auto& n : ar.names()) std::printf(" %s", n.c_str());
  std::printf("\n");
  wf->Plan(64, Shape{784});
  CHECK(wf->order().size() == wf->Size());
  CHECK(wf->output_shape().size() == 1 && wf->output_shape()[0] == 10);
  size_t naive = 0;
  for (const auto& n : wf->memory_plan()) naive += n.size;
  CHECK(wf->arena_bytes() <= naive && wf->arena_bytes() > 0);
  std::printf("arena %zu bytes (sum of buffers %zu)\n", wf->arena_bytes(), naive);
}

// logger: thresholds (global and per domain), sink callback, level parsing, long records, class-name domains
struct SinkRecord {
  int level;
  std::string domain, message;
};
static std::vector<SinkRecord> g_sink;
static void test_sink(int level, const char* domain, const char* message, void* user) {
  (void)user;
  g_sink.push_back({level, domain, message});
}

static void test_logger() {
  using veles::Logger;
  using veles::LogLevel;
  LogLevel l;
  CHECK(Logger::ParseLevel("DEBUG", &l) && l == LogLevel::kDebug);
  CHECK(Logger::ParseLevel("warn", &l) && l == LogLevel::kWarning);
  CHECK(Logger::ParseLevel("1", &l) && l == LogLevel::kError);
  CHECK(!Logger::ParseLevel("loud", &l));
  CHECK(std::string(Logger::LevelName(LogLevel::kCritical)) == "CRITICAL");
  CHECK(Logger::Demangle(typeid(veles::WorkflowLoader).name()) == "WorkflowLoader");
  const LogLevel before = Logger::Level();
  Logger::SetSink(test_sink, nullptr);
  Logger::SetLevel(LogLevel::kWarning);
  Logger a("Alpha", veles::kLogColorCyan), b("Beta");
  VELES_INF(a, "hidden %d", 1);
  VELES_WRN(a, "shown %d", 2);
  CHECK(g_sink.size() == 1 && g_sink[0].level == 2 && g_sink[0].domain == "Alpha" && g_sink[0].message == "shown 2");
  Logger::SetDomainLevel("Beta", LogLevel::kDebug);
  VELES_DBG(b, "beta debug");
  VELES_DBG(a, "alpha debug");
  CHECK(g_sink.size() == 2 && g_sink[1].domain == "Beta" && g_sink[1].level == 4);
  CHECK(b.enabled(LogLevel::kDebug) && !a.enabled(LogLevel::kInfo) && a.enabled(LogLevel::kError));
  const std::string big(2000, 'x');
  VELES_ERR(a, "%s!", big.c_str());
  CHECK(g_sink.size() == 3 && g_sink[2].message.size() == 2001 && g_sink[2].message.back() == '!');
  Logger::ClearDomainLevels();
  VELES_DBG(b, "gone");
  CHECK(g_sink.size() == 3);
  Logger::SetSink(nullptr, nullptr);
  Logger::SetLevel(before);
}

int main(int argc, char** argv) {
  test_logger();
  test_json();
  test_npy();
  test_memory_optimizer();
  test_factory();
  for (int i = 1; i < argc; ++i) test_package(argv[i]);
  std::printf("%d checks, %d failed\n", g_checks, g_failed);
  return g_failed ? 1 : 0;
}
