#include "veles/workflow_loader.h"

#include <stdexcept>

#include "veles/logger.h"
#include "veles/unit_factory.h"

namespace veles {

static const Logger& loader_log() {
  static const Logger log("WorkflowLoader", kLogColorCyan);
  return log;
}

Property WorkflowLoader::ToProperty(const JsonValue& v, const WorkflowArchive& archive) {
  Property p;
  switch (v.type) {
    case JsonValue::Bool: p.kind = Property::Bool; p.b = v.b; break;
    case JsonValue::Number: p.kind = Property::Number; p.num = v.num; break;
    case JsonValue::String:
      if (!v.str.empty() && v.str[0] == '@') {  // "@0001_100x784" -> 0001_100x784.npy
        p.kind = Property::Array;
        p.array = std::make_shared<NumpyArray>(load_npy(archive.get(v.str.substr(1) + ".npy")));
      } else {
        p.kind = Property::String;
        p.str = v.str;
      }
      break;
    case JsonValue::Array:
      p.kind = Property::List;
      for (const auto& e : v.arr) p.list.push_back(e.is_number() ? e.num : (e.is_bool() ? (e.b ? 1.0 : 0.0) : 0.0));
      break;
    default: break;
  }
  return p;
}

std::unique_ptr<Workflow> WorkflowLoader::FromArchive(const WorkflowArchive& archive) {
  const JsonValue doc = parse_json(archive.get("contents.json"));
  std::unique_ptr<Workflow> wf(new Workflow());
  if (const JsonValue* n = doc.find("workflow")) wf->set_name(n->str);
  if (const JsonValue* c = doc.find("checksum")) wf->set_checksum(c->str);
  const JsonValue& units = doc.at("units");
  if (!units.is_array() || units.arr.empty()) throw std::runtime_error("contents.json has no units");
  for (const JsonValue& u : units.arr) {
    const JsonValue& cls = u.at("class");
    std::unique_ptr<Unit> unit = UnitFactory::Instance().Create(cls.at("uuid").str, cls.at("name").str);
    VELES_DBG(loader_log(), "unit %zu: %s (%s)", wf->Size(), cls.at("name").str.c_str(), cls.at("uuid").str.c_str());
    // arrays last: scalar properties (include_bias, geometry, ...) must be known when weights arrive
    const JsonValue& data = u.at("data");
    for (int pass = 0; pass < 2; ++pass)
      for (const auto& kv : data.obj) {
        const bool is_array = kv.second.is_string() && !kv.second.str.empty() && kv.second.str[0] == '@';
        if (is_array == (pass == 1)) unit->SetParameter(kv.first, ToProperty(kv.second, archive));
      }
    wf->AddUnit(std::move(unit));
  }
  for (size_t i = 0; i < units.arr.size(); ++i) {
    const JsonValue* links = units.arr[i].find("links");
    if (!links) continue;
    for (const JsonValue& l : links->arr) {
      const size_t j = static_cast<size_t>(l.num);
      if (j >= wf->Size()) throw std::runtime_error("contents.json links to a unit that does not exist");
      wf->Get(j)->LinkFrom(wf->Get(i));
    }
  }
  VELES_INF(loader_log(), "loaded workflow \"%s\" (%zu units, checksum %s)", wf->name().c_str(), wf->Size(),
            wf->checksum().empty() ? "-" : wf->checksum().c_str());
  return wf;
}

std::unique_ptr<Workflow> WorkflowLoader::Load(const std::string& package_path) {
  VELES_DBG(loader_log(), "opening %s", package_path.c_str());
  return FromArchive(WorkflowArchive::Load(package_path));
}

}  // namespace veles
