#include "veles/unit_factory.h"

#include <stdexcept>

namespace veles {

UnitFactory& UnitFactory::Instance() {
  static UnitFactory f;
  return f;
}

void UnitFactory::Register(const std::string& uuid, const std::string& name, Constructor ctor) {
  by_uuid_[uuid] = ctor;
  by_name_[name] = ctor;
}

std::unique_ptr<Unit> UnitFactory::Create(const std::string& uuid, const std::string& name) const {
  auto it = by_uuid_.find(uuid);
  if (it != by_uuid_.end()) return it->second();
  auto jt = by_name_.find(name);
  if (jt != by_name_.end()) return jt->second();
  throw std::runtime_error("no native unit is registered for uuid " + uuid + " (class " + name + ")");
}

std::vector<std::string> UnitFactory::Names() const {
  std::vector<std::string> out;
  for (const auto& kv : by_name_) out.push_back(kv.first);
  return out;
}

}  // namespace veles
