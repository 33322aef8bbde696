// veles_b200 native runtime — UnitFactory: class UUID -> constructor, filled by static registration
// (reference: libVeles/inc/veles/unit_factory.h:51-117, REGISTER_UNIT macro).
#pragma once
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <vector>

#include "veles/unit.h"

namespace veles {

class UnitFactory {
 public:
  using Constructor = std::function<std::unique_ptr<Unit>()>;
  static UnitFactory& Instance();
  void Register(const std::string& uuid, const std::string& name, Constructor ctor);
  // Looks up by UUID first, then by class name (packages written by other builds of the exporter).
  std::unique_ptr<Unit> Create(const std::string& uuid, const std::string& name = "") const;
  bool Has(const std::string& uuid) const { return by_uuid_.count(uuid) != 0; }
  std::vector<std::string> Names() const;

 private:
  std::map<std::string, Constructor> by_uuid_;
  std::map<std::string, Constructor> by_name_;
};

template <class T>
struct UnitRegistrar {
  UnitRegistrar(const char* uuid, const char* name) {
    UnitFactory::Instance().Register(uuid, name, [] { return std::unique_ptr<Unit>(new T()); });
  }
};

#define VELES_REGISTER_UNIT(T, uuid, name) static ::veles::UnitRegistrar<T> veles_registrar_##T(uuid, name)

}  // namespace veles
