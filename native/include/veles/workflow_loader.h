// veles_b200 native runtime — WorkflowLoader: package (zip / tgz with contents.json + NNNN_shape.npy) -> Workflow
// (reference: libVeles/src/workflow_loader.cc:41-131, main_file_loader.cc:119-176; format written by
// veles_b200.workflow.Workflow.package_export and byte-compatible with the reference's).
#pragma once
#include <memory>
#include <string>

#include "veles/archive.h"
#include "veles/json.h"
#include "veles/workflow.h"

namespace veles {

class WorkflowLoader {
 public:
  static std::unique_ptr<Workflow> Load(const std::string& package_path);
  static std::unique_ptr<Workflow> FromArchive(const WorkflowArchive& archive);
  static Property ToProperty(const JsonValue& v, const WorkflowArchive& archive);
};

}  // namespace veles
