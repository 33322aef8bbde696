// veles_b200 native runtime — package archive reader: zip (stored/deflate) and tar(.gz) via zlib.
// Replaces the libarchive wrapper of the reference (libVeles/src/workflow_archive.cc:54-168).
#pragma once
#include <map>
#include <string>
#include <vector>

namespace veles {

class WorkflowArchive {
 public:
  // Reads every member of a .zip / .tar / .tar.gz / .tgz package into memory.
  static WorkflowArchive Load(const std::string& path);
  bool has(const std::string& name) const { return files_.count(name) != 0; }
  const std::string& get(const std::string& name) const;
  std::vector<std::string> names() const;

 private:
  std::map<std::string, std::string> files_;
  void load_zip(const std::string& blob);
  void load_tar(const std::string& blob);
};

std::string read_file(const std::string& path);
std::string gunzip(const std::string& blob);
std::string inflate_raw(const char* data, size_t size, size_t expected);

}  // namespace veles
