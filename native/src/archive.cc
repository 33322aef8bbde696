#include "veles/archive.h"

#include <zlib.h>

#include <cstdint>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace veles {

std::string read_file(const std::string& path) {
  std::ifstream f(path, std::ios::binary);
  if (!f) throw std::runtime_error("cannot open " + path);
  std::ostringstream ss;
  ss << f.rdbuf();
  return ss.str();
}

std::string gunzip(const std::string& blob) {
  z_stream zs;
  std::memset(&zs, 0, sizeof(zs));
  if (inflateInit2(&zs, 16 + MAX_WBITS) != Z_OK) throw std::runtime_error("inflateInit2 failed");
  zs.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(blob.data()));
  zs.avail_in = static_cast<uInt>(blob.size());
  std::string out;
  char buf[1 << 16];
  int rc;
  do {
    zs.next_out = reinterpret_cast<Bytef*>(buf);
    zs.avail_out = sizeof(buf);
    rc = inflate(&zs, Z_NO_FLUSH);
    if (rc != Z_OK && rc != Z_STREAM_END) {
      inflateEnd(&zs);
      throw std::runtime_error("gzip stream is corrupt");
    }
    out.append(buf, sizeof(buf) - zs.avail_out);
  } while (rc != Z_STREAM_END);
  inflateEnd(&zs);
  return out;
}

std::string inflate_raw(const char* data, size_t size, size_t expected) {
  z_stream zs;
  std::memset(&zs, 0, sizeof(zs));
  if (inflateInit2(&zs, -MAX_WBITS) != Z_OK) throw std::runtime_error("inflateInit2 failed");
  std::string out(expected, '\0');
  zs.next_in = reinterpret_cast<Bytef*>(const_cast<char*>(data));
  zs.avail_in = static_cast<uInt>(size);
  zs.next_out = reinterpret_cast<Bytef*>(&out[0]);
  zs.avail_out = static_cast<uInt>(expected);
  const int rc = inflate(&zs, Z_FINISH);
  inflateEnd(&zs);
  if (rc != Z_STREAM_END) throw std::runtime_error("deflate member is corrupt");
  return out;
}

static uint32_t rd32(const std::string& b, size_t o) {
  return static_cast<uint8_t>(b[o]) | (static_cast<uint8_t>(b[o + 1]) << 8) | (static_cast<uint8_t>(b[o + 2]) << 16) |
         (static_cast<uint32_t>(static_cast<uint8_t>(b[o + 3])) << 24);
}
static uint16_t rd16(const std::string& b, size_t o) {
  return static_cast<uint16_t>(static_cast<uint8_t>(b[o]) | (static_cast<uint8_t>(b[o + 1]) << 8));
}

void WorkflowArchive::load_zip(const std::string& blob) {
  // locate the end-of-central-directory record
  if (blob.size() < 22) throw std::runtime_error("zip: file too small");
  size_t eocd = std::string::npos;
  for (size_t i = blob.size() - 22;; --i) {
    if (rd32(blob, i) == 0x06054b50u) {
      eocd = i;
      break;
    }
    if (i == 0 || blob.size() - i > 65557) break;
  }
  if (eocd == std::string::npos) throw std::runtime_error("zip: end of central directory not found");
  const uint16_t entries = rd16(blob, eocd + 10);
  size_t p = rd32(blob, eocd + 16);
  for (uint16_t e = 0; e < entries; ++e) {
    if (p + 46 > blob.size() || rd32(blob, p) != 0x02014b50u) throw std::runtime_error("zip: bad central directory");
    const uint16_t method = rd16(blob, p + 10);
    const uint32_t csize = rd32(blob, p + 20), usize = rd32(blob, p + 24);
    const uint16_t nlen = rd16(blob, p + 28), xlen = rd16(blob, p + 30), clen = rd16(blob, p + 32);
    const uint32_t local = rd32(blob, p + 42);
    const std::string name = blob.substr(p + 46, nlen);
    p += 46 + nlen + xlen + clen;
    if (!name.empty() && name.back() == '/') continue;
    if (local + 30 > blob.size() || rd32(blob, local) != 0x04034b50u) throw std::runtime_error("zip: bad local header");
    const size_t data = local + 30 + rd16(blob, local + 26) + rd16(blob, local + 28);
    if (data + csize > blob.size()) throw std::runtime_error("zip: member exceeds the file");
    if (method == 0) files_[name] = blob.substr(data, csize);
    else if (method == 8) files_[name] = inflate_raw(blob.data() + data, csize, usize);
    else throw std::runtime_error("zip: unsupported compression method " + std::to_string(method));
  }
}

void WorkflowArchive::load_tar(const std::string& blob) {
  size_t p = 0;
  while (p + 512 <= blob.size()) {
    const char* h = blob.data() + p;
    if (h[0] == '\0') break;
    std::string name(h, strnlen(h, 100));
    const std::string prefix(h + 345, strnlen(h + 345, 155));
    if (!prefix.empty() && std::memcmp(h + 257, "ustar", 5) == 0) name = prefix + "/" + name;
    const size_t size = std::strtoull(std::string(h + 124, 12).c_str(), nullptr, 8);
    const char type = h[156];
    p += 512;
    if (p + size > blob.size()) throw std::runtime_error("tar: member exceeds the file");
    if (type == '0' || type == '\0') {
      if (name.rfind("./", 0) == 0) name = name.substr(2);
      files_[name] = blob.substr(p, size);
    }
    p += (size + 511) / 512 * 512;
  }
}

WorkflowArchive WorkflowArchive::Load(const std::string& path) {
  WorkflowArchive a;
  std::string blob = read_file(path);
  if (blob.size() >= 4 && rd32(blob, 0) == 0x04034b50u) {
    a.load_zip(blob);
  } else {
    if (blob.size() >= 2 && static_cast<uint8_t>(blob[0]) == 0x1f && static_cast<uint8_t>(blob[1]) == 0x8b)
      blob = gunzip(blob);
    a.load_tar(blob);
  }
  if (a.files_.empty()) throw std::runtime_error("package " + path + " is empty or has an unknown format");
  return a;
}

const std::string& WorkflowArchive::get(const std::string& name) const {
  auto it = files_.find(name);
  if (it == files_.end()) throw std::runtime_error("package has no member \"" + name + "\"");
  return it->second;
}

std::vector<std::string> WorkflowArchive::names() const {
  std::vector<std::string> out;
  for (const auto& kv : files_) out.push_back(kv.first);
  return out;
}

}  // namespace veles
