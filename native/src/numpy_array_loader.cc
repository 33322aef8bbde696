#include "veles/numpy_array_loader.h"

#include <cstring>
#include <stdexcept>

namespace veles {

float half_to_float(uint16_t h) {
  const uint32_t sign = (h & 0x8000u) << 16;
  uint32_t exp = (h >> 10) & 0x1F;
  uint32_t man = h & 0x3FF;
  uint32_t bits;
  if (exp == 0) {
    if (man == 0) {
      bits = sign;
    } else {  // subnormal
      exp = 127 - 15 + 1;
      while (!(man & 0x400)) {
        man <<= 1;
        --exp;
      }
      bits = sign | (exp << 23) | ((man & 0x3FF) << 13);
    }
  } else if (exp == 31) {
    bits = sign | 0x7F800000u | (man << 13);
  } else {
    bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
  }
  float f;
  std::memcpy(&f, &bits, 4);
  return f;
}

static std::string dict_value(const std::string& header, const std::string& key) {
  size_t p = header.find("'" + key + "'");
  if (p == std::string::npos) throw std::runtime_error("npy: header lacks '" + key + "'");
  p = header.find(':', p) + 1;
  while (p < header.size() && header[p] == ' ') ++p;
  size_t e = p;
  if (header[p] == '(') e = header.find(')', p) + 1;
  else if (header[p] == '\'') e = header.find('\'', p + 1) + 1;
  else while (e < header.size() && header[e] != ',' && header[e] != '}') ++e;
  return header.substr(p, e - p);
}

NumpyArray load_npy(const std::string& blob) {
  if (blob.size() < 10 || std::memcmp(blob.data(), "\x93NUMPY", 6) != 0) throw std::runtime_error("npy: bad magic");
  const int major = static_cast<uint8_t>(blob[6]);
  size_t hlen, hoff;
  if (major == 1) {
    hlen = static_cast<uint8_t>(blob[8]) | (static_cast<uint8_t>(blob[9]) << 8);
    hoff = 10;
  } else {
    hlen = static_cast<uint8_t>(blob[8]) | (static_cast<uint8_t>(blob[9]) << 8) | (static_cast<uint8_t>(blob[10]) << 16) |
           (static_cast<size_t>(static_cast<uint8_t>(blob[11])) << 24);
    hoff = 12;
  }
  if (hoff + hlen > blob.size()) throw std::runtime_error("npy: truncated header");
  const std::string header = blob.substr(hoff, hlen);
  std::string descr = dict_value(header, "descr");
  descr = descr.substr(1, descr.size() - 2);
  const bool fortran = dict_value(header, "fortran_order").find("True") != std::string::npos;
  const std::string shp = dict_value(header, "shape");
  NumpyArray arr;
  for (size_t i = 1; i < shp.size();) {
    while (i < shp.size() && (shp[i] == ' ' || shp[i] == ',')) ++i;
    if (i >= shp.size() || shp[i] == ')') break;
    arr.shape.push_back(std::strtoll(shp.c_str() + i, nullptr, 10));
    while (i < shp.size() && shp[i] >= '0' && shp[i] <= '9') ++i;
  }
  if (descr.size() < 3) throw std::runtime_error("npy: bad descr " + descr);
  if (descr[0] == '>') throw std::runtime_error("npy: big-endian arrays are not supported");
  const char kind = descr[1];
  const int isz = std::atoi(descr.c_str() + 2);
  const int64_t n = arr.size();
  const char* src = blob.data() + hoff + hlen;
  if (hoff + hlen + static_cast<size_t>(n) * isz > blob.size()) throw std::runtime_error("npy: truncated data");
  std::vector<float> flat(n);
  for (int64_t i = 0; i < n; ++i) {
    const char* e = src + i * isz;
    float v;
    if (kind == 'f' && isz == 4) { std::memcpy(&v, e, 4); }
    else if (kind == 'f' && isz == 2) { uint16_t h; std::memcpy(&h, e, 2); v = half_to_float(h); }
    else if (kind == 'f' && isz == 8) { double d; std::memcpy(&d, e, 8); v = static_cast<float>(d); }
    else if (kind == 'i' && isz == 4) { int32_t t; std::memcpy(&t, e, 4); v = static_cast<float>(t); }
    else if (kind == 'i' && isz == 2) { int16_t t; std::memcpy(&t, e, 2); v = static_cast<float>(t); }
    else if (kind == 'i' && isz == 8) { int64_t t; std::memcpy(&t, e, 8); v = static_cast<float>(t); }
    else if (kind == 'u' && isz == 1) { v = static_cast<float>(static_cast<uint8_t>(*e)); }
    else if (kind == 'b' && isz == 1) { v = *e ? 1.0f : 0.0f; }
    else throw std::runtime_error("npy: unsupported dtype " + descr);
    flat[i] = v;
  }
  if (!fortran || arr.shape.size() < 2) {
    arr.data.swap(flat);
    return arr;
  }
  // Fortran order -> C order
  arr.data.resize(n);
  const size_t nd = arr.shape.size();
  std::vector<int64_t> idx(nd, 0);
  for (int64_t c = 0; c < n; ++c) {
    int64_t f = 0, stride = 1;
    for (size_t d = 0; d < nd; ++d) {
      f += idx[d] * stride;
      stride *= arr.shape[d];
    }
    arr.data[c] = flat[f];
    for (size_t d = nd; d-- > 0;) {
      if (++idx[d] < arr.shape[d]) break;
      idx[d] = 0;
    }
  }
  return arr;
}

}  // namespace veles
