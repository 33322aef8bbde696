// veles_b200 native runtime — .npy reader: header parse, f16/f32/f64/i16/i32/u8 -> float32 conversion,
// optional Fortran-order -> C-order transposition (reference: libVeles/src/numpy_array_loader.cc:38-246).
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace veles {

struct NumpyArray {
  std::vector<int64_t> shape;
  std::vector<float> data;  // always C order, float32
  int64_t size() const {
    int64_t n = 1;
    for (auto d : shape) n *= d;
    return n;
  }
};

NumpyArray load_npy(const std::string& blob);
float half_to_float(uint16_t h);

}  // namespace veles
