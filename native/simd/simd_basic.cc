// Host array primitives: memory helpers, dtype conversion, element-wise arithmetic, small dense matrices.
// Behavioural counterpart of the reference's libSimd memory.c / arithmetic.h / mathfun.h / matrix.c; the loops are
// written so that g++ -O3 vectorises them for whatever ISA the node has.
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "veles/simd.h"

extern "C" {

void* vs_malloc_aligned(size_t bytes) {
  void* p = nullptr;
  if (posix_memalign(&p, 64, bytes ? bytes : 64) != 0) return nullptr;
  return p;
}

float* vs_mallocf(size_t length) { return static_cast<float*>(vs_malloc_aligned(length * sizeof(float))); }

void vs_free(void* ptr) { free(ptr); }

int vs_align_complement_f32(const float* ptr) {
  const uintptr_t a = reinterpret_cast<uintptr_t>(ptr);
  const uintptr_t rem = a & 63u;
  return rem == 0 ? 0 : static_cast<int>((64u - rem) / sizeof(float));
}

void vs_memsetf(float* ptr, float value, size_t length) {
  for (size_t i = 0; i < length; ++i) ptr[i] = value;
}

size_t vs_next_highest_power_of_2(size_t value) {
  size_t p = 1;
  while (p < value) p <<= 1;
  return p;
}

float* vs_zeropaddingex(const float* ptr, size_t length, size_t* new_length, size_t additional_length) {
  const size_t nl = 2 * vs_next_highest_power_of_2(length) + additional_length;
  float* out = vs_mallocf(nl);
  if (out == nullptr) return nullptr;
  memcpy(out, ptr, length * sizeof(float));
  memset(out + length, 0, (nl - length) * sizeof(float));
  *new_length = nl;
  return out;
}

float* vs_zeropadding(const float* ptr, size_t length, size_t* new_length) {
  return vs_zeropaddingex(ptr, length, new_length, 0);
}

float* vs_rmemcpyf(float* dest, const float* src, size_t length) {
  for (size_t i = 0; i < length; ++i) dest[i] = src[length - 1 - i];
  return dest;
}

float* vs_crmemcpyf(float* dest, const float* src, size_t length) {
  // length floats = length / 2 complex numbers; the pairs are reversed, each pair keeps its (re, im) order
  const size_t n = length / 2;
  for (size_t i = 0; i < n; ++i) {
    dest[2 * i] = src[2 * (n - 1 - i)];
    dest[2 * i + 1] = src[2 * (n - 1 - i) + 1];
  }
  return dest;
}

// ---- conversions ------------------------------------------------------------------------------------------------
void vs_int16_to_float(const int16_t* data, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = static_cast<float>(data[i]);
}
void vs_float_to_int16(const float* data, size_t length, int16_t* res) {
  for (size_t i = 0; i < length; ++i) res[i] = static_cast<int16_t>(data[i]);
}
void vs_int32_to_float(const int32_t* data, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = static_cast<float>(data[i]);
}
void vs_float_to_int32(const float* data, size_t length, int32_t* res) {
  for (size_t i = 0; i < length; ++i) res[i] = static_cast<int32_t>(data[i]);
}
void vs_int16_to_int32(const int16_t* data, size_t length, int32_t* res) {
  for (size_t i = 0; i < length; ++i) res[i] = data[i];
}
void vs_int32_to_int16(const int32_t* data, size_t length, int16_t* res) {
  for (size_t i = 0; i < length; ++i) res[i] = static_cast<int16_t>(data[i]);
}

void vs_float16_to_float(const uint16_t* data, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) {
    const uint32_t h = data[i];
    const uint32_t sign = (h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu;
    uint32_t man = h & 0x3ffu;
    uint32_t bits;
    if (exp == 0) {
      if (man == 0) {
        bits = sign;
      } else {  // subnormal half -> normal float
        int e = -1;
        do {
          ++e;
          man <<= 1;
        } while ((man & 0x400u) == 0);
        bits = sign | (static_cast<uint32_t>(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
      }
    } else if (exp == 31) {
      bits = sign | 0x7f800000u | (man << 13);
    } else {
      bits = sign | ((exp + 127 - 15) << 23) | (man << 13);
    }
    memcpy(&res[i], &bits, sizeof(float));
  }
}

void vs_int16_multiply(const int16_t* a, const int16_t* b, size_t length, int32_t* res) {
  for (size_t i = 0; i < length; ++i) res[i] = static_cast<int32_t>(a[i]) * static_cast<int32_t>(b[i]);
}

// ---- element-wise -------------------------------------------------------------------------------------------------
void vs_real_multiply_array(const float* a, const float* b, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = a[i] * b[i];
}
void vs_real_multiply_scalar(const float* array, size_t length, float value, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = array[i] * value;
}
void vs_complex_multiply(const float* a, const float* b, size_t length, float* res) {
  for (size_t i = 0; i + 1 < length; i += 2) {
    const float re1 = a[i], im1 = a[i + 1], re2 = b[i], im2 = b[i + 1];
    res[i] = re1 * re2 - im1 * im2;
    res[i + 1] = re1 * im2 + re2 * im1;
  }
}
void vs_complex_multiply_conjugate(const float* a, const float* b, size_t length, float* res) {
  for (size_t i = 0; i + 1 < length; i += 2) {
    const float re1 = a[i], im1 = a[i + 1], re2 = b[i], im2 = -b[i + 1];
    res[i] = re1 * re2 - im1 * im2;
    res[i + 1] = re1 * im2 + re2 * im1;
  }
}
void vs_complex_conjugate(const float* array, size_t length, float* res) {
  for (size_t i = 0; i + 1 < length; i += 2) {
    res[i] = array[i];
    res[i + 1] = -array[i + 1];
  }
}
float vs_sum_elements(const float* input, size_t length) {
  float s = 0.f;
  for (size_t i = 0; i < length; ++i) s += input[i];
  return s;
}
void vs_add_to_all(float* input, size_t length, float value) {
  for (size_t i = 0; i < length; ++i) input[i] += value;
}

void vs_sin_psv(const float* src, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = sinf(src[i]);
}
void vs_cos_psv(const float* src, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = cosf(src[i]);
}
void vs_log_psv(const float* src, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = logf(src[i]);
}
void vs_exp_psv(const float* src, size_t length, float* res) {
  for (size_t i = 0; i < length; ++i) res[i] = expf(src[i]);
}

// ---- matrices -----------------------------------------------------------------------------------------------------
void vs_matrix_add(const float* m1, const float* m2, size_t w, size_t h, float* res) {
  const size_t n = w * h;
  for (size_t i = 0; i < n; ++i) res[i] = m1[i] + m2[i];
}
void vs_matrix_sub(const float* m1, const float* m2, size_t w, size_t h, float* res) {
  const size_t n = w * h;
  for (size_t i = 0; i < n; ++i) res[i] = m1[i] - m2[i];
}

void vs_matrix_multiply(const float* m1, const float* m2, size_t w1, size_t h1, size_t w2, size_t h2, float* res) {
  (void)h2;  // == w1
  // i-k-j order: the inner loop streams a row of m2 and a row of res (unit stride, vectorises)
  for (size_t i = 0; i < h1; ++i) {
    float* out = res + i * w2;
    for (size_t j = 0; j < w2; ++j) out[j] = 0.f;
    for (size_t k = 0; k < w1; ++k) {
      const float a = m1[i * w1 + k];
      const float* row = m2 + k * w2;
      for (size_t j = 0; j < w2; ++j) out[j] += a * row[j];
    }
  }
}

void vs_matrix_multiply_transposed(const float* m1, const float* m2, size_t w1, size_t h1, size_t w2, size_t h2,
                                   float* res) {
  (void)w2;  // == w1
  for (size_t i = 0; i < h1; ++i) {
    const float* a = m1 + i * w1;
    for (size_t j = 0; j < h2; ++j) {
      const float* b = m2 + j * w1;
      // eight independent partial sums: strict IEEE ordering would otherwise keep this dot product scalar
      float p[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      size_t k = 0;
      for (; k + 8 <= w1; k += 8)
        for (int u = 0; u < 8; ++u) p[u] += a[k + u] * b[k + u];
      float s = ((p[0] + p[1]) + (p[2] + p[3])) + ((p[4] + p[5]) + (p[6] + p[7]));
      for (; k < w1; ++k) s += a[k] * b[k];
      res[i * h2 + j] = s;
    }
  }
}

}  // extern "C"
