// Host signal processing: FFT, convolution / cross-correlation (brute force, one-shot FFT, overlap-save),
// 8-bit plane normalisation, min/max, peak detection.  Behavioural counterpart of the reference's libSimd
// convolve.c / correlate.c / normalize.c / detect_peaks.c (the reference leans on the external FFTF library for its
// transforms; the radix-2 transform here is self-contained).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "veles/simd.h"

namespace {

// In-place iterative radix-2 FFT on interleaved (re, im) floats; twiddles evaluated in double.
void fft_inplace(float* d, size_t n, int direction) {
  for (size_t i = 1, j = 0; i < n; ++i) {
    size_t bit = n >> 1;
    for (; j & bit; bit >>= 1) j ^= bit;
    j ^= bit;
    if (i < j) {
      std::swap(d[2 * i], d[2 * j]);
      std::swap(d[2 * i + 1], d[2 * j + 1]);
    }
  }
  const double sign = direction >= 0 ? -1.0 : 1.0;
  for (size_t len = 2; len <= n; len <<= 1) {
    const double ang = sign * 2.0 * M_PI / static_cast<double>(len);
    const size_t half = len >> 1;
    std::vector<float> tw(2 * half);
    for (size_t k = 0; k < half; ++k) {
      tw[2 * k] = static_cast<float>(std::cos(ang * static_cast<double>(k)));
      tw[2 * k + 1] = static_cast<float>(std::sin(ang * static_cast<double>(k)));
    }
    for (size_t i = 0; i < n; i += len) {
      for (size_t k = 0; k < half; ++k) {
        const float wr = tw[2 * k], wi = tw[2 * k + 1];
        float* a = d + 2 * (i + k);
        float* b = d + 2 * (i + k + half);
        const float tr = b[0] * wr - b[1] * wi;
        const float ti = b[0] * wi + b[1] * wr;
        b[0] = a[0] - tr;
        b[1] = a[1] - ti;
        a[0] += tr;
        a[1] += ti;
      }
    }
  }
}

bool is_pow2(size_t n) { return n != 0 && (n & (n - 1)) == 0; }

}  // namespace

struct VsConvolution {
  int algorithm;
  int reverse;        // cross-correlation: h is applied back to front
  size_t x_length;
  size_t h_length;
  size_t fft_size;    // one-shot FFT: M;  overlap-save: L
  std::vector<float> H;      // spectrum of h (interleaved), fft_size complex
  std::vector<float> work;   // fft_size complex
};

extern "C" {

int vs_fft_complex(float* data, size_t n, int direction) {
  if (!is_pow2(n)) return -1;
  fft_inplace(data, n, direction);
  return 0;
}

void vs_convolve_simd(const float* x, size_t x_length, const float* h, size_t h_length, float* result) {
  const long xl = static_cast<long>(x_length), hl = static_cast<long>(h_length);
  for (long n = 0; n < xl + hl - 1; ++n) {
    const long beg = n < xl ? 0 : n - xl + 1;
    const long end = std::min(n + 1, hl);
    float sum = 0.f;
    for (long m = beg; m < end; ++m) sum += h[m] * x[n - m];
    result[n] = sum;
  }
}

void vs_cross_correlate_simd(const float* x, size_t x_length, const float* h, size_t h_length, float* result) {
  const long xl = static_cast<long>(x_length), hl = static_cast<long>(h_length);
  // result[k] = sum_m x[m] * h[m + (hl - 1 - k)]   (numpy.correlate(x, h, "full"))
  for (long k = 0; k < xl + hl - 1; ++k) {
    const long n = hl - 1 - k;
    const long beg = n <= 0 ? -n : 0;
    const long end = std::min(hl - n, xl);
    float sum = 0.f;
    for (long m = beg; m < end; ++m) sum += x[m] * h[n + m];
    result[k] = sum;
  }
}

int vs_convolve_choose(size_t x_length, size_t h_length) {
  // the reference's heuristic (src/convolve.c:347-363): long signal with a short filter -> overlap-save,
  // comparable lengths -> one FFT, small problems -> direct sums
  if (x_length > h_length * 2) return x_length > 200 ? VS_CONV_OVERLAP_SAVE : VS_CONV_BRUTE;
  return x_length > 350 ? VS_CONV_FFT : VS_CONV_BRUTE;
}

static VsConvolution* make_handle(size_t x_length, size_t h_length, int algorithm, int reverse) {
  if (x_length == 0 || h_length == 0) return nullptr;
  if (algorithm == VS_CONV_OVERLAP_SAVE && !(h_length < x_length / 2)) algorithm = VS_CONV_FFT;
  VsConvolution* c = new VsConvolution();
  c->algorithm = algorithm;
  c->reverse = reverse;
  c->x_length = x_length;
  c->h_length = h_length;
  c->fft_size = 0;
  if (algorithm == VS_CONV_FFT) {
    c->fft_size = vs_next_highest_power_of_2(x_length + h_length - 1);
  } else if (algorithm == VS_CONV_OVERLAP_SAVE) {
    // block of 4 * 2^floor(log2 M) points: at least 2 M, so every block yields more than M new outputs
    size_t L = h_length;
    int log = 2;
    while (L >>= 1) ++log;
    c->fft_size = static_cast<size_t>(1) << log;
  }
  if (c->fft_size) {
    c->H.assign(2 * c->fft_size, 0.f);
    c->work.assign(2 * c->fft_size, 0.f);
  }
  return c;
}

VsConvolution* vs_convolve_initialize_algorithm(size_t x_length, size_t h_length, int algorithm) {
  return make_handle(x_length, h_length, algorithm, 0);
}
VsConvolution* vs_convolve_initialize(size_t x_length, size_t h_length) {
  return make_handle(x_length, h_length, vs_convolve_choose(x_length, h_length), 0);
}
VsConvolution* vs_cross_correlate_initialize_algorithm(size_t x_length, size_t h_length, int algorithm) {
  return make_handle(x_length, h_length, algorithm, 1);
}
VsConvolution* vs_cross_correlate_initialize(size_t x_length, size_t h_length) {
  return make_handle(x_length, h_length, vs_convolve_choose(x_length, h_length), 1);
}
int vs_convolve_algorithm(const VsConvolution* handle) { return handle ? handle->algorithm : -1; }
void vs_convolve_finalize(VsConvolution* handle) { delete handle; }

static void load_filter_spectrum(VsConvolution* c, const float* h) {
  std::fill(c->H.begin(), c->H.end(), 0.f);
  const size_t M = c->h_length;
  for (size_t i = 0; i < M; ++i) c->H[2 * i] = c->reverse ? h[M - 1 - i] : h[i];
  fft_inplace(c->H.data(), c->fft_size, +1);
}

static void run_fft(VsConvolution* c, const float* x, const float* h, float* result) {
  const size_t N = c->fft_size, out_len = c->x_length + c->h_length - 1;
  load_filter_spectrum(c, h);
  float* w = c->work.data();
  std::fill(c->work.begin(), c->work.end(), 0.f);
  for (size_t i = 0; i < c->x_length; ++i) w[2 * i] = x[i];
  fft_inplace(w, N, +1);
  vs_complex_multiply(w, c->H.data(), 2 * N, w);
  fft_inplace(w, N, -1);
  const float scale = 1.f / static_cast<float>(N);
  for (size_t i = 0; i < out_len; ++i) result[i] = w[2 * i] * scale;
}

static void run_overlap_save(VsConvolution* c, const float* x, const float* h, float* result) {
  const size_t L = c->fft_size, M = c->h_length, X = c->x_length;
  const size_t out_len = X + M - 1, step = L - (M - 1);
  load_filter_spectrum(c, h);
  float* w = c->work.data();
  const float scale = 1.f / static_cast<float>(L);
  for (size_t out0 = 0; out0 < out_len; out0 += step) {
    // block covers signal positions [out0 - (M - 1), out0 - (M - 1) + L); outside the signal it reads zeros
    const long base = static_cast<long>(out0) - static_cast<long>(M - 1);
    for (size_t i = 0; i < L; ++i) {
      const long p = base + static_cast<long>(i);
      w[2 * i] = (p >= 0 && p < static_cast<long>(X)) ? x[p] : 0.f;
      w[2 * i + 1] = 0.f;
    }
    fft_inplace(w, L, +1);
    vs_complex_multiply(w, c->H.data(), 2 * L, w);
    fft_inplace(w, L, -1);
    const size_t n = std::min(step, out_len - out0);
    for (size_t i = 0; i < n; ++i) result[out0 + i] = w[2 * (M - 1 + i)] * scale;   // the alias-free tail
  }
}

static void run(VsConvolution* c, const float* x, const float* h, float* result) {
  switch (c->algorithm) {
    case VS_CONV_FFT: run_fft(c, x, h, result); break;
    case VS_CONV_OVERLAP_SAVE: run_overlap_save(c, x, h, result); break;
    default:
      if (c->reverse) vs_cross_correlate_simd(x, c->x_length, h, c->h_length, result);
      else vs_convolve_simd(x, c->x_length, h, c->h_length, result);
  }
}

void vs_convolve(VsConvolution* handle, const float* x, const float* h, float* result) { run(handle, x, h, result); }
void vs_cross_correlate(VsConvolution* handle, const float* x, const float* h, float* result) {
  run(handle, x, h, result);
}

// ---- normalisation ------------------------------------------------------------------------------------------------
void vs_minmax2D(const uint8_t* src, int src_stride, int width, int height, uint8_t* min, uint8_t* max) {
  uint8_t lo = 255, hi = 0;
  for (int y = 0; y < height; ++y) {
    const uint8_t* row = src + static_cast<size_t>(y) * src_stride;
    for (int x = 0; x < width; ++x) {
      lo = std::min(lo, row[x]);
      hi = std::max(hi, row[x]);
    }
  }
  if (min) *min = lo;
  if (max) *max = hi;
}

void vs_normalize2D_minmax(uint8_t min, uint8_t max, const uint8_t* src, int src_stride, int width, int height,
                           float* dst, int dst_stride) {
  if (max == min) {   // a flat plane carries no contrast: all zeros
    for (int y = 0; y < height; ++y) vs_memsetf(dst + static_cast<size_t>(y) * dst_stride, 0.f, width);
    return;
  }
  const float diff = (static_cast<float>(max) - static_cast<float>(min)) / 2.f;
  for (int y = 0; y < height; ++y) {
    const uint8_t* s = src + static_cast<size_t>(y) * src_stride;
    float* d = dst + static_cast<size_t>(y) * dst_stride;
    for (int x = 0; x < width; ++x) d[x] = (static_cast<float>(s[x]) - static_cast<float>(min)) / diff - 1.0f;
  }
}

void vs_normalize2D(const uint8_t* src, int src_stride, int width, int height, float* dst, int dst_stride) {
  uint8_t lo, hi;
  vs_minmax2D(src, src_stride, width, height, &lo, &hi);
  vs_normalize2D_minmax(lo, hi, src, src_stride, width, height, dst, dst_stride);
}

void vs_minmax1D(const float* src, int length, float* min, float* max) {
  float lo = src[0], hi = src[0];
  for (int i = 1; i < length; ++i) {
    lo = std::min(lo, src[i]);
    hi = std::max(hi, src[i]);
  }
  if (min) *min = lo;
  if (max) *max = hi;
}

// ---- peaks --------------------------------------------------------------------------------------------------------
void vs_detect_peaks(const float* data, size_t size, int type, VsExtremumPoint** results, size_t* results_length) {
  *results = nullptr;
  *results_length = 0;
  if (size < 3) return;
  std::vector<VsExtremumPoint> found;
  for (size_t i = 1; i + 1 < size; ++i) {
    const float d1 = data[i] - data[i - 1];
    const float d2 = data[i] - data[i + 1];
    if (d1 * d2 > 0) {
      if ((d1 > 0 && (type & VS_EXTREMUM_MAXIMUM)) || (d1 < 0 && (type & VS_EXTREMUM_MINIMUM))) {
        found.push_back(VsExtremumPoint{static_cast<int>(i), data[i]});
      }
    }
  }
  if (found.empty()) return;
  auto* out = static_cast<VsExtremumPoint*>(vs_malloc_aligned(found.size() * sizeof(VsExtremumPoint)));
  memcpy(out, found.data(), found.size() * sizeof(VsExtremumPoint));
  *results = out;
  *results_length = found.size();
}

}  // extern "C"
