// Host discrete wavelet transforms (decimated and stationary) over the Daubechies / coiflet / symlet filter banks of
// wavelet_tables.h (computed by scripts/gen_wavelet_tables.py).  Behavioural counterpart of the reference's
// libSimd wavelet.c: one analysis step splits a signal into a highpass and a lowpass half,
//     out[d] = sum_j taps[j] * ext(src)[2 d + j],
// with the signal continued on the right by one of four rules.  The reference unrolls per-order SSE/AVX/NEON
// variants; the filter length is a run-time loop here and the fast path is the batched device kernel.
#include <cmath>
#include <cstring>
#include <vector>

#include "veles/simd.h"
#include "veles/wavelet_tables.h"

namespace {

const veles::wavelets::Filter* find_filter(int type, int order) {
  for (int i = 0; i < veles::wavelets::kFilterCount; ++i) {
    const veles::wavelets::Filter& f = veles::wavelets::kFilters[i];
    if (f.type == type && f.order == order) return &f;
  }
  return nullptr;
}

inline float extended(const float* src, long length, long index, int ext) {
  if (index < length) return src[index];
  const long i = index - length;
  switch (ext) {
    case VS_EXT_PERIODIC: return src[i % length];
    case VS_EXT_MIRROR: return src[length - 1 - (i % length)];
    case VS_EXT_CONSTANT: return src[length - 1];
    default: return 0.f;
  }
}

}  // namespace

extern "C" {

int vs_wavelet_validate_order(int type, int order) { return find_filter(type, order) != nullptr ? 1 : 0; }

int vs_wavelet_filters(int type, int order, int reference_scale, float* lowpass, float* highpass) {
  const veles::wavelets::Filter* f = find_filter(type, order);
  if (f == nullptr) return -1;
  // the tables are orthonormal (sum sqrt 2); the reference's coiflet and symlet tables are scaled to sum 1
  const double scale = (reference_scale && type != VS_WAVELET_DAUBECHIES) ? 1.0 / std::sqrt(2.0) : 1.0;
  for (int i = 0; i < order; ++i) {
    const float val = static_cast<float>(f->taps[i] * scale);
    if (lowpass) lowpass[i] = val;
    if (highpass) highpass[order - 1 - i] = (i & 1) ? val : -val;   // quadrature mirror of the lowpass
  }
  return 0;
}

float* vs_wavelet_prepare_array(int /*order*/, const float* src, size_t length) {
  float* out = vs_mallocf(length);
  if (out) memcpy(out, src, length * sizeof(float));
  return out;
}

float* vs_wavelet_allocate_destination(int /*order*/, size_t source_length) { return vs_mallocf(source_length / 2); }

void vs_wavelet_recycle_source(int /*order*/, float* src, size_t length, float** desthihi, float** desthilo,
                               float** destlohi, float** destlolo) {
  const size_t lq = length / 4;
  *desthihi = src;
  *desthilo = src + lq;
  *destlohi = src + 2 * lq;
  *destlolo = src + 3 * lq;
}

// Both transforms run tap-outermost over an explicitly extended copy of the signal: every tap then is one
// contiguous multiply-add sweep over the outputs (the compiler vectorises it), and the border needs no special case.
// For the decimated transform the extended signal is first split into its even and odd samples, which turns the
// stride-2 reads into unit-stride ones.
static void extend_signal(const float* src, long n, long extra, int ext, std::vector<float>* out) {
  out->resize(static_cast<size_t>(n + extra));
  memcpy(out->data(), src, static_cast<size_t>(n) * sizeof(float));
  for (long i = 0; i < extra; ++i) (*out)[n + i] = extended(src, n, n + i, ext);
}

void vs_wavelet_apply_ex(int type, int order, int ext, int reference_scale, const float* src, size_t length,
                         float* desthi, float* destlo) {
  std::vector<float> lo(order), hi(order);
  if (vs_wavelet_filters(type, order, reference_scale, lo.data(), hi.data()) != 0) return;
  const long n = static_cast<long>(length), half = n / 2;
  if (order <= 16) {
    // short filters: one pass over the signal with both dot products in registers beats the extra copies below
    long d = 0;
    for (; 2 * d + order <= n; ++d) {
      const float* s = src + 2 * d;
      float rh = 0.f, rl = 0.f;
      for (int j = 0; j < order; ++j) {
        rh += hi[j] * s[j];
        rl += lo[j] * s[j];
      }
      desthi[d] = rh;
      destlo[d] = rl;
    }
    for (; d < half; ++d) {
      float rh = 0.f, rl = 0.f;
      for (int j = 0; j < order; ++j) {
        const float v = extended(src, n, 2 * d + j, ext);
        rh += hi[j] * v;
        rl += lo[j] * v;
      }
      desthi[d] = rh;
      destlo[d] = rl;
    }
    return;
  }
  std::vector<float> full;
  extend_signal(src, n, order, ext, &full);
  const long m = (n + order + 1) / 2;
  std::vector<float> even(m + 1, 0.f), odd(m + 1, 0.f);
  for (long k = 0; 2 * k < n + order; ++k) even[k] = full[2 * k];
  for (long k = 0; 2 * k + 1 < n + order; ++k) odd[k] = full[2 * k + 1];
  constexpr long kBlock = 2048;     // outputs per block: the two accumulator rows stay in L1 across the tap sweeps
  for (long d0 = 0; d0 < half; d0 += kBlock) {
    const long nb = (half - d0 < kBlock) ? half - d0 : kBlock;
    float* ph = desthi + d0;
    float* pl = destlo + d0;
    for (long d = 0; d < nb; ++d) ph[d] = pl[d] = 0.f;
    for (int j = 0; j < order; ++j) {
      // ext[2 d + j] = even[d + j / 2] (j even) or odd[d + (j - 1) / 2] (j odd)
      const float* s = ((j & 1) ? odd.data() + (j - 1) / 2 : even.data() + j / 2) + d0;
      const float th = hi[j], tl = lo[j];
      for (long d = 0; d < nb; ++d) {
        ph[d] += th * s[d];
        pl[d] += tl * s[d];
      }
    }
  }
}

void vs_wavelet_apply(int type, int order, int ext, const float* src, size_t length, float* desthi, float* destlo) {
  vs_wavelet_apply_ex(type, order, ext, 1, src, length, desthi, destlo);
}

void vs_stationary_wavelet_apply_ex(int type, int order, int level, int ext, int reference_scale, const float* src,
                                    size_t length, float* desthi, float* destlo) {
  std::vector<float> lo(order), hi(order);
  if (vs_wavelet_filters(type, order, reference_scale, lo.data(), hi.data()) != 0 || level < 1) return;
  const long stride = 1L << (level - 1);     // taps are spread `stride` samples apart ("a trous")
  const long n = static_cast<long>(length);
  std::vector<float> full;
  extend_signal(src, n, static_cast<long>(order) * stride, ext, &full);
  constexpr long kBlock = 2048;
  for (long i0 = 0; i0 < n; i0 += kBlock) {
    const long nb = (n - i0 < kBlock) ? n - i0 : kBlock;
    float* ph = desthi + i0;
    float* pl = destlo + i0;
    for (long i = 0; i < nb; ++i) ph[i] = pl[i] = 0.f;
    for (int j = 0; j < order; ++j) {
      const float* s = full.data() + j * stride + i0;
      const float th = hi[j], tl = lo[j];
      for (long i = 0; i < nb; ++i) {
        ph[i] += th * s[i];
        pl[i] += tl * s[i];
      }
    }
  }
}

void vs_stationary_wavelet_apply(int type, int order, int level, int ext, const float* src, size_t length,
                                 float* desthi, float* destlo) {
  vs_stationary_wavelet_apply_ex(type, order, level, ext, 1, src, length, desthi, destlo);
}

}  // extern "C"
