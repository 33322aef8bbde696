// veles_b200 native signal / array primitives (host side).
//
// Equivalent of the reference's libSimd (libVeles/simd/inc/simd/{memory,arithmetic,mathfun,matrix,convolve,
// correlate,normalize,wavelet,detect_peaks}.h): the C API the SoundFeatureExtraction-style audio path and the
// package loader lean on.  Written as plain C++17 loops the compiler vectorises (no per-ISA intrinsics: the device
// versions in veles_b200/csrc/signal.cu are the fast path on a B200 node; these are the host oracle and the
// fallback for small inputs).  Every function is extern "C" and thread-safe unless it takes a handle.
#pragma once
#include <cstddef>
#include <cstdint>

extern "C" {

// ---- memory (reference: inc/simd/memory.h:51-178) -------------------------------------------------------------
void* vs_malloc_aligned(size_t bytes);                       // 64-byte aligned, release with vs_free
float* vs_mallocf(size_t length);
void vs_free(void* ptr);
int vs_align_complement_f32(const float* ptr);               // elements until the next 64-byte boundary
void vs_memsetf(float* ptr, float value, size_t length);
size_t vs_next_highest_power_of_2(size_t value);
// new block of 2 * (smallest power of two >= length) (+ additional) floats, the tail zero-filled
float* vs_zeropadding(const float* ptr, size_t length, size_t* new_length);
float* vs_zeropaddingex(const float* ptr, size_t length, size_t* new_length, size_t additional_length);
float* vs_rmemcpyf(float* dest, const float* src, size_t length);    // reversed copy
float* vs_crmemcpyf(float* dest, const float* src, size_t length);   // reversed copy of (re, im) pairs

// ---- arithmetic (reference: inc/simd/arithmetic.h) -------------------------------------------------------------
void vs_int16_to_float(const int16_t* data, size_t length, float* res);
void vs_float_to_int16(const float* data, size_t length, int16_t* res);     // truncation, like the reference
void vs_int32_to_float(const int32_t* data, size_t length, float* res);
void vs_float_to_int32(const float* data, size_t length, int32_t* res);
void vs_int16_to_int32(const int16_t* data, size_t length, int32_t* res);
void vs_int32_to_int16(const int32_t* data, size_t length, int16_t* res);
void vs_float16_to_float(const uint16_t* data, size_t length, float* res);
void vs_int16_multiply(const int16_t* a, const int16_t* b, size_t length, int32_t* res);
void vs_real_multiply_array(const float* a, const float* b, size_t length, float* res);
void vs_real_multiply_scalar(const float* array, size_t length, float value, float* res);
// length counts floats (2 per complex number), interleaved (re, im)
void vs_complex_multiply(const float* a, const float* b, size_t length, float* res);
void vs_complex_multiply_conjugate(const float* a, const float* b, size_t length, float* res);
void vs_complex_conjugate(const float* array, size_t length, float* res);
float vs_sum_elements(const float* input, size_t length);
void vs_add_to_all(float* input, size_t length, float value);

// ---- mathfun (reference: inc/simd/mathfun.h:142-205) -----------------------------------------------------------
void vs_sin_psv(const float* src, size_t length, float* res);
void vs_cos_psv(const float* src, size_t length, float* res);
void vs_log_psv(const float* src, size_t length, float* res);
void vs_exp_psv(const float* src, size_t length, float* res);

// ---- matrix (reference: inc/simd/matrix.h:47-87); row-major, w = columns, h = rows -----------------------------
void vs_matrix_add(const float* m1, const float* m2, size_t w, size_t h, float* res);
void vs_matrix_sub(const float* m1, const float* m2, size_t w, size_t h, float* res);
// res[h1 x w2] = m1[h1 x w1] * m2[h2 x w2], w1 == h2
void vs_matrix_multiply(const float* m1, const float* m2, size_t w1, size_t h1, size_t w2, size_t h2, float* res);
// res[h1 x h2] = m1[h1 x w1] * m2[h2 x w2]^T, w1 == w2
void vs_matrix_multiply_transposed(const float* m1, const float* m2, size_t w1, size_t h1, size_t w2, size_t h2,
                                   float* res);

// ---- convolution / cross-correlation (reference: inc/simd/convolve.h:57-125, correlate.h:58-135) ---------------
// result has x_length + h_length - 1 elements.  Algorithms: 0 brute force, 1 one-shot FFT, 2 overlap-save.
enum { VS_CONV_BRUTE = 0, VS_CONV_FFT = 1, VS_CONV_OVERLAP_SAVE = 2 };
typedef struct VsConvolution VsConvolution;
void vs_convolve_simd(const float* x, size_t x_length, const float* h, size_t h_length, float* result);
void vs_cross_correlate_simd(const float* x, size_t x_length, const float* h, size_t h_length, float* result);
int vs_convolve_choose(size_t x_length, size_t h_length);           // the reference's size heuristic
VsConvolution* vs_convolve_initialize(size_t x_length, size_t h_length);
VsConvolution* vs_convolve_initialize_algorithm(size_t x_length, size_t h_length, int algorithm);
VsConvolution* vs_cross_correlate_initialize(size_t x_length, size_t h_length);
VsConvolution* vs_cross_correlate_initialize_algorithm(size_t x_length, size_t h_length, int algorithm);
int vs_convolve_algorithm(const VsConvolution* handle);
void vs_convolve(VsConvolution* handle, const float* x, const float* h, float* result);
void vs_cross_correlate(VsConvolution* handle, const float* x, const float* h, float* result);
void vs_convolve_finalize(VsConvolution* handle);
// in-place complex FFT of n = 2^k points on interleaved data (direction +1 forward, -1 inverse unscaled)
int vs_fft_complex(float* data, size_t n, int direction);

// ---- normalisation (reference: inc/simd/normalize.h:48-89) -----------------------------------------------------
void vs_minmax2D(const uint8_t* src, int src_stride, int width, int height, uint8_t* min, uint8_t* max);
void vs_normalize2D_minmax(uint8_t min, uint8_t max, const uint8_t* src, int src_stride, int width, int height,
                           float* dst, int dst_stride);             // (v - min) / ((max - min) / 2) - 1; flat -> 0
void vs_normalize2D(const uint8_t* src, int src_stride, int width, int height, float* dst, int dst_stride);
void vs_minmax1D(const float* src, int length, float* min, float* max);

// ---- wavelets (reference: inc/simd/wavelet.h:45-157, wavelet_types.h) ------------------------------------------
enum { VS_WAVELET_DAUBECHIES = 0, VS_WAVELET_COIFLET = 1, VS_WAVELET_SYMLET = 2 };
enum { VS_EXT_PERIODIC = 0, VS_EXT_MIRROR = 1, VS_EXT_CONSTANT = 2, VS_EXT_ZERO = 3 };
int vs_wavelet_validate_order(int type, int order);
// lowpass / highpass analysis taps as the transform applies them.  reference_scale != 0 reproduces the reference's
// tables (Daubechies sum sqrt 2, coiflets and symlets sum 1); 0 gives orthonormal filters for every family.
int vs_wavelet_filters(int type, int order, int reference_scale, float* lowpass, float* highpass);
float* vs_wavelet_prepare_array(int order, const float* src, size_t length);            // a copy (no alignment games)
float* vs_wavelet_allocate_destination(int order, size_t source_length);
void vs_wavelet_recycle_source(int order, float* src, size_t length, float** desthihi, float** desthilo,
                               float** destlohi, float** destlolo);
// desthi / destlo get length / 2 elements each: out[d] = sum_j taps[j] * ext(src)[2 d + j]
void vs_wavelet_apply(int type, int order, int ext, const float* src, size_t length, float* desthi, float* destlo);
// undecimated transform at `level` (taps dilated by 2^(level-1)); desthi / destlo get `length` elements
void vs_stationary_wavelet_apply(int type, int order, int level, int ext, const float* src, size_t length,
                                 float* desthi, float* destlo);
void vs_wavelet_apply_ex(int type, int order, int ext, int reference_scale, const float* src, size_t length,
                         float* desthi, float* destlo);
void vs_stationary_wavelet_apply_ex(int type, int order, int level, int ext, int reference_scale, const float* src,
                                    size_t length, float* desthi, float* destlo);

// ---- peaks (reference: inc/simd/detect_peaks.h) ----------------------------------------------------------------
enum { VS_EXTREMUM_MAXIMUM = 1, VS_EXTREMUM_MINIMUM = 2, VS_EXTREMUM_BOTH = 3 };
typedef struct {
  int position;
  float value;
} VsExtremumPoint;
// strict interior extrema in index order; *results is vs_malloc_aligned memory (vs_free), nullptr when none
void vs_detect_peaks(const float* data, size_t size, int type, VsExtremumPoint** results, size_t* results_length);

}  // extern "C"
