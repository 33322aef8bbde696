// veles_b200 — batched 1-D signal kernels: direct convolution / cross-correlation, decimated and stationary wavelet
// analysis steps, min/max + 8-bit plane normalisation, peak detection, spectrum products, sample-format conversion.
//
// Device counterpart of the reference's CPU SIMD library (libVeles/simd/src/{convolve,correlate,wavelet,normalize,
// detect_peaks}.c, inc/simd/arithmetic.h), which served one signal per call on SSE/AVX/NEON lanes.  Here every
// entry point takes `batch` equally long rows, so a whole minibatch of audio frames / image planes is one launch,
// and the data stays in HBM between feature-extraction stages.  None of this is GEMM-shaped (filters are 2..76 taps,
// arithmetic intensity < 40 flop/byte only for the longest ones): the kernels are organised around coalesced
// streaming and shared-memory reuse of the sliding window, not tensor cores.
#include "common.cuh"

namespace vb {

// ---------------------------------------------------------------------------------------------------------------
// Direct convolution:  out[n] = sum_m h'[m] * x[n - m],  n in [0, X + M - 1),  h' = h or h reversed (correlation).
// A block owns CONV_TILE consecutive outputs of one row; taps are consumed in chunks of CONV_HC held in shared
// memory next to the matching window of x (zero-filled outside the signal).  Thread t computes CONV_ITEMS
// CONSECUTIVE outputs: 8 taps x 8 outputs need only 15 window samples, which live in registers, so a step is
// 15 + 2 shared loads for 64 FMAs.  The window is stored with one pad word per 32 (index j + j / 32): the
// stride-8 accesses of neighbouring threads then fall into distinct banks.
// ---------------------------------------------------------------------------------------------------------------
constexpr int CONV_THREADS = 256;
constexpr int CONV_ITEMS = 8;
constexpr int CONV_TILE = CONV_THREADS * CONV_ITEMS;
constexpr int CONV_HC = 256;
constexpr int CONV_XS = CONV_TILE + CONV_HC;

__device__ __forceinline__ int conv_pad(int j) { return j + (j >> 5); }

__global__ void __launch_bounds__(CONV_THREADS)
sig_conv_kernel(const float* __restrict__ x, long long x_pitch, int X, const float* __restrict__ h, long long h_pitch,
                int M, int reverse, float* __restrict__ out, long long out_pitch) {
  __shared__ __align__(16) float h_s[CONV_HC];
  __shared__ float x_s[CONV_XS + CONV_XS / 32 + 1];
  const int row = blockIdx.y;
  const int n0 = blockIdx.x * CONV_TILE;
  const int out_len = X + M - 1;
  const float* xr = x + row * x_pitch;
  const float* hr = h + row * h_pitch;
  float acc[CONV_ITEMS];
#pragma unroll
  for (int i = 0; i < CONV_ITEMS; ++i) acc[i] = 0.f;
  // only taps that can meet the signal for this tile:  n - m in [0, X)  ->  m in [n0 - X + 1, n0 + TILE - 1]
  int m_begin = n0 - X + 1;
  if (m_begin < 0) m_begin = 0;
  int m_end = n0 + CONV_TILE;
  if (m_end > M) m_end = M;
  for (int m0 = m_begin; m0 < m_end; m0 += CONV_HC) {
    const int hc = min(CONV_HC, m_end - m0);
    for (int i = threadIdx.x; i < CONV_HC; i += CONV_THREADS) {
      const int m = m0 + i;
      h_s[i] = (i < hc) ? hr[reverse ? (M - 1 - m) : m] : 0.f;
    }
    // window sample j = x[n0 - m0 - (CONV_HC - 1) + j]
    const int xbase = n0 - m0 - (CONV_HC - 1);
    for (int j = threadIdx.x; j < CONV_XS; j += CONV_THREADS) {
      const int p = xbase + j;
      x_s[conv_pad(j)] = (p >= 0 && p < X) ? xr[p] : 0.f;
    }
    __syncthreads();
    const int hc8 = (hc + 7) & ~7;     // the tail of h_s is zero
    for (int i0 = 0; i0 < hc8; i0 += 8) {
      // output k, tap i0 + i reads window sample  t*8 + k - i0 - i + HC - 1  =  base + (k - i + 7)
      const int base = threadIdx.x * CONV_ITEMS - i0 + (CONV_HC - 1) - 7;
      float v[15];
#pragma unroll
      for (int j = 0; j < 15; ++j) v[j] = x_s[conv_pad(base + j)];
      const float4 ha = *reinterpret_cast<const float4*>(h_s + i0);
      const float4 hb = *reinterpret_cast<const float4*>(h_s + i0 + 4);
      const float taps[8] = {ha.x, ha.y, ha.z, ha.w, hb.x, hb.y, hb.z, hb.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
#pragma unroll
        for (int k = 0; k < CONV_ITEMS; ++k) acc[k] = fmaf(taps[i], v[k - i + 7], acc[k]);
      }
    }
    __syncthreads();
  }
  float* outr = out + row * out_pitch;
  const int nb = n0 + threadIdx.x * CONV_ITEMS;
#pragma unroll
  for (int k = 0; k < CONV_ITEMS; ++k) {
    if (nb + k < out_len) outr[nb + k] = acc[k];
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Wavelet analysis step.  taps: lowpass then highpass, `order` each (host builds them, reference scaling included).
//   decimated : out[d] = sum_j taps[j] * ext(x)[2 d + j],          d < length / 2
//   stationary: out[i] = sum_j taps[j] * ext(x)[i + j * stride],   i < length
// ext continues the row on the right: 0 periodic, 1 mirror, 2 constant, 3 zero.
// ---------------------------------------------------------------------------------------------------------------
constexpr int DWT_MAX_ORDER = 76;
struct DwtTaps {
  float lo[DWT_MAX_ORDER];
  float hi[DWT_MAX_ORDER];
};

__device__ __forceinline__ float dwt_ext(const float* __restrict__ xr, int length, int index, int ext) {
  if (index < length) return __ldg(xr + index);
  const int i = index - length;
  switch (ext) {
    case 0: return __ldg(xr + (i % length));
    case 1: return __ldg(xr + (length - 1 - (i % length)));
    case 2: return __ldg(xr + (length - 1));
    default: return 0.f;
  }
}

constexpr int DWT_THREADS = 256;
constexpr int DWT_ITEMS = 4;                       // outputs per thread (d0 + t + k * 256)
constexpr int DWT_TILE = DWT_THREADS * DWT_ITEMS;  // outputs per block
constexpr int DWT_WIN = 2 * DWT_TILE + DWT_MAX_ORDER + 2;

__global__ void __launch_bounds__(DWT_THREADS)
sig_dwt_kernel(const float* __restrict__ x, long long x_pitch, int length, const __grid_constant__ DwtTaps taps,
               int order, int ext, float* __restrict__ hi, float* __restrict__ lo, long long out_pitch) {
  // window of the block: inputs [2 d0, 2 d0 + 2 * TILE + order), stored with one pad word per 32 so that the
  // stride-2 reads of neighbouring threads hit distinct banks.  Each thread keeps 8+ independent global loads in
  // flight while filling it (the kernel is bound by bytes in flight, not by arithmetic, for short filters).
  __shared__ float win[DWT_WIN + DWT_WIN / 32 + 1];
  const int row = blockIdx.y;
  const int d0 = blockIdx.x * DWT_TILE;
  const float* xr = x + row * x_pitch;
  const int need = 2 * DWT_TILE + order;
  const int first = 2 * d0;
  if (first + need <= length) {      // interior block: no extension logic
#pragma unroll 4
    for (int j = threadIdx.x; j < need; j += DWT_THREADS) win[j + (j >> 5)] = __ldg(xr + first + j);
  } else {
    for (int j = threadIdx.x; j < need; j += DWT_THREADS) win[j + (j >> 5)] = dwt_ext(xr, length, first + j, ext);
  }
  __syncthreads();
  float rh[DWT_ITEMS], rl[DWT_ITEMS];
#pragma unroll
  for (int k = 0; k < DWT_ITEMS; ++k) rh[k] = rl[k] = 0.f;
  for (int j = 0; j < order; ++j) {
    const float th = taps.hi[j], tl = taps.lo[j];
#pragma unroll
    for (int k = 0; k < DWT_ITEMS; ++k) {
      const int idx = 2 * (threadIdx.x + k * DWT_THREADS) + j;
      const float v = win[idx + (idx >> 5)];
      rh[k] = fmaf(th, v, rh[k]);
      rl[k] = fmaf(tl, v, rl[k]);
    }
  }
#pragma unroll
  for (int k = 0; k < DWT_ITEMS; ++k) {
    const int d = d0 + threadIdx.x + k * DWT_THREADS;
    if (d < length / 2) {
      hi[row * out_pitch + d] = rh[k];
      lo[row * out_pitch + d] = rl[k];
    }
  }
}

__global__ void __launch_bounds__(DWT_THREADS)
sig_swt_kernel(const float* __restrict__ x, long long x_pitch, int length, const __grid_constant__ DwtTaps taps,
               int order, int stride, int ext, float* __restrict__ hi, float* __restrict__ lo, long long out_pitch) {
  const int row = blockIdx.y;
  const int i = blockIdx.x * DWT_THREADS + threadIdx.x;
  if (i >= length) return;
  const float* xr = x + row * x_pitch;
  float rh = 0.f, rl = 0.f;
  for (int j = 0; j < order; ++j) {
    // neighbouring threads read neighbouring samples for every tap: coalesced, and L1 keeps the overlap
    const float v = dwt_ext(xr, length, i + j * stride, ext);
    rh = fmaf(taps.hi[j], v, rh);
    rl = fmaf(taps.lo[j], v, rl);
  }
  hi[row * out_pitch + i] = rh;
  lo[row * out_pitch + i] = rl;
}

// ---------------------------------------------------------------------------------------------------------------
// Row min / max (float rows, uint8 planes) and plane normalisation  dst = (v - min) / ((max - min) / 2) - 1.
// One block per row: rows are frames / image planes (<= a few hundred thousand samples), the batch fills the GPU.
// ---------------------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void block_minmax(T& lo, T& hi) {
  __shared__ T s_lo[32], s_hi[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const T a = __shfl_xor_sync(0xffffffffu, lo, o);
    const T b = __shfl_xor_sync(0xffffffffu, hi, o);
    lo = a < lo ? a : lo;
    hi = b > hi ? b : hi;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
    s_lo[warp] = lo;
    s_hi[warp] = hi;
  }
  __syncthreads();
  if (warp == 0) {
    const int nw = blockDim.x >> 5;
    lo = s_lo[lane < nw ? lane : 0];
    hi = s_hi[lane < nw ? lane : 0];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const T a = __shfl_xor_sync(0xffffffffu, lo, o);
      const T b = __shfl_xor_sync(0xffffffffu, hi, o);
      lo = a < lo ? a : lo;
      hi = b > hi ? b : hi;
    }
  }
}

__global__ void __launch_bounds__(512) sig_minmax_f32_kernel(const float* __restrict__ x, long long pitch, int length,
                                                             float* __restrict__ out) {
  const float* xr = x + blockIdx.x * pitch;
  float lo = __ldg(xr), hi = lo;
  int done = 0;
  if ((reinterpret_cast<uintptr_t>(xr) & 15) == 0) {
    // 16-byte loads, four of them in flight per thread
    const float4* x4 = reinterpret_cast<const float4*>(xr);
    const int n4 = length >> 2;
    int i = threadIdx.x;
    for (; i + 3 * 512 < n4; i += 4 * 512) {
      const float4 a = __ldg(x4 + i), b = __ldg(x4 + i + 512), c = __ldg(x4 + i + 1024), d = __ldg(x4 + i + 1536);
      lo = fminf(fminf(fminf(lo, fminf(a.x, a.y)), fminf(fminf(a.z, a.w), fminf(b.x, b.y))),
                 fminf(fminf(fminf(b.z, b.w), fminf(c.x, c.y)), fminf(fminf(c.z, c.w), fminf(fminf(d.x, d.y), fminf(d.z, d.w)))));
      hi = fmaxf(fmaxf(fmaxf(hi, fmaxf(a.x, a.y)), fmaxf(fmaxf(a.z, a.w), fmaxf(b.x, b.y))),
                 fmaxf(fmaxf(fmaxf(b.z, b.w), fmaxf(c.x, c.y)), fmaxf(fmaxf(c.z, c.w), fmaxf(fmaxf(d.x, d.y), fmaxf(d.z, d.w)))));
    }
    for (; i < n4; i += 512) {
      const float4 a = __ldg(x4 + i);
      lo = fminf(fminf(lo, fminf(a.x, a.y)), fminf(a.z, a.w));
      hi = fmaxf(fmaxf(hi, fmaxf(a.x, a.y)), fmaxf(a.z, a.w));
    }
    done = n4 << 2;
  }
  for (int i = done + threadIdx.x; i < length; i += blockDim.x) {
    const float v = __ldg(xr + i);
    lo = fminf(lo, v);
    hi = fmaxf(hi, v);
  }
  block_minmax(lo, hi);
  if (threadIdx.x == 0) {
    out[2 * blockIdx.x] = lo;
    out[2 * blockIdx.x + 1] = hi;
  }
}

// planes: [batch][height rows of `stride` bytes]; out: int32 (min, max) per plane
__global__ void __launch_bounds__(512) sig_minmax_u8_kernel(const uint8_t* __restrict__ src, long long plane_pitch,
                                                            int stride, int width, int height, int* __restrict__ out) {
  const uint8_t* p = src + blockIdx.x * plane_pitch;
  int lo = 255, hi = 0;
  if (((reinterpret_cast<uintptr_t>(p) | static_cast<uintptr_t>(stride) | static_cast<uintptr_t>(width)) & 3) == 0) {
    const int w4 = width >> 2, total4 = w4 * height;
    for (int i = threadIdx.x; i < total4; i += blockDim.x) {
      const int y = i / w4, xx = i - y * w4;
      const uchar4 v = __ldg(reinterpret_cast<const uchar4*>(p + static_cast<long long>(y) * stride) + xx);
      lo = min(min(lo, min(v.x, v.y)), min(v.z, v.w));
      hi = max(max(hi, max(v.x, v.y)), max(v.z, v.w));
    }
  } else {
    const int total = width * height;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
      const int y = i / width, xx = i - y * width;
      const int v = p[static_cast<long long>(y) * stride + xx];
      lo = min(lo, v);
      hi = max(hi, v);
    }
  }
  block_minmax(lo, hi);
  if (threadIdx.x == 0) {
    out[2 * blockIdx.x] = lo;
    out[2 * blockIdx.x + 1] = hi;
  }
}

__global__ void __launch_bounds__(256)
sig_normalize_u8_kernel(const uint8_t* __restrict__ src, long long plane_pitch, int stride, int width, int height,
                        const int* __restrict__ minmax, float* __restrict__ dst, long long dst_plane_pitch,
                        int dst_stride) {
  const int plane = blockIdx.y;
  const int lo = minmax[2 * plane], hi = minmax[2 * plane + 1];
  // flat plane (max == min): all zeros
  const float inv = hi > lo ? 2.f / static_cast<float>(hi - lo) : 0.f;
  const float off = hi > lo ? -static_cast<float>(lo) * inv - 1.f : 0.f;
  const uint8_t* p = src + plane * plane_pitch;
  float* d = dst + plane * dst_plane_pitch;
  const bool vec = ((reinterpret_cast<uintptr_t>(p) | static_cast<uintptr_t>(stride) | static_cast<uintptr_t>(width)) & 3) == 0 &&
                   ((reinterpret_cast<uintptr_t>(d) & 15) | (static_cast<uintptr_t>(dst_stride) & 3)) == 0;
  if (vec) {
    const int w4 = width >> 2, total4 = w4 * height;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total4; i += gridDim.x * blockDim.x) {
      const int y = i / w4, xx = i - y * w4;
      const uchar4 v = __ldg(reinterpret_cast<const uchar4*>(p + static_cast<long long>(y) * stride) + xx);
      float4 r;
      r.x = fmaf(static_cast<float>(v.x), inv, off);
      r.y = fmaf(static_cast<float>(v.y), inv, off);
      r.z = fmaf(static_cast<float>(v.z), inv, off);
      r.w = fmaf(static_cast<float>(v.w), inv, off);
      reinterpret_cast<float4*>(d + static_cast<long long>(y) * dst_stride)[xx] = r;
    }
  } else {
    const int total = width * height;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
      const int y = i / width, xx = i - y * width;
      const float v = static_cast<float>(p[static_cast<long long>(y) * stride + xx]);
      d[static_cast<long long>(y) * dst_stride + xx] = fmaf(v, inv, off);
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// Peak detection: strict interior extrema (d1 = x[i] - x[i-1], d2 = x[i] - x[i+1], d1 * d2 > 0), reported in index
// order.  Pass 1 counts per 1024-sample block; pass 2 re-evaluates the flags, ranks them with a block scan on top of
// the sum of the preceding blocks' counts and writes (position, value) pairs.  Rows are independent (grid.y).
// ---------------------------------------------------------------------------------------------------------------
constexpr int PEAK_THREADS = 256;
constexpr int PEAK_ITEMS = 4;
constexpr int PEAK_TILE = PEAK_THREADS * PEAK_ITEMS;

__device__ __forceinline__ bool is_peak(const float* __restrict__ xr, int length, int i, int type) {
  if (i < 1 || i + 1 >= length) return false;
  const float c = __ldg(xr + i);
  const float d1 = c - __ldg(xr + i - 1);
  const float d2 = c - __ldg(xr + i + 1);
  if (!(d1 * d2 > 0.f)) return false;
  return (d1 > 0.f && (type & 1)) || (d1 < 0.f && (type & 2));
}

__global__ void __launch_bounds__(PEAK_THREADS)
sig_peaks_count_kernel(const float* __restrict__ x, long long pitch, int length, int type, int* __restrict__ counts,
                       int blocks_per_row) {
  const float* xr = x + blockIdx.y * pitch;
  // thread t owns PEAK_ITEMS consecutive samples so that the in-block rank order equals the index order
  const int base = blockIdx.x * PEAK_TILE + threadIdx.x * PEAK_ITEMS;
  int c = 0;
#pragma unroll
  for (int k = 0; k < PEAK_ITEMS; ++k) c += is_peak(xr, length, base + k, type) ? 1 : 0;
  __shared__ int s[PEAK_THREADS / 32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < PEAK_THREADS / 32; ++w) t += s[w];
    counts[blockIdx.y * blocks_per_row + blockIdx.x] = t;
  }
}

__global__ void __launch_bounds__(PEAK_THREADS)
sig_peaks_write_kernel(const float* __restrict__ x, long long pitch, int length, int type,
                       const int* __restrict__ counts, int blocks_per_row, int* __restrict__ positions,
                       float* __restrict__ values, int capacity, int* __restrict__ totals) {
  const int row = blockIdx.y;
  const float* xr = x + row * pitch;
  const int* cr = counts + row * blocks_per_row;
  __shared__ int s_base;
  __shared__ int s_warp[PEAK_THREADS / 32];
  // peaks in the blocks before this one
  int part = 0;
  for (int b = threadIdx.x; b < blockIdx.x; b += PEAK_THREADS) part += cr[b];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if ((threadIdx.x & 31) == 0) s_warp[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < PEAK_THREADS / 32; ++w) t += s_warp[w];
    s_base = t;
  }
  __syncthreads();
  const int base = blockIdx.x * PEAK_TILE + threadIdx.x * PEAK_ITEMS;
  bool f[PEAK_ITEMS];
  int mine = 0;
#pragma unroll
  for (int k = 0; k < PEAK_ITEMS; ++k) {
    f[k] = is_peak(xr, length, base + k, type);
    mine += f[k] ? 1 : 0;
  }
  // exclusive scan of `mine` over the block
  int incl = mine;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  __syncthreads();   // s_warp is reused
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int warp_off = 0;
  for (int w = 0; w < warp; ++w) warp_off += s_warp[w];
  int rank = s_base + warp_off + incl - mine;
#pragma unroll
  for (int k = 0; k < PEAK_ITEMS; ++k) {
    if (f[k]) {
      if (rank < capacity) {
        positions[static_cast<long long>(row) * capacity + rank] = base + k;
        values[static_cast<long long>(row) * capacity + rank] = __ldg(xr + base + k);
      }
      ++rank;
    }
  }
  if (blockIdx.x == gridDim.x - 1 && threadIdx.x == PEAK_THREADS - 1) totals[row] = rank;
}

// ---------------------------------------------------------------------------------------------------------------
// Spectrum product for FFT convolution: out[r][k] = a[r][k] * (conj?) b[r or 0][k] * scale, interleaved complex.
// ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sig_complex_mul_kernel(const float2* __restrict__ a, long long a_pitch, const float2* __restrict__ b, long long b_pitch,
                       float2* __restrict__ out, long long out_pitch, int n, int conj_b, float scale) {
  const int row = blockIdx.y;
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < n; k += gridDim.x * blockDim.x) {
    const float2 u = a[row * a_pitch + k];
    float2 v = b[row * b_pitch + k];
    if (conj_b) v.y = -v.y;
    float2 r;
    r.x = (u.x * v.x - u.y * v.y) * scale;
    r.y = (u.x * v.y + u.y * v.x) * scale;
    out[row * out_pitch + k] = r;
  }
}

// Sample-format conversion: kind 0 int16 -> float, 1 float -> int16 (truncate), 2 int32 -> float, 3 float -> int32,
// 4 half -> float.
__global__ void __launch_bounds__(256) sig_convert_kernel(const void* __restrict__ src, void* __restrict__ dst,
                                                          long long n, int kind) {
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    switch (kind) {
      case 0: static_cast<float*>(dst)[i] = static_cast<float>(static_cast<const int16_t*>(src)[i]); break;
      case 1: static_cast<int16_t*>(dst)[i] = static_cast<int16_t>(static_cast<const float*>(src)[i]); break;
      case 2: static_cast<float*>(dst)[i] = static_cast<float>(static_cast<const int32_t*>(src)[i]); break;
      case 3: static_cast<int32_t*>(dst)[i] = static_cast<int32_t>(static_cast<const float*>(src)[i]); break;
      default: static_cast<float*>(dst)[i] = __half2float(static_cast<const __half*>(src)[i]); break;
    }
  }
}

}  // namespace vb

using namespace vb;

extern "C" int vb_sig_convolve(const float* x, long long x_pitch, int x_length, const float* h, long long h_pitch,
                               int h_length, int reverse, float* out, long long out_pitch, int batch,
                               cudaStream_t s) {
  VB_REQUIRE(x_length > 0 && h_length > 0 && batch > 0, "vb_sig_convolve: empty problem");
  VB_REQUIRE(batch <= 65535, "vb_sig_convolve: at most 65535 rows per call");
  const int out_len = x_length + h_length - 1;
  dim3 grid(ceil_div(out_len, CONV_TILE), batch);
  sig_conv_kernel<<<grid, CONV_THREADS, 0, s>>>(x, x_pitch, x_length, h, h_pitch, h_length, reverse, out, out_pitch);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_sig_wavelet(const float* x, long long x_pitch, int length, const float* lowpass,
                              const float* highpass, int order, int level, int stationary, int ext, float* hi,
                              float* lo, long long out_pitch, int batch, cudaStream_t s) {
  VB_REQUIRE(order >= 2 && order <= DWT_MAX_ORDER, "vb_sig_wavelet: order %d outside 2..%d", order, DWT_MAX_ORDER);
  VB_REQUIRE(length >= 2 && batch > 0 && batch <= 65535, "vb_sig_wavelet: bad length %d / batch %d", length, batch);
  VB_REQUIRE(ext >= 0 && ext <= 3, "vb_sig_wavelet: unknown extension %d", ext);
  DwtTaps taps;
  memset(&taps, 0, sizeof(taps));
  for (int i = 0; i < order; ++i) {   // host pointers: the taps travel as a kernel parameter
    taps.lo[i] = lowpass[i];
    taps.hi[i] = highpass[i];
  }
  if (!stationary) {
    VB_REQUIRE(length % 2 == 0, "vb_sig_wavelet: the decimated transform needs an even length, got %d", length);
    dim3 grid(ceil_div(length / 2, DWT_TILE), batch);
    sig_dwt_kernel<<<grid, DWT_THREADS, 0, s>>>(x, x_pitch, length, taps, order, ext, hi, lo, out_pitch);
  } else {
    VB_REQUIRE(level >= 1 && level <= 24, "vb_sig_wavelet: level %d outside 1..24", level);
    dim3 grid(ceil_div(length, DWT_THREADS), batch);
    sig_swt_kernel<<<grid, DWT_THREADS, 0, s>>>(x, x_pitch, length, taps, order, 1 << (level - 1), ext, hi, lo,
                                                out_pitch);
  }
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_sig_minmax_f32(const float* x, long long pitch, int length, int batch, float* out, cudaStream_t s) {
  VB_REQUIRE(length > 0 && batch > 0, "vb_sig_minmax_f32: empty input");
  sig_minmax_f32_kernel<<<batch, 512, 0, s>>>(x, pitch, length, out);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_sig_normalize_u8(const uint8_t* src, long long plane_pitch, int stride, int width, int height,
                                   int batch, int* minmax, int have_minmax, float* dst, long long dst_plane_pitch,
                                   int dst_stride, cudaStream_t s) {
  VB_REQUIRE(width > 0 && height > 0 && batch > 0 && batch <= 65535, "vb_sig_normalize_u8: bad shape");
  VB_REQUIRE(stride >= width && dst_stride >= width, "vb_sig_normalize_u8: stride smaller than width");
  if (!have_minmax) sig_minmax_u8_kernel<<<batch, 512, 0, s>>>(src, plane_pitch, stride, width, height, minmax);
  if (dst != nullptr) {
    dim3 grid(min(ceil_div(width * height, 256), 64), batch);
    sig_normalize_u8_kernel<<<grid, 256, 0, s>>>(src, plane_pitch, stride, width, height, minmax, dst,
                                                 dst_plane_pitch, dst_stride);
  }
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// counts: int32 scratch of batch * ceil(length / 1024); positions / values: [batch][capacity]; totals: [batch]
extern "C" int vb_sig_detect_peaks(const float* x, long long pitch, int length, int batch, int type, int* counts,
                                   int* positions, float* values, int capacity, int* totals, cudaStream_t s) {
  VB_REQUIRE(length > 0 && batch > 0 && batch <= 65535, "vb_sig_detect_peaks: bad shape");
  VB_REQUIRE(type >= 1 && type <= 3, "vb_sig_detect_peaks: type must be 1 (max), 2 (min) or 3 (both)");
  const int blocks = ceil_div(length, PEAK_TILE);
  dim3 grid(blocks, batch);
  sig_peaks_count_kernel<<<grid, PEAK_THREADS, 0, s>>>(x, pitch, length, type, counts, blocks);
  sig_peaks_write_kernel<<<grid, PEAK_THREADS, 0, s>>>(x, pitch, length, type, counts, blocks, positions, values,
                                                      capacity, totals);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_sig_complex_mul(const float* a, long long a_pitch, const float* b, long long b_pitch, float* out,
                                  long long out_pitch, int n_complex, int batch, int conj_b, float scale,
                                  cudaStream_t s) {
  VB_REQUIRE(n_complex > 0 && batch > 0 && batch <= 65535, "vb_sig_complex_mul: bad shape");
  dim3 grid(min(ceil_div(n_complex, 256), 1024), batch);
  sig_complex_mul_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const float2*>(a), a_pitch,
                                              reinterpret_cast<const float2*>(b), b_pitch,
                                              reinterpret_cast<float2*>(out), out_pitch, n_complex, conj_b, scale);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_sig_convert(const void* src, void* dst, long long n, int kind, cudaStream_t s) {
  VB_REQUIRE(kind >= 0 && kind <= 4, "vb_sig_convert: unknown conversion %d", kind);
  if (n == 0) return 0;
  const int blocks = static_cast<int>(ceil_div_ll(n, 256) < 4096 ? ceil_div_ll(n, 256) : 4096);
  sig_convert_kernel<<<blocks, 256, 0, s>>>(src, dst, n, kind);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
