// veles_b200 — memory-bound kernels: minibatch gather (+cast +normalise),
// mean/dispersion normaliser, joiner copy, stand-alone activations, dropout
// with in-kernel xorshift128+, and the single-GPU fused gradient-descent update.
//
// Reference kernels these replace (all plain one-thread-per-element SIMT):
//   cuda/fullbatch_loader.cu:9-61   fill_minibatch_data_labels / fill_minibatch_target
//   cuda/mean_disp_normalizer.cu:11-19  (launched with block (1,1,1))
//   cuda/join.jcu:10-35             (one thread per sample, serial copy loop)
// Znicz activation / dropout / gradient-descent kernels are absent from the
// reference checkout; their semantics are pinned by
// veles_b200/znicz/*_oracle numpy code.
#include "common.cuh"

namespace vb {

enum DType : int { DT_F32 = 0, DT_BF16 = 1, DT_U8 = 2, DT_I32 = 3 };

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <> __device__ __forceinline__ float to_f<uint8_t>(uint8_t v) { return static_cast<float>(v); }

template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

// ---------------------------------------------------------------------------
// K1/K2/K4: gather rows by shuffled index, cast, optional (x - mean) * rdisp,
// -1 / zero padding for the tail of a short minibatch.
// ---------------------------------------------------------------------------
// A thread owns 4 consecutive COLUMNS and walks GATHER_ROWS_PER_BLOCK rows: the per-feature mean / 1/dispersion of
// its columns are read once and stay in registers.  (Row-major thread assignment re-read both fp32 vectors for every
// row: for a 256 x 154587 uint8 minibatch that is 316 MB of L2 traffic next to 40 MB of image bytes, and it bounded
// the kernel.)
constexpr int GATHER_ROWS_PER_BLOCK = 32;

template <typename SRC, typename DST>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const SRC* __restrict__ src, long long src_ld, const int* __restrict__ indices, int start, int count,
                   int max_rows, int row_size, DST* __restrict__ dst, long long dst_ld,
                   const float* __restrict__ mean, const float* __restrict__ rdisp) {
  vb::pdl_sync();
  const int c = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (c >= row_size) return;
  const int nc = min(4, row_size - c);
  float m[4] = {0.f, 0.f, 0.f, 0.f}, d[4] = {1.f, 1.f, 1.f, 1.f};
  if (mean != nullptr) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < nc) {
        m[j] = __ldg(mean + c + j);
        d[j] = __ldg(rdisp + c + j);
      }
  }
  const int r0 = blockIdx.y * GATHER_ROWS_PER_BLOCK;
  const int r1 = min(max_rows, r0 + GATHER_ROWS_PER_BLOCK);
#pragma unroll 4
  for (int row = r0; row < r1; ++row) {
    DST* drow = dst + static_cast<long long>(row) * dst_ld + c;
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    if (row < count) {
      const SRC* srow = src + static_cast<long long>(__ldg(indices + start + row)) * src_ld + c;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j < nc) v[j] = (to_f<SRC>(srow[j]) - m[j]) * d[j];
    }
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (j < nc) drow[j] = from_f<DST>(v[j]);
  }
}

// ---------------------------------------------------------------------------
// Minibatch gather FUSED with the first layer's space-to-depth transform: rows picked by shuffled index, cast,
// (x - mean) * rdisp, and written straight in the [B, ceil(H/sy), ceil(W/sx), sy*sx*C] bf16 layout the strided
// few-channel convolution reads (AlexNet conv1: 227x227x3 -> 57x57x48).  The plain NHWC minibatch (79 MB per 256
// images) is then neither written nor read again.
// One block = one output row h2 of GS2D_ROWS consecutive minibatch rows.  The sy input rows that feed it are contiguous
// strips of W*C elements: threads read them element by element (coalesced whatever the row alignment - image rows of
// 3-channel uint8 pixels are 1-byte aligned), GS2D_GROUP minibatch rows at a time so that each thread has 4 x 11
// loads in flight (one row at a time was latency-bound: 2x slower than the two kernels it replaces); uint8 sources are
// fetched as aligned words, 8 rows at a time.  Each strip
// element's mean / 1/dispersion and its position in the output strip are tabulated in shared memory once per block;
// converted values are scattered into shared-memory images of the output strips, which leave as aligned 16-byte
// vectors.  Positions past the image edge stay zero.
// ---------------------------------------------------------------------------
constexpr int GS2D_ROWS = 32;     // minibatch rows per block
constexpr int GS2D_GROUP = 4;     // rows whose loads are in flight together
constexpr int GS2D_EPT = 12;      // strip elements per thread: sy * W * C <= 256 * 12

template <typename SRC>
__global__ void __launch_bounds__(256, 3)
gather_s2d_kernel(const SRC* __restrict__ src, long long src_ld, const int* __restrict__ indices, int start, int count,
                  int max_rows, int H, int W, int C, int sy, int sx, int H2, int W2, __nv_bfloat16* __restrict__ xs,
                  const float* __restrict__ mean, const float* __restrict__ rdisp) {
  vb::pdl_sync();
  constexpr int GROUP = GS2D_GROUP;
  extern __shared__ __align__(16) unsigned char gs2d_smem[];
  const int C2 = sy * sx * C;
  const int out_elems = W2 * C2;                 // one output strip (multiple of 8)
  const int row_elems = W * C;
  const int strip_cap = sy * row_elems;
  // [GROUP output strips][(mean, 1/disp) per strip element][output position per strip element]
  __nv_bfloat16* obuf = reinterpret_cast<__nv_bfloat16*>(gs2d_smem);
  float2* md = reinterpret_cast<float2*>(gs2d_smem + static_cast<size_t>(GROUP) * out_elems * 2);
  unsigned short* pos = reinterpret_cast<unsigned short*>(md + strip_cap);
  const int h2 = blockIdx.x;
  const int rows_in = min(sy, H - h2 * sy);      // input rows that exist for this strip
  const int strip = rows_in * row_elems;
  const long long col0 = static_cast<long long>(h2) * sy * row_elems;   // first feature of the strip
  for (int e = threadIdx.x; e < strip; e += 256) {
    const int dy = e / row_elems, rem = e - dy * row_elems;
    const int w = rem / C, c = rem - w * C;
    const int w2 = w / sx, dx = w - w2 * sx;
    pos[e] = static_cast<unsigned short>(w2 * C2 + (dy * sx + dx) * C + c);
    md[e] = mean != nullptr ? make_float2(__ldg(mean + col0 + e), __ldg(rdisp + col0 + e)) : make_float2(0.0f, 1.0f);
  }
  for (int i = threadIdx.x; i < GROUP * out_elems; i += 256)      // padding positions are never written again
    obuf[i] = __float2bfloat16(0.0f);
  __syncthreads();
  const int r0 = blockIdx.y * GS2D_ROWS;
  const int r1 = min(max_rows, r0 + GS2D_ROWS);
  const int vecs = out_elems >> 3;
  for (int g0 = r0; g0 < r1; g0 += GROUP) {
    const SRC* srow[GROUP];
#pragma unroll
    for (int r = 0; r < GROUP; ++r) {
      const int row = g0 + r;
      srow[r] = (row < r1 && row < count)
                    ? src + static_cast<long long>(__ldg(indices + start + row)) * src_ld + col0 : nullptr;
    }
    SRC v[GROUP][GS2D_EPT];
#pragma unroll
    for (int r = 0; r < GROUP; ++r)
#pragma unroll
      for (int i = 0; i < GS2D_EPT; ++i) {
        const int e = threadIdx.x + 256 * i;
        if (e < strip && srow[r] != nullptr) v[r][i] = srow[r][e];
      }
#pragma unroll
    for (int i = 0; i < GS2D_EPT; ++i) {
      const int e = threadIdx.x + 256 * i;
      if (e < strip) {
        const float2 t = md[e];
        const int p = pos[e];
#pragma unroll
        for (int r = 0; r < GROUP; ++r)
          obuf[r * out_elems + p] = __float2bfloat16(srow[r] != nullptr ? (to_f<SRC>(v[r][i]) - t.x) * t.y : 0.0f);
      }
    }
    __syncthreads();
    const int live = min(GROUP, r1 - g0);
    for (int i = threadIdx.x; i < live * vecs; i += 256) {
      const int r = i / vecs, k = i - r * vecs;
      reinterpret_cast<uint4*>(xs + (static_cast<long long>(g0 + r) * H2 + h2) * out_elems)[k] =
          reinterpret_cast<const uint4*>(obuf + r * out_elems)[k];
    }
    __syncthreads();
  }
}

// The same for uint8 sources whose segments are 12 elements long (sx * C = 12: RGB at stride 4, AlexNet's first
// layer).  A thread owns one (dy, w2) segment: 12 CONSECUTIVE source bytes that become 12 CONSECUTIVE output
// elements.  It fetches the (at most four) aligned 32-bit words covering them, shifts the row's misalignment out,
// and writes the 24 output bytes to the shared strip image with three 8-byte stores (element-wise 2-byte stores at
// this lane stride of 96 bytes are 8-way bank conflicts: they made the first form of this kernel conversion-bound);
// mean / 1/dispersion of its 12 features stay in registers for all rows of the block.  A row's last word may reach
// up to 3 bytes past the strip: inside the source tensor except for its very last word, which the allocator's
// granularity covers.
__global__ void __launch_bounds__(256, 3)
gather_s2d_seg12_kernel(const unsigned char* __restrict__ src, long long src_ld, const int* __restrict__ indices,
                        int start, int count, int max_rows, int H, int W, int C, int sy, int H2, int W2,
                        __nv_bfloat16* __restrict__ xs, const float* __restrict__ mean,
                        const float* __restrict__ rdisp) {
  vb::pdl_sync();
  constexpr int GROUP = GS2D_GROUP;
  constexpr int SEG = 12;
  extern __shared__ __align__(16) unsigned char gs2d_smem[];
  const int C2 = sy * SEG;
  const int out_elems = W2 * C2;
  const int row_elems = W * C;
  __nv_bfloat16* obuf = reinterpret_cast<__nv_bfloat16*>(gs2d_smem);
  const int h2 = blockIdx.x;
  const int rows_in = min(sy, H - h2 * sy);
  const long long col0 = static_cast<long long>(h2) * sy * row_elems;
  const int dy = threadIdx.x / W2, w2 = threadIdx.x - dy * W2;
  const int e0 = dy * row_elems + w2 * SEG;                        // first strip element of this thread's segment
  const int n = (dy < rows_in) ? min(SEG, row_elems - w2 * SEG) : 0;   // 0: no segment (or the image ended above)
  float m[SEG], d[SEG];
#pragma unroll
  for (int k = 0; k < SEG; ++k) {
    m[k] = 0.0f;
    d[k] = (k < n) ? 1.0f : 0.0f;                                 // past the image edge: (x - 0) * 0 = 0
    if (k < n && mean != nullptr) {
      m[k] = __ldg(mean + col0 + e0 + k);
      d[k] = __ldg(rdisp + col0 + e0 + k);
    }
  }
  for (int i = threadIdx.x; i < GROUP * out_elems; i += 256) obuf[i] = __float2bfloat16(0.0f);
  __syncthreads();
  const int r0 = blockIdx.y * GS2D_ROWS;
  const int r1 = min(max_rows, r0 + GS2D_ROWS);
  const int vecs = out_elems >> 3;
  const bool owner = dy < sy;                                      // threads past sy * W2 only help with the stores
  for (int g0 = r0; g0 < r1; g0 += GROUP) {
    uint32_t wd[GROUP][4];
    unsigned sh[GROUP];
#pragma unroll
    for (int r = 0; r < GROUP; ++r) {
      const int row = g0 + r;
      sh[r] = 0;
#pragma unroll
      for (int j = 0; j < 4; ++j) wd[r][j] = 0;
      if (n > 0 && row < r1 && row < count) {
        const unsigned char* pb = src + static_cast<long long>(__ldg(indices + start + row)) * src_ld + col0 + e0;
        const unsigned a = static_cast<unsigned>(reinterpret_cast<uintptr_t>(pb) & 3);
        const uint32_t* pw = reinterpret_cast<const uint32_t*>(pb - a);
        const int nw = (static_cast<int>(a) + n + 3) >> 2;
        sh[r] = a * 8;
#pragma unroll
        for (int j = 0; j < 4; ++j)
          if (j < nw) wd[r][j] = __ldg(pw + j);
      }
    }
    if (owner) {
#pragma unroll
      for (int r = 0; r < GROUP; ++r) {
        const bool live = (g0 + r < count);
        uint32_t o[SEG / 2];
#pragma unroll
        for (int j = 0; j < SEG / 4; ++j) {
          const uint32_t word = __funnelshift_r(wd[r][j], wd[r][j + 1], sh[r]);
          float f[4];
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int k = 4 * j + q;
            f[q] = live ? (static_cast<float>((word >> (8 * q)) & 0xffu) - m[k]) * d[k] : 0.0f;
          }
          o[2 * j] = pack_bf16x2(f[0], f[1]);
          o[2 * j + 1] = pack_bf16x2(f[2], f[3]);
        }
        uint2* dst = reinterpret_cast<uint2*>(obuf + r * out_elems + w2 * C2 + dy * SEG);
        dst[0] = make_uint2(o[0], o[1]);
        dst[1] = make_uint2(o[2], o[3]);
        dst[2] = make_uint2(o[4], o[5]);
      }
    }
    __syncthreads();
    const int live_rows = min(GROUP, r1 - g0);
    for (int i = threadIdx.x; i < live_rows * vecs; i += 256) {
      const int r = i / vecs, k = i - r * vecs;
      reinterpret_cast<uint4*>(xs + (static_cast<long long>(g0 + r) * H2 + h2) * out_elems)[k] =
          reinterpret_cast<const uint4*>(obuf + r * out_elems)[k];
    }
    __syncthreads();
  }
}

__global__ void gather_labels_kernel(const int* __restrict__ labels, const int* __restrict__ indices,
                                     int start, int count, int max_rows, int* __restrict__ mb_labels,
                                     int* __restrict__ mb_indices) {
  vb::pdl_sync();
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= max_rows) return;
  if (i < count) {
    const int idx = indices[start + i];
    if (mb_indices) mb_indices[i] = idx;
    if (mb_labels && labels) mb_labels[i] = labels[idx];
  } else {
    if (mb_indices) mb_indices[i] = -1;
    if (mb_labels) mb_labels[i] = -1;
  }
}

template <typename SRC>
static int launch_gather(int dst_dtype, const void* src, long long src_ld, const int* indices, int start,
                         int count, int max_rows, int row_size, void* dst, long long dst_ld,
                         const float* mean, const float* rdisp, cudaStream_t s) {
  dim3 block(256);
  dim3 grid(ceil_div(row_size, 256 * 4), ceil_div(max_rows, GATHER_ROWS_PER_BLOCK));
  if (dst_dtype == DT_F32)
    vb::launch_k(gather_rows_kernel<SRC, float>, dim3(grid), dim3(block), 0, s, static_cast<const SRC*>(src), src_ld, indices, start, count,
                                                          max_rows, row_size, static_cast<float*>(dst), dst_ld, mean, rdisp);
  else
    vb::launch_k(gather_rows_kernel<SRC, __nv_bfloat16>, dim3(grid), dim3(block), 0, s, static_cast<const SRC*>(src), src_ld, indices, start,
                                                                  count, max_rows, row_size,
                                                                  static_cast<__nv_bfloat16*>(dst), dst_ld, mean, rdisp);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------
// Generic strided 2-D copy with cast (InputJoiner column blocks, pitch changes)
// ---------------------------------------------------------------------------
template <typename SRC, typename DST>
__global__ void copy2d_kernel(const SRC* __restrict__ src, long long src_ld, DST* __restrict__ dst,
                              long long dst_ld, int rows, int cols) {
  vb::pdl_sync();
  const long long total = static_cast<long long>(rows) * cols;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / cols);
    const int c = static_cast<int>(i - static_cast<long long>(r) * cols);
    dst[r * dst_ld + c] = from_f<DST>(to_f<SRC>(src[r * src_ld + c]));
  }
}

// ---------------------------------------------------------------------------
// A1: stand-alone activations.  forward: y = f(x); backward: err_in = err_out * f'(.)
// tanh/sigmoid/relu/strict-relu differentiate through the OUTPUT y; log, tanhlog,
// sincos and mul differentiate through the INPUT x (Znicz activation units keep both).
// ---------------------------------------------------------------------------
__device__ __forceinline__ float act_forward_full(int act, float x, float factor, int idx) {
  switch (act) {
    case ACT_LOG:
      return logf(x + sqrtf(x * x + 1.0f));
    case ACT_TANHLOG: {
      // tanh branch inside |x| <= d, log continuation outside (d = 3, scaled tanh)
      const float d = 3.0f, a = 0.242528761112f, b = 305.459953195f;
      if (x > d) return logf(x * b) * a;
      if (x < -d) return -logf(-x * b) * a;
      return 1.7159f * tanhf(0.6666f * x);
    }
    case ACT_SINCOS:
      return (idx & 1) ? sinf(x) : cosf(x);
    case ACT_MUL:
      return x * factor;
    default:
      return act_forward(act, x);
  }
}
__device__ __forceinline__ float act_backward_full(int act, float x, float y, float factor, int idx) {
  switch (act) {
    case ACT_LOG:
      return rsqrtf(x * x + 1.0f);
    case ACT_TANHLOG: {
      const float d = 3.0f, a = 0.242528761112f;
      if (x > d || x < -d) return a / fabsf(x);
      return y * y * (-0.388484177f) + 1.14381894f;
    }
    case ACT_SINCOS:
      return (idx & 1) ? cosf(x) : -sinf(x);
    case ACT_MUL:
      return factor;
    default:
      return act_deriv_from_output(act, y);
  }
}

template <typename T>
__global__ void act_forward_kernel(int act, const T* __restrict__ x, T* __restrict__ y, long long n, float factor) {
  vb::pdl_sync();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    y[i] = from_f<T>(act_forward_full(act, to_f<T>(x[i]), factor, static_cast<int>(i & 1)));
}
template <typename T>
__global__ void act_backward_kernel(int act, const T* __restrict__ err_out, const T* __restrict__ x,
                                    const T* __restrict__ y, T* __restrict__ err_in, long long n, float factor) {
  vb::pdl_sync();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float xv = x ? to_f<T>(x[i]) : 0.0f;
    const float yv = y ? to_f<T>(y[i]) : 0.0f;
    err_in[i] = from_f<T>(to_f<T>(err_out[i]) * act_backward_full(act, xv, yv, factor, static_cast<int>(i & 1)));
  }
}

// ---------------------------------------------------------------------------
// D1: dropout.  Each thread owns one xorshift128+ state (reference generator:
// cuda/random.cu:107-128) and walks a contiguous chunk; the 0/scale mask is
// regenerated in backward from the saved pre-forward states, so neither random
// bytes nor the mask ever reach HBM.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint64_t xs128p_next(uint64_t& s0, uint64_t& s1) {
  uint64_t x = s0;
  const uint64_t y = s1;
  s0 = y;
  x ^= x << 23;
  s1 = x ^ y ^ (x >> 17) ^ (y >> 26);
  return s1 + y;
}

template <typename T>
__global__ void dropout_kernel(const T* __restrict__ in, T* __restrict__ out, long long n,
                               const ulonglong2* __restrict__ states_in, ulonglong2* __restrict__ states_out,
                               int n_states, uint32_t threshold, float scale) {
  vb::pdl_sync();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_states) return;
  ulonglong2 st = states_in[t];
  uint64_t s0 = st.x, s1 = st.y;
  // interleaved assignment keeps the warp's accesses coalesced:
  // element i is handled by thread (i / 2) % n_states, two per 64-bit draw
  // The generator chain is serial per state but the loads are not: draw UNR numbers, then issue all the
  // loads of those iterations together so their latencies overlap (same stream -> same mask as the oracle).
  constexpr int UNR = 4;
  const long long step = 2LL * n_states;
  long long base = 2LL * t;
  for (; base + (UNR - 1) * step + 1 < n; base += UNR * step) {
    uint64_t r[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) r[u] = xs128p_next(s0, s1);
    float a[UNR], b[UNR];
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      a[u] = to_f<T>(in[base + u * step]);
      b[u] = to_f<T>(in[base + u * step + 1]);
    }
#pragma unroll
    for (int u = 0; u < UNR; ++u) {
      const uint32_t r0 = static_cast<uint32_t>(r[u]), r1 = static_cast<uint32_t>(r[u] >> 32);
      out[base + u * step] = from_f<T>(r0 >= threshold ? a[u] * scale : 0.0f);
      out[base + u * step + 1] = from_f<T>(r1 >= threshold ? b[u] * scale : 0.0f);
    }
  }
  for (; base < n; base += step) {
    const uint64_t r = xs128p_next(s0, s1);
    const uint32_t r0 = static_cast<uint32_t>(r), r1 = static_cast<uint32_t>(r >> 32);
    out[base] = from_f<T>(r0 >= threshold ? to_f<T>(in[base]) * scale : 0.0f);
    if (base + 1 < n) out[base + 1] = from_f<T>(r1 >= threshold ? to_f<T>(in[base + 1]) * scale : 0.0f);
  }
  if (states_out) states_out[t] = make_ulonglong2(s0, s1);
}

// ---------------------------------------------------------------------------
// U1: fused Znicz gradient-descent step (is the world_size==1 case of C1).
//   g   = grad*grad_scale + wd * ((1-l1_vs_l2)*W + 0.5*l1_vs_l2*sign(W))
//   sgd     : v = moment*v - lr*g ;           W += v
//   adagrad : G += g^2 ; W -= lr*g/sqrt(G+eps)
//   adadelta: G = rho*G+(1-rho)g^2 ; d = -sqrt(v+eps)/sqrt(G+eps)*g ; v = rho*v+(1-rho)d^2 ; W += d
// Also emits the bf16 shadow copy the next forward GEMM TMA-loads, and may zero
// the fp32 gradient so split-K atomics can accumulate into it next step.
// ---------------------------------------------------------------------------
struct GDParams {
  float lr, wd, l1_vs_l2, moment, grad_scale, eps, rho;
  int solver;     // 0 sgd(+momentum), 1 adagrad, 2 adadelta
  int zero_grad;  // write 0 back to grad after consuming it
};

__device__ __forceinline__ float gd_step(const GDParams& p, float w, float g, float& v, float& G) {
  const float sgn = (w > 0.0f) ? 1.0f : ((w < 0.0f) ? -1.0f : 0.0f);
  g = g * p.grad_scale + p.wd * ((1.0f - p.l1_vs_l2) * w + 0.5f * p.l1_vs_l2 * sgn);
  if (p.solver == 0) {
    v = p.moment * v - p.lr * g;
    return w + v;
  } else if (p.solver == 1) {
    G += g * g;
    return w - p.lr * g * rsqrtf(G + p.eps);
  } else {
    G = p.rho * G + (1.0f - p.rho) * g * g;
    const float d = -sqrtf(v + p.eps) * rsqrtf(G + p.eps) * g;
    v = p.rho * v + (1.0f - p.rho) * d * d;
    return w + d;
  }
}

__global__ void gd_update_kernel(float* __restrict__ w, float* __restrict__ grad, float* __restrict__ vel,
                                 float* __restrict__ acc, __nv_bfloat16* __restrict__ w_bf16, long long n,
                                 const GDParams p) {
  vb::pdl_sync();
  const long long n4 = n >> 2;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n4;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    float4 wv = reinterpret_cast<float4*>(w)[i];
    float4 gv = reinterpret_cast<float4*>(grad)[i];
    float4 vv = vel ? reinterpret_cast<float4*>(vel)[i] : make_float4(0, 0, 0, 0);
    float4 av = acc ? reinterpret_cast<float4*>(acc)[i] : make_float4(0, 0, 0, 0);
    wv.x = gd_step(p, wv.x, gv.x, vv.x, av.x);
    wv.y = gd_step(p, wv.y, gv.y, vv.y, av.y);
    wv.z = gd_step(p, wv.z, gv.z, vv.z, av.z);
    wv.w = gd_step(p, wv.w, gv.w, vv.w, av.w);
    reinterpret_cast<float4*>(w)[i] = wv;
    if (vel) reinterpret_cast<float4*>(vel)[i] = vv;
    if (acc) reinterpret_cast<float4*>(acc)[i] = av;
    if (p.zero_grad) reinterpret_cast<float4*>(grad)[i] = make_float4(0, 0, 0, 0);
    if (w_bf16) {
      uint2 o;
      o.x = pack_bf16x2(wv.x, wv.y);
      o.y = pack_bf16x2(wv.z, wv.w);
      reinterpret_cast<uint2*>(w_bf16)[i] = o;
    }
  }
  // scalar tail
  const long long tail0 = n4 << 2;
  const long long t = tail0 + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (t < n) {
    float vv = vel ? vel[t] : 0.0f, av = acc ? acc[t] : 0.0f;
    const float nw = gd_step(p, w[t], grad[t], vv, av);
    w[t] = nw;
    if (vel) vel[t] = vv;
    if (acc) acc[t] = av;
    if (p.zero_grad) grad[t] = 0.0f;
    if (w_bf16) w_bf16[t] = __float2bfloat16(nw);
  }
}

// fp32 -> bf16 (weight shadow initialisation) and bf16 -> fp32
__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n) {
  vb::pdl_sync();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = __float2bfloat16(in[i]);
}
__global__ void cast_bf16_f32_kernel(const __nv_bfloat16* __restrict__ in, float* __restrict__ out, long long n) {
  vb::pdl_sync();
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x)
    out[i] = __bfloat162float(in[i]);
}

static inline int grid_for(long long n, int block, int per_thread = 1) {
  long long g = ceil_div_ll(n, static_cast<long long>(block) * per_thread);
  const long long cap = static_cast<long long>(device_sm_count()) * 16;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

}  // namespace vb

using namespace vb;

extern "C" int vb_gather_minibatch(const void* src, int src_dtype, long long src_ld, const int* indices,
                                   int start, int count, int max_rows, int row_size, void* dst,
                                   int dst_dtype, long long dst_ld, const float* mean, const float* rdisp,
                                   const int* labels, int* mb_labels, int* mb_indices, cudaStream_t s) {
  VB_REQUIRE(count >= 0 && count <= max_rows, "vb_gather_minibatch: count %d out of [0,%d]", count, max_rows);
  VB_REQUIRE(dst_dtype == DT_F32 || dst_dtype == DT_BF16, "vb_gather_minibatch: dst dtype must be f32/bf16");
  if (max_rows == 0) return 0;
  int rc;
  switch (src_dtype) {
    case DT_U8: rc = launch_gather<uint8_t>(dst_dtype, src, src_ld, indices, start, count, max_rows, row_size, dst, dst_ld, mean, rdisp, s); break;
    case DT_F32: rc = launch_gather<float>(dst_dtype, src, src_ld, indices, start, count, max_rows, row_size, dst, dst_ld, mean, rdisp, s); break;
    case DT_BF16: rc = launch_gather<__nv_bfloat16>(dst_dtype, src, src_ld, indices, start, count, max_rows, row_size, dst, dst_ld, mean, rdisp, s); break;
    default: VB_REQUIRE(false, "vb_gather_minibatch: unsupported source dtype %d", src_dtype);
  }
  if (rc) return rc;
  if (mb_labels || mb_indices) {
    vb::launch_k(gather_labels_kernel, dim3(ceil_div(max_rows, 256)), dim3(256), 0, s, labels, indices, start, count, max_rows, mb_labels, mb_indices);
    VB_CHECK_CUDA(cudaGetLastError());
  }
  return 0;
}

// Fused gather + space-to-depth (see gather_s2d_kernel).  Returns 1 when the geometry is not covered (run
// vb_gather_minibatch and vb_space_to_depth instead).  xs: bf16 [max_rows, ceil(H/sy), ceil(W/sx), sy*sx*C], 16-byte
// aligned; labels / indices are gathered like vb_gather_minibatch does.
extern "C" int vb_gather_s2d(const void* src, int src_dtype, long long src_ld, const int* indices, int start,
                             int count, int max_rows, int H, int W, int C, int sy, int sx, void* xs, const float* mean,
                             const float* rdisp, const int* labels, int* mb_labels, int* mb_indices, cudaStream_t s) {
  VB_REQUIRE(count >= 0 && count <= max_rows, "vb_gather_s2d: count %d out of [0,%d]", count, max_rows);
  if (max_rows == 0) return 0;
  const int C2 = sy * sx * C;
  const int H2 = (H + sy - 1) / sy, W2 = (W + sx - 1) / sx;
  // GROUP output strips + (mean, 1/disp) and output position per strip element
  long long smem = static_cast<long long>(GS2D_GROUP) * W2 * C2 * 2 + static_cast<long long>(sy) * W * C * 10;
  if (C2 % 8 != 0 || sy * W * C > 256 * GS2D_EPT || W2 * C2 > 65535 || smem > 96 * 1024 ||
      (reinterpret_cast<uintptr_t>(xs) & 15) != 0)
    return 1;
  if (smem > 48 * 1024) {
    static bool attr_set = false;
    if (!attr_set) {
      VB_CHECK_CUDA(cudaFuncSetAttribute(gather_s2d_kernel<uint8_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      VB_CHECK_CUDA(cudaFuncSetAttribute(gather_s2d_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      VB_CHECK_CUDA(cudaFuncSetAttribute(gather_s2d_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      attr_set = true;
    }
  }
  dim3 grid(H2, ceil_div(max_rows, GS2D_ROWS));
  __nv_bfloat16* out = static_cast<__nv_bfloat16*>(xs);
  if (src_dtype == DT_U8 && sx * C == 12 && sy * W2 <= 256) {
    smem = static_cast<long long>(GS2D_GROUP) * W2 * C2 * 2;
    vb::launch_k(gather_s2d_seg12_kernel, grid, dim3(256), smem, s, static_cast<const unsigned char*>(src), src_ld,
                 indices, start, count, max_rows, H, W, C, sy, H2, W2, out, mean, rdisp);
  } else {
    switch (src_dtype) {
      case DT_U8:
        vb::launch_k(gather_s2d_kernel<uint8_t>, grid, dim3(256), smem, s, static_cast<const uint8_t*>(src), src_ld,
                     indices, start, count, max_rows, H, W, C, sy, sx, H2, W2, out, mean, rdisp);
        break;
      case DT_F32:
        vb::launch_k(gather_s2d_kernel<float>, grid, dim3(256), smem, s, static_cast<const float*>(src), src_ld,
                     indices, start, count, max_rows, H, W, C, sy, sx, H2, W2, out, mean, rdisp);
        break;
      case DT_BF16:
        vb::launch_k(gather_s2d_kernel<__nv_bfloat16>, grid, dim3(256), smem, s,
                     static_cast<const __nv_bfloat16*>(src), src_ld, indices, start, count, max_rows, H, W, C, sy, sx,
                     H2, W2, out, mean, rdisp);
        break;
      default: VB_REQUIRE(false, "vb_gather_s2d: unsupported source dtype %d", src_dtype);
    }
  }
  VB_CHECK_CUDA(cudaGetLastError());
  if (mb_labels || mb_indices) {
    vb::launch_k(gather_labels_kernel, dim3(ceil_div(max_rows, 256)), dim3(256), 0, s, labels, indices, start, count,
                 max_rows, mb_labels, mb_indices);
    VB_CHECK_CUDA(cudaGetLastError());
  }
  return 0;
}

namespace vb {
template <typename SRC, typename DST>
__global__ void mean_disp_kernel(const SRC* __restrict__ in, const float* __restrict__ mean,
                                 const float* __restrict__ rdisp, DST* __restrict__ out, int batch, int size) {
  vb::pdl_sync();
  const long long total = static_cast<long long>(batch) * size;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int j = static_cast<int>(i % size);
    out[i] = from_f<DST>((to_f<SRC>(in[i]) - mean[j]) * rdisp[j]);
  }
}
}  // namespace vb

// out[i,j] = ((float)in[i,j] - mean[j]) * rdisp[j]   (ref: cuda/mean_disp_normalizer.cu:11-19)
extern "C" int vb_mean_disp_normalize(const void* in, int in_dtype, const float* mean, const float* rdisp,
                                      void* out, int out_dtype, int batch, int size, cudaStream_t s) {
  const long long total = static_cast<long long>(batch) * size;
  if (total == 0) return 0;
  const int g = grid_for(total, 256, 4);
#define VB_MD(SRC, DST) vb::launch_k(mean_disp_kernel<SRC, DST>, dim3(g), dim3(256), 0, s, static_cast<const SRC*>(in), mean, rdisp, static_cast<DST*>(out), batch, size)
  if (in_dtype == DT_U8 && out_dtype == DT_F32) VB_MD(uint8_t, float);
  else if (in_dtype == DT_U8 && out_dtype == DT_BF16) VB_MD(uint8_t, __nv_bfloat16);
  else if (in_dtype == DT_F32 && out_dtype == DT_F32) VB_MD(float, float);
  else if (in_dtype == DT_F32 && out_dtype == DT_BF16) VB_MD(float, __nv_bfloat16);
  else if (in_dtype == DT_BF16 && out_dtype == DT_BF16) VB_MD(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == DT_BF16 && out_dtype == DT_F32) VB_MD(__nv_bfloat16, float);
  else VB_REQUIRE(false, "vb_mean_disp_normalize: unsupported dtype pair %d -> %d", in_dtype, out_dtype);
#undef VB_MD
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_copy_2d(const void* src, int src_dtype, long long src_ld, void* dst, int dst_dtype,
                          long long dst_ld, int rows, int cols, cudaStream_t s) {
  const long long total = static_cast<long long>(rows) * cols;
  if (total == 0) return 0;
  const int g = grid_for(total, 256, 4);
#define VB_CP(SRC, DST) vb::launch_k(copy2d_kernel<SRC, DST>, dim3(g), dim3(256), 0, s, static_cast<const SRC*>(src), src_ld, static_cast<DST*>(dst), dst_ld, rows, cols)
  if (src_dtype == DT_F32 && dst_dtype == DT_F32) VB_CP(float, float);
  else if (src_dtype == DT_F32 && dst_dtype == DT_BF16) VB_CP(float, __nv_bfloat16);
  else if (src_dtype == DT_BF16 && dst_dtype == DT_BF16) VB_CP(__nv_bfloat16, __nv_bfloat16);
  else if (src_dtype == DT_BF16 && dst_dtype == DT_F32) VB_CP(__nv_bfloat16, float);
  else if (src_dtype == DT_U8 && dst_dtype == DT_F32) VB_CP(uint8_t, float);
  else if (src_dtype == DT_U8 && dst_dtype == DT_BF16) VB_CP(uint8_t, __nv_bfloat16);
  else VB_REQUIRE(false, "vb_copy_2d: unsupported dtype pair %d -> %d", src_dtype, dst_dtype);
#undef VB_CP
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_activation_forward(int act, const void* x, void* y, long long n, int dtype, float factor,
                                     cudaStream_t s) {
  if (n == 0) return 0;
  const int g = grid_for(n, 256, 4);
  if (dtype == DT_F32) vb::launch_k(act_forward_kernel<float>, dim3(g), dim3(256), 0, s, act, static_cast<const float*>(x), static_cast<float*>(y), n, factor);
  else if (dtype == DT_BF16) vb::launch_k(act_forward_kernel<__nv_bfloat16>, dim3(g), dim3(256), 0, s, act, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), n, factor);
  else VB_REQUIRE(false, "vb_activation_forward: dtype must be f32/bf16");
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_activation_backward(int act, const void* err_out, const void* x, const void* y, void* err_in,
                                      long long n, int dtype, float factor, cudaStream_t s) {
  if (n == 0) return 0;
  const int g = grid_for(n, 256, 4);
  if (dtype == DT_F32)
    vb::launch_k(act_backward_kernel<float>, dim3(g), dim3(256), 0, s, act, static_cast<const float*>(err_out), static_cast<const float*>(x),
                                                 static_cast<const float*>(y), static_cast<float*>(err_in), n, factor);
  else if (dtype == DT_BF16)
    vb::launch_k(act_backward_kernel<__nv_bfloat16>, dim3(g), dim3(256), 0, s, act, static_cast<const __nv_bfloat16*>(err_out),
                                                         static_cast<const __nv_bfloat16*>(x),
                                                         static_cast<const __nv_bfloat16*>(y),
                                                         static_cast<__nv_bfloat16*>(err_in), n, factor);
  else VB_REQUIRE(false, "vb_activation_backward: dtype must be f32/bf16");
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Dropout forward AND backward (backward = same call on err_output with the saved states).
extern "C" int vb_dropout(const void* in, void* out, long long n, int dtype, const void* states_in,
                          void* states_out, int n_states, float ratio, cudaStream_t s) {
  if (n == 0) return 0;
  VB_REQUIRE(ratio >= 0.0f && ratio < 1.0f, "vb_dropout: ratio must be in [0,1)");
  const double th = static_cast<double>(ratio) * 4294967296.0;
  const uint32_t threshold = th >= 4294967295.0 ? 0xFFFFFFFFu : static_cast<uint32_t>(th);
  const float scale = 1.0f / (1.0f - ratio);
  const int g = ceil_div(n_states, 128);
  if (dtype == DT_F32)
    vb::launch_k(dropout_kernel<float>, dim3(g), dim3(128), 0, s, static_cast<const float*>(in), static_cast<float*>(out), n,
                                            static_cast<const ulonglong2*>(states_in), static_cast<ulonglong2*>(states_out),
                                            n_states, threshold, scale);
  else if (dtype == DT_BF16)
    vb::launch_k(dropout_kernel<__nv_bfloat16>, dim3(g), dim3(128), 0, s, static_cast<const __nv_bfloat16*>(in), static_cast<__nv_bfloat16*>(out), n,
                                                    static_cast<const ulonglong2*>(states_in),
                                                    static_cast<ulonglong2*>(states_out), n_states, threshold, scale);
  else VB_REQUIRE(false, "vb_dropout: dtype must be f32/bf16");
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_gd_update(float* w, float* grad, float* vel, float* acc, void* w_bf16, long long n, float lr,
                            float wd, float l1_vs_l2, float moment, float grad_scale, int solver, float eps,
                            float rho, int zero_grad, cudaStream_t s) {
  if (n == 0) return 0;
  VB_REQUIRE((reinterpret_cast<uintptr_t>(w) & 15) == 0 && (reinterpret_cast<uintptr_t>(grad) & 15) == 0,
             "vb_gd_update: weights/grad must be 16-byte aligned");
  GDParams p{lr, wd, l1_vs_l2, moment, grad_scale, eps, rho, solver, zero_grad};
  const int g = grid_for(n, 256, 4);
  vb::launch_k(gd_update_kernel, dim3(g), dim3(256), 0, s, w, grad, vel, acc, static_cast<__nv_bfloat16*>(w_bf16), n, p);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------
// Second phase of a split-K layer GEMM: out = f(acc + bias) [* f'(z)], acc is the fp32 matrix the K-splits
// reduce-added into; it is zeroed again on the way out so the next step can accumulate into it.
// ---------------------------------------------------------------------------
template <typename DST>
__global__ void bias_act_kernel(float* __restrict__ acc, long long ld_acc, const float* __restrict__ bias, int act,
                                const __nv_bfloat16* __restrict__ z, long long ldz, int deriv, DST* __restrict__ out,
                                long long ld_out, int M, int N, int zero_acc) {
  vb::pdl_sync();
  const int n4 = N >> 2;   // N % 4 == 0 (checked by the launcher)
  const long long total = static_cast<long long>(M) * n4;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(t / n4);
    const int c = static_cast<int>(t - static_cast<long long>(r) * n4) << 2;
    float4* ap = reinterpret_cast<float4*>(acc + r * ld_acc + c);
    float4 v = *ap;
    if (zero_acc) *ap = make_float4(0.f, 0.f, 0.f, 0.f);
    float f[4] = {v.x, v.y, v.z, v.w};
    if (bias != nullptr) {
      const float4 b = __ldg(reinterpret_cast<const float4*>(bias + c));
      f[0] += b.x; f[1] += b.y; f[2] += b.z; f[3] += b.w;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) f[q] = act_forward(act, f[q]);
    if (z != nullptr) {
#pragma unroll
      for (int q = 0; q < 4; ++q) f[q] *= act_deriv_from_output(deriv, __bfloat162float(z[r * ldz + c + q]));
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) out[r * ld_out + c + q] = from_f<DST>(f[q]);
  }
}

extern "C" int vb_bias_act_2d(float* acc, long long ld_acc, const float* bias, int act, const void* z, long long ldz,
                              int deriv, void* out, int out_dtype, long long ld_out, int M, int N, int zero_acc,
                              cudaStream_t s) {
  if (M == 0 || N == 0) return 0;
  VB_REQUIRE(N % 4 == 0 && ld_acc % 4 == 0 && (reinterpret_cast<uintptr_t>(acc) & 15) == 0 &&
                 (bias == nullptr || (reinterpret_cast<uintptr_t>(bias) & 15) == 0),
             "vb_bias_act_2d: N and the accumulator pitch must be multiples of 4, pointers 16-byte aligned");
  const int g = grid_for(static_cast<long long>(M) * (N >> 2), 256, 2);
  if (out_dtype == DT_F32)
    vb::launch_k(bias_act_kernel<float>, dim3(g), dim3(256), 0, s, acc, ld_acc, bias, act, static_cast<const __nv_bfloat16*>(z), ldz, deriv,
                                             static_cast<float*>(out), ld_out, M, N, zero_acc);
  else if (out_dtype == DT_BF16)
    vb::launch_k(bias_act_kernel<__nv_bfloat16>, dim3(g), dim3(256), 0, s, acc, ld_acc, bias, act, static_cast<const __nv_bfloat16*>(z), ldz,
                                                     deriv, static_cast<__nv_bfloat16*>(out), ld_out, M, N, zero_acc);
  else
    VB_REQUIRE(false, "vb_bias_act_2d: output dtype must be f32/bf16");
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_cast(const void* in, int in_dtype, void* out, int out_dtype, long long n, cudaStream_t s) {
  if (n == 0) return 0;
  const int g = grid_for(n, 256, 4);
  if (in_dtype == DT_F32 && out_dtype == DT_BF16)
    vb::launch_k(cast_f32_bf16_kernel, dim3(g), dim3(256), 0, s, static_cast<const float*>(in), static_cast<__nv_bfloat16*>(out), n);
  else if (in_dtype == DT_BF16 && out_dtype == DT_F32)
    vb::launch_k(cast_bf16_f32_kernel, dim3(g), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(in), static_cast<float*>(out), n);
  else VB_REQUIRE(false, "vb_cast: unsupported dtype pair %d -> %d", in_dtype, out_dtype);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// factor_ortho regulariser of the Znicz GD units: grad[i,j] += f * (sums[j] - w[i,j])  (transposed: sums[i]).
// `sums` holds the column (row) sums of the fp32 weight matrix; f already carries factor_ortho / n_rows.
namespace vb {
__global__ void ortho_grad_kernel(const float* __restrict__ w, float* __restrict__ grad, const float* __restrict__ sums,
                                  long long rows, long long cols, long long ldw, long long ldg, float f,
                                  int by_row) {
  vb::pdl_sync();
  const long long n = rows * cols;
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < n;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = i / cols, c = i - r * cols;
    grad[r * ldg + c] += f * (sums[by_row ? r : c] - w[r * ldw + c]);
  }
}
}  // namespace vb

extern "C" int vb_ortho_grad(const float* w, long long ldw, float* grad, long long ldg, const float* sums, long long rows,
                             long long cols, float f, int by_row, cudaStream_t s) {
  VB_REQUIRE(rows > 0 && cols > 0, "vb_ortho_grad: empty matrix");
  const long long n = rows * cols;
  const long long cap = static_cast<long long>(vb::device_sm_count()) * 16;
  long long blocks = (n + 255) / 256;
  if (blocks > cap) blocks = cap;
  vb::launch_k(vb::ortho_grad_kernel, dim3(static_cast<int>(blocks)), dim3(256), 0, s, w, grad, sums, rows, cols, ldw, ldg, f, by_row);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
