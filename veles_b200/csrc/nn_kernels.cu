// veles_b200 — spatial / reduction kernels of the Znicz layer set, NHWC layout:
// im2col / col2im ("unpack" for the conv GEMMs), max / maxabs / avg / stochastic
// pooling forward+backward, cross-channel LRN forward+backward, fused softmax +
// evaluator, MSE evaluator, row/column reductions.
//
// The Znicz sources are absent from the reference checkout; layer
// semantics follow docs/source/manualrst_veles_workflow_parameters.rst:469-541
// and are pinned by the numpy oracles in veles_b200/znicz/.  matrix_reduce
// replaces cuda/matrix_reduce.cu:35-81 (shared-memory tree) with warp shuffles.
#include "common.cuh"
#include "fast_paths.cuh"

namespace vb {

enum DType2 : int { D_F32 = 0, D_BF16 = 1 };

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

static inline int grid_cap(long long n, int block) {
  long long g = (n + block - 1) / block;
  const long long cap = static_cast<long long>(device_sm_count()) * 32;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

// ---------------------------------------------------------------------------
// im2col: x[B,H,W,C] -> cols[B*OH*OW, ld] with column index
//   g*(ky*kx*Cg) + (i*kx + j)*Cg + cg        (c = g*Cg + cg)
// Vectorised over 8 channels (16 B) when Cg % 8 == 0, scalar otherwise.
// Columns in [groups*ky*kx*Cg, ld) are zero-filled so the GEMM's K tail is clean.
// ---------------------------------------------------------------------------
struct ConvGeom {
  int B, H, W, C, OH, OW, ky, kx, sy, sx, pt, pl, groups;
  long long ld;  // row pitch of cols (elements)
};

__global__ void im2col_vec8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols,
                                   const ConvGeom g) {
  vb::pdl_sync();
  const int Cg = g.C / g.groups;
  const int kk = g.ky * g.kx;
  const int vec_per_row = static_cast<int>(g.ld >> 3);
  const int used = g.groups * kk * Cg;
  const long long rows = static_cast<long long>(g.B) * g.OH * g.OW;
  const long long total = rows * vec_per_row;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long row = t / vec_per_row;
    const int col = static_cast<int>(t - row * vec_per_row) << 3;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (col < used) {
      const int grp = col / (kk * Cg);
      const int rem = col - grp * kk * Cg;
      const int tap = rem / Cg;
      const int cg = rem - tap * Cg;
      const int i = tap / g.kx, j = tap - i * g.kx;
      const int ow = static_cast<int>(row % g.OW);
      const long long r2 = row / g.OW;
      const int oh = static_cast<int>(r2 % g.OH);
      const int b = static_cast<int>(r2 / g.OH);
      const int ih = oh * g.sy - g.pt + i, iw = ow * g.sx - g.pl + j;
      if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W)
        v = *reinterpret_cast<const uint4*>(x + ((static_cast<long long>(b) * g.H + ih) * g.W + iw) * g.C + grp * Cg + cg);
    }
    *reinterpret_cast<uint4*>(cols + row * g.ld + col) = v;
  }
}

__global__ void im2col_scalar_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols,
                                     const ConvGeom g) {
  vb::pdl_sync();
  const int Cg = g.C / g.groups;
  const int kk = g.ky * g.kx;
  const int used = g.groups * kk * Cg;
  const long long rows = static_cast<long long>(g.B) * g.OH * g.OW;
  const long long total = rows * g.ld;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long row = t / g.ld;
    const int col = static_cast<int>(t - row * g.ld);
    __nv_bfloat16 v = __float2bfloat16(0.0f);
    if (col < used) {
      const int grp = col / (kk * Cg);
      const int rem = col - grp * kk * Cg;
      const int tap = rem / Cg;
      const int cg = rem - tap * Cg;
      const int i = tap / g.kx, j = tap - i * g.kx;
      const int ow = static_cast<int>(row % g.OW);
      const long long r2 = row / g.OW;
      const int oh = static_cast<int>(r2 % g.OH);
      const int b = static_cast<int>(r2 / g.OH);
      const int ih = oh * g.sy - g.pt + i, iw = ow * g.sx - g.pl + j;
      if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W)
        v = x[((static_cast<long long>(b) * g.H + ih) * g.W + iw) * g.C + grp * Cg + cg];
    }
    cols[t] = v;
  }
}

// col2im in gather form (deterministic, no atomics, no pre-zeroing):
// dx[b,ih,iw,c] = sum over taps (i,j) with (ih+pt-i) % sy == 0 ... of dcols[row(oh,ow), col(i,j,c)]
// optional fused activation derivative through the saved forward output z of the
// PREVIOUS layer (dx *= f'(z)), which is what Znicz GD units apply to err_output.
__global__ void col2im_kernel(const __nv_bfloat16* __restrict__ dcols, __nv_bfloat16* __restrict__ dx,
                              const ConvGeom g, const __nv_bfloat16* __restrict__ z, int deriv) {
  vb::pdl_sync();
  const int Cg = g.C / g.groups;
  const int kk = g.ky * g.kx;
  const long long total = static_cast<long long>(g.B) * g.H * g.W * g.C;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % g.C);
    long long r = t / g.C;
    const int iw = static_cast<int>(r % g.W);
    r /= g.W;
    const int ih = static_cast<int>(r % g.H);
    const int b = static_cast<int>(r / g.H);
    const int grp = c / Cg, cg = c - grp * Cg;
    float acc = 0.0f;
    for (int i = 0; i < g.ky; ++i) {
      const int nh = ih + g.pt - i;
      if (nh < 0 || nh % g.sy) continue;
      const int oh = nh / g.sy;
      if (oh >= g.OH) continue;
      for (int j = 0; j < g.kx; ++j) {
        const int nw = iw + g.pl - j;
        if (nw < 0 || nw % g.sx) continue;
        const int ow = nw / g.sx;
        if (ow >= g.OW) continue;
        const long long row = (static_cast<long long>(b) * g.OH + oh) * g.OW + ow;
        acc += __bfloat162float(dcols[row * g.ld + grp * kk * Cg + (i * g.kx + j) * Cg + cg]);
      }
    }
    if (z != nullptr) acc *= act_deriv_from_output(deriv, __bfloat162float(z[t]));
    dx[t] = __float2bfloat16(acc);
  }
}

// ---------------------------------------------------------------------------
// Pooling (NHWC).  kind: 0 max, 1 maxabs, 2 avg, 3 stochastic, 4 stochastic-abs.
// Windows are clipped at the right/bottom border (Znicz semantics: the last
// window may be partial).  `offs` records the flat input index of the chosen
// element for max-like kinds so backward is an exact routing.
// ---------------------------------------------------------------------------
struct PoolGeom {
  int B, H, W, C, OH, OW, ky, kx, sy, sx;
};

template <typename T>
__global__ void pool_forward_kernel(int kind, const T* __restrict__ x, T* __restrict__ y, int* __restrict__ offs,
                                    const PoolGeom g, const uint16_t* __restrict__ rnd) {
  vb::pdl_sync();
  const long long total = static_cast<long long>(g.B) * g.OH * g.OW * g.C;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % g.C);
    long long r = t / g.C;
    const int ow = static_cast<int>(r % g.OW);
    r /= g.OW;
    const int oh = static_cast<int>(r % g.OH);
    const int b = static_cast<int>(r / g.OH);
    const int h0 = oh * g.sy, w0 = ow * g.sx;
    const int h1 = min(h0 + g.ky, g.H), w1 = min(w0 + g.kx, g.W);
    const long long base = static_cast<long long>(b) * g.H * g.W * g.C + c;
    if (kind == 2) {
      float s = 0.0f;
      for (int h = h0; h < h1; ++h)
        for (int w = w0; w < w1; ++w) s += ldf<T>(x + base + (static_cast<long long>(h) * g.W + w) * g.C);
      stf<T>(y + t, s / static_cast<float>((h1 - h0) * (w1 - w0)));
    } else if (kind <= 1) {
      float best = 0.0f, bestkey = -INFINITY;
      long long bo = base + (static_cast<long long>(h0) * g.W + w0) * g.C;
      for (int h = h0; h < h1; ++h)
        for (int w = w0; w < w1; ++w) {
          const long long o = base + (static_cast<long long>(h) * g.W + w) * g.C;
          const float v = ldf<T>(x + o);
          const float key = kind == 1 ? fabsf(v) : v;
          if (key > bestkey) {
            bestkey = key;
            best = v;
            bo = o;
          }
        }
      stf<T>(y + t, best);
      if (offs) offs[t] = static_cast<int>(bo);
    } else {
      // stochastic: pick element with probability proportional to max(v,0) (or |v|)
      float sum = 0.0f;
      for (int h = h0; h < h1; ++h)
        for (int w = w0; w < w1; ++w) {
          const float v = ldf<T>(x + base + (static_cast<long long>(h) * g.W + w) * g.C);
          sum += kind == 4 ? fabsf(v) : fmaxf(v, 0.0f);
        }
      const float pos = sum * (static_cast<float>(rnd[t]) * (1.0f / 65536.0f));
      float run = 0.0f, best = 0.0f;
      long long bo = base + (static_cast<long long>(h0) * g.W + w0) * g.C;
      bool found = false;
      for (int h = h0; h < h1 && !found; ++h)
        for (int w = w0; w < w1; ++w) {
          const long long o = base + (static_cast<long long>(h) * g.W + w) * g.C;
          const float v = ldf<T>(x + o);
          run += kind == 4 ? fabsf(v) : fmaxf(v, 0.0f);
          if (sum > 0.0f && pos <= run) {
            best = v;
            bo = o;
            found = true;
            break;
          }
        }
      stf<T>(y + t, best);
      if (offs) offs[t] = static_cast<int>(bo);
    }
  }
}

// Backward, gather form: every input element scans the (few) windows covering it.
template <typename T>
__global__ void pool_backward_kernel(int kind, const T* __restrict__ dy, const int* __restrict__ offs,
                                     T* __restrict__ dx, const PoolGeom g) {
  vb::pdl_sync();
  const long long total = static_cast<long long>(g.B) * g.H * g.W * g.C;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % g.C);
    long long r = t / g.C;
    const int w = static_cast<int>(r % g.W);
    r /= g.W;
    const int h = static_cast<int>(r % g.H);
    const int b = static_cast<int>(r / g.H);
    const int oh_lo = max(0, (h - g.ky + g.sy) / g.sy), oh_hi = min(g.OH - 1, h / g.sy);
    const int ow_lo = max(0, (w - g.kx + g.sx) / g.sx), ow_hi = min(g.OW - 1, w / g.sx);
    float acc = 0.0f;
    for (int oh = oh_lo; oh <= oh_hi; ++oh)
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const long long o = ((static_cast<long long>(b) * g.OH + oh) * g.OW + ow) * g.C + c;
        if (kind == 2) {
          const int hh = min(oh * g.sy + g.ky, g.H) - oh * g.sy, ww = min(ow * g.sx + g.kx, g.W) - ow * g.sx;
          acc += ldf<T>(dy + o) / static_cast<float>(hh * ww);
        } else if (offs[o] == static_cast<int>(t)) {
          acc += ldf<T>(dy + o);
        }
      }
    stf<T>(dx + t, acc);
  }
}

// ---------------------------------------------------------------------------
// LRN across channels (AlexNet): y = x * (k + alpha * sum_{window n} x^2)^(-beta)
// One thread per element; the n-wide window is re-read through L1 (channels are
// contiguous in NHWC).  Backward follows the closed form
//   dx_i = dy_i * s_i^-beta - 2*alpha*beta * x_i * sum_{j: i in win(j)} dy_j * x_j * s_j^(-beta-1)
// ---------------------------------------------------------------------------
template <typename T>
__global__ void lrn_forward_kernel(const T* __restrict__ x, T* __restrict__ y, long long pixels, int C, int n,
                                   float alpha, float beta, float k) {
  vb::pdl_sync();
  const long long total = pixels * C;
  const int half = n / 2;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % C);
    const T* px = x + (t - c);
    float s = 0.0f;
    const int lo = max(0, c - half), hi = min(C - 1, c + half);
    for (int j = lo; j <= hi; ++j) {
      const float v = ldf<T>(px + j);
      s += v * v;
    }
    s = k + alpha * s;
    stf<T>(y + t, ldf<T>(px + c) * __powf(s, -beta));
  }
}

template <typename T>
__global__ void lrn_backward_kernel(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx,
                                    long long pixels, int C, int n, float alpha, float beta, float k) {
  vb::pdl_sync();
  const long long total = pixels * C;
  const int half = n / 2;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % C);
    const T* px = x + (t - c);
    const T* pdy = dy + (t - c);
    const int lo = max(0, c - half), hi = min(C - 1, c + half);
    float acc = 0.0f, si = 0.0f;
    for (int j = lo; j <= hi; ++j) {
      // s_j over its own window
      const int lo2 = max(0, j - half), hi2 = min(C - 1, j + half);
      float s = 0.0f;
      for (int q = lo2; q <= hi2; ++q) {
        const float v = ldf<T>(px + q);
        s += v * v;
      }
      s = k + alpha * s;
      if (j == c) si = s;
      acc += ldf<T>(pdy + j) * ldf<T>(px + j) * __powf(s, -beta - 1.0f);
    }
    const float xi = ldf<T>(px + c);
    stf<T>(dx + t, ldf<T>(pdy + c) * __powf(si, -beta) - 2.0f * alpha * beta * xi * acc);
  }
}

// ---------------------------------------------------------------------------
// S1: softmax + argmax + EvaluatorSoftmax in one pass, one warp per sample.
//   y = softmax(s)              max_idx[b] = argmax
//   err_output = (y - onehot(label)) * err_scale for label >= 0, else 0
//   n_err[0] += (argmax != label), n_err[1] += valid; confusion[argmax*N + label]++
//   max_err_sum = max over samples of sum_j |err_output|; loss_sum += -log(y[label])
// ---------------------------------------------------------------------------
template <typename TIN, typename TOUT>
__global__ void softmax_eval_kernel(const TIN* __restrict__ s, long long lds, TOUT* __restrict__ y, long long ldy,
                                    TOUT* __restrict__ err, long long lde, const int* __restrict__ labels,
                                    int* __restrict__ max_idx, int batch, int N, float err_scale,
                                    int* __restrict__ n_err, int* __restrict__ confusion,
                                    float* __restrict__ max_err_sum, float* __restrict__ loss_sum) {
  vb::pdl_sync();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= batch) return;
  const TIN* row = s + static_cast<long long>(warp) * lds;
  float m = -INFINITY;
  int mi = 0;
  for (int j = lane; j < N; j += 32) {
    const float v = ldf<TIN>(row + j);
    if (v > m) {
      m = v;
      mi = j;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, m, o);
    const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
    if (om > m || (om == m && oi < mi)) {
      m = om;
      mi = oi;
    }
  }
  float sum = 0.0f;
  for (int j = lane; j < N; j += 32) sum += __expf(ldf<TIN>(row + j) - m);
  sum = warp_sum(sum);
  const float inv = 1.0f / sum;
  const int label = labels ? labels[warp] : -1;
  float esum = 0.0f;
  for (int j = lane; j < N; j += 32) {
    const float p = __expf(ldf<TIN>(row + j) - m) * inv;
    if (y) stf<TOUT>(y + static_cast<long long>(warp) * ldy + j, p);
    if (err) {
      const float e = label >= 0 ? (p - (j == label ? 1.0f : 0.0f)) * err_scale : 0.0f;
      stf<TOUT>(err + static_cast<long long>(warp) * lde + j, e);
      esum += fabsf(e);
    }
  }
  esum = warp_sum(esum);
  if (lane == 0) {
    if (max_idx) max_idx[warp] = mi;
    if (label >= 0) {
      if (n_err) {
        if (mi != label) atomicAdd(n_err, 1);
        atomicAdd(n_err + 1, 1);
      }
      if (confusion) atomicAdd(confusion + mi * N + label, 1);
      if (max_err_sum) atomicMax(reinterpret_cast<int*>(max_err_sum), __float_as_int(esum));  // esum >= 0
      if (loss_sum) {
        const float pl = __expf(ldf<TIN>(row + label) - m) * inv;
        atomicAdd(loss_sum, -__logf(fmaxf(pl, 1e-30f)));
      }
    }
  }
}

// EvaluatorMSE: err = (y - target) * err_scale for valid rows (row < batch_valid);
// metrics[0] += sum_b rmse_b, metrics[1] = max rmse_b, metrics[2] = min rmse_b.
template <typename T>
__global__ void mse_eval_kernel(const T* __restrict__ y, long long ldy, const T* __restrict__ target, long long ldt,
                                T* __restrict__ err, long long lde, int batch, int batch_valid, int N,
                                float err_scale, float* __restrict__ metrics, float* __restrict__ mse_per_sample) {
  vb::pdl_sync();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= batch) return;
  const bool valid = warp < batch_valid;
  float ss = 0.0f;
  const T* yr = y + static_cast<long long>(warp) * ldy;
  const T* tr = target + static_cast<long long>(warp) * ldt;
  T* er = err ? err + static_cast<long long>(warp) * lde : nullptr;
  int j0 = 0;
  if (sizeof(T) == 2 && ((ldy | ldt | lde) & 7) == 0 &&
      ((reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(target) | reinterpret_cast<uintptr_t>(err)) & 15) == 0) {
    // bf16 rows with 16-byte alignment: 8 elements per lane per step
    const int nv = N & ~7;
    for (int j = lane * 8; j < nv; j += 256) {
      const uint4 a = valid ? __ldg(reinterpret_cast<const uint4*>(yr + j)) : make_uint4(0, 0, 0, 0);
      const uint4 b = valid ? __ldg(reinterpret_cast<const uint4*>(tr + j)) : make_uint4(0, 0, 0, 0);
      const uint32_t au[4] = {a.x, a.y, a.z, a.w}, bu[4] = {b.x, b.y, b.z, b.w};
      uint32_t ou[4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float e0 = bf16_lo(au[q]) - bf16_lo(bu[q]), e1 = bf16_hi(au[q]) - bf16_hi(bu[q]);
        ss += e0 * e0 + e1 * e1;
        ou[q] = pack_bf16x2(e0 * err_scale, e1 * err_scale);
      }
      if (er) *reinterpret_cast<uint4*>(er + j) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
    }
    j0 = nv;
  }
  for (int j = j0 + lane; j < N; j += 32) {
    float e = 0.0f;
    if (valid) e = ldf<T>(yr + j) - ldf<T>(tr + j);
    ss += e * e;
    if (er) stf<T>(er + j, e * err_scale);
  }
  ss = warp_sum(ss);
  if (lane == 0) {
    const float rmse = sqrtf(ss / static_cast<float>(N));
    if (mse_per_sample) mse_per_sample[warp] = valid ? rmse : 0.0f;
    if (valid && metrics) {
      atomicAdd(metrics, rmse);
      atomicMax(reinterpret_cast<int*>(metrics + 1), __float_as_int(rmse));
      atomicMin(reinterpret_cast<int*>(metrics + 2), __float_as_int(rmse));
    }
  }
}

// ---------------------------------------------------------------------------
// R1: matrix_reduce.  Row sums: one warp per row.  Column sums (bias gradient
// = sum over the batch of err_output): two-phase — a 2-D grid writes per-slab
// partial sums into `scratch[slabs, cols]`, then one pass folds the slabs.
// Deterministic (no atomics).
// ---------------------------------------------------------------------------
template <typename T>
__global__ void row_sum_kernel(const T* __restrict__ a, long long lda, int rows, int cols, float* __restrict__ out) {
  vb::pdl_sync();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= rows) return;
  float s = 0.0f;
  for (int j = lane; j < cols; j += 32) s += ldf<T>(a + static_cast<long long>(warp) * lda + j);
  s = warp_sum(s);
  if (lane == 0) out[warp] = s;
}

constexpr int COLSUM_ROWS_PER_SLAB = 64;

// Rows per slab actually used: at least COLSUM_ROWS_PER_SLAB, grown so that the second pass folds
// at most ~512 partial rows (a 774400-row conv err_output would otherwise leave 12100 of them).
static inline int colsum_rows_per_slab(int rows) {
  int rps = COLSUM_ROWS_PER_SLAB;
  const int want = (rows + 511) / 512;
  if (want > rps) rps = (want + COLSUM_ROWS_PER_SLAB - 1) / COLSUM_ROWS_PER_SLAB * COLSUM_ROWS_PER_SLAB;
  return rps;
}

template <typename T>
__global__ void col_sum_partial_kernel(const T* __restrict__ a, long long lda, int rows, int cols,
                                       float* __restrict__ scratch, int rows_per_slab) {
  vb::pdl_sync();
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  const int slab = blockIdx.y;
  if (col >= cols) return;
  const int r0 = slab * rows_per_slab, r1 = min(rows, r0 + rows_per_slab);
  float s = 0.0f;
  for (int r = r0; r < r1; ++r) s += ldf<T>(a + static_cast<long long>(r) * lda + col);
  scratch[static_cast<long long>(slab) * cols + col] = s;
}

// bf16 fast path of the first phase: 16-byte loads.  A block is RL row-lanes x VCB 8-column vectors
// (VCB = min(cols/8, 32), RL = 256 / VCB); each thread accumulates 8 columns over every RL-th row of
// the slab, the row-lanes are folded through shared memory.  Same scratch layout as the scalar kernel.
__global__ void __launch_bounds__(256)
col_sum_partial_vec8_kernel(const __nv_bfloat16* __restrict__ a, long long lda, int rows, int cols,
                            float* __restrict__ scratch, int rows_per_slab, int vcb, int rl) {
  vb::pdl_sync();
  __shared__ float red[256 * 8];
  const int vc_local = threadIdx.x % vcb;
  const int lane_r = threadIdx.x / vcb;
  const int vc = blockIdx.x * vcb + vc_local;
  const int slab = blockIdx.y;
  const int r0 = slab * rows_per_slab, r1 = min(rows, r0 + rows_per_slab);
  float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (lane_r < rl && vc * 8 < cols) {
    const __nv_bfloat16* base = a + vc * 8;
    int r = r0 + lane_r;
    // four independent 16-byte loads in flight per thread (a slab of a large err_output is ~70 rows per thread:
    // one load at a time leaves the kernel latency-bound at a third of the copy bandwidth)
    for (; r + 3 * rl < r1; r += 4 * rl) {
      uint4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldg(reinterpret_cast<const uint4*>(base + static_cast<long long>(r + u * rl) * lda));
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        acc[0] += bf16_lo(v[u].x); acc[1] += bf16_hi(v[u].x); acc[2] += bf16_lo(v[u].y); acc[3] += bf16_hi(v[u].y);
        acc[4] += bf16_lo(v[u].z); acc[5] += bf16_hi(v[u].z); acc[6] += bf16_lo(v[u].w); acc[7] += bf16_hi(v[u].w);
      }
    }
    for (; r < r1; r += rl) {
      const uint4 v = __ldg(reinterpret_cast<const uint4*>(base + static_cast<long long>(r) * lda));
      acc[0] += bf16_lo(v.x); acc[1] += bf16_hi(v.x); acc[2] += bf16_lo(v.y); acc[3] += bf16_hi(v.y);
      acc[4] += bf16_lo(v.z); acc[5] += bf16_hi(v.z); acc[6] += bf16_lo(v.w); acc[7] += bf16_hi(v.w);
    }
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) red[threadIdx.x * 8 + q] = acc[q];
  __syncthreads();
  // thread t < vcb*8 folds column t of this block over the row-lanes
  for (int t = threadIdx.x; t < vcb * 8; t += blockDim.x) {
    const int col = blockIdx.x * vcb * 8 + t;
    if (col >= cols) continue;
    float s2 = 0.0f;
    for (int l = 0; l < rl; ++l) s2 += red[(l * vcb + (t >> 3)) * 8 + (t & 7)];
    scratch[static_cast<long long>(slab) * cols + col] = s2;
  }
}

// Optional gradient-descent step applied by the second phase to the vector the column sums are the gradient of (a
// layer's bias): saves the separate, latency-bound update launch per layer.
struct ColSumGD {
  float* w;       // null: plain column sum
  float* vel;
  float lr, wd, l12, moment;
};

// Second phase: 8 columns x 32 slab-lanes per block (the partial rows are folded 32 at a time: with up to ~512
// slabs the former 8-lane version was a chain of 64 dependent loads per thread).
__global__ void __launch_bounds__(256)
col_sum_final_kernel(const float* __restrict__ scratch, int slabs, int cols, float* __restrict__ out, float alpha,
                     int accumulate, const ColSumGD gd) {
  vb::pdl_sync();
  __shared__ float red[32][9];
  const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;
  const int col = blockIdx.x * 8 + tx;
  float s = 0.0f;
  if (col < cols)
    for (int k = ty; k < slabs; k += 32) s += scratch[static_cast<long long>(k) * cols + col];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && col < cols) {
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 32; ++k) t += red[k][tx];
    t *= alpha;
    if (gd.w != nullptr) {
      // the bias-gradient buffer stays zero (what the separate update leaves behind with zero_grad)
      const float g = (accumulate ? out[col] + t : t);
      const float w = gd.w[col];
      const float sgn = (w > 0.0f) ? 1.0f : ((w < 0.0f) ? -1.0f : 0.0f);
      const float gg = g + gd.wd * ((1.0f - gd.l12) * w + 0.5f * gd.l12 * sgn);
      const float v = gd.moment * gd.vel[col] - gd.lr * gg;
      gd.vel[col] = v;
      gd.w[col] = w + v;
      if (accumulate) out[col] = 0.0f;
    } else {
      out[col] = accumulate ? out[col] + t : t;
    }
  }
}

// Launches both phases (shared by vb_col_sum and vb_col_sum_static).
static int launch_col_sum(const void* a, int dtype, long long lda, int rows, int cols, float* out, float alpha,
                          int accumulate, float* scratch, int rps, cudaStream_t s, ColSumGD gd = ColSumGD{}) {
  const int slabs = rows > 0 ? ceil_div(rows, rps) : 0;
  if (slabs > 0) {
    if (dtype == D_BF16 && cols % 8 == 0 && lda % 8 == 0 && (reinterpret_cast<uintptr_t>(a) & 15) == 0) {
      const int vcols = cols / 8;
      // 32 vector-columns (512 B of a row) per block: small-batch FC layers then still launch >= one block per SM
      const int vcb = vcols < 32 ? vcols : 32;
      const int rl = 256 / vcb;
      dim3 grid(ceil_div(vcols, vcb), slabs);
      vb::launch_k(col_sum_partial_vec8_kernel, dim3(grid), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(a), lda, rows, cols, scratch,
                                                       rps, vcb, rl);
    } else {
      dim3 grid(ceil_div(cols, 128), slabs);
      if (dtype == D_F32)
        vb::launch_k(col_sum_partial_kernel<float>, dim3(grid), dim3(128), 0, s, static_cast<const float*>(a), lda, rows, cols, scratch, rps);
      else
        vb::launch_k(col_sum_partial_kernel<__nv_bfloat16>, dim3(grid), dim3(128), 0, s, static_cast<const __nv_bfloat16*>(a), lda, rows, cols, scratch, rps);
    }
    VB_CHECK_CUDA(cudaGetLastError());
  }
  vb::launch_k(col_sum_final_kernel, dim3(ceil_div(cols, 8)), dim3(256), 0, s, scratch, slabs, cols, out, alpha, accumulate, gd);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

}  // namespace vb

using namespace vb;

static int check_geom(const ConvGeom& g) {
  VB_REQUIRE(g.groups >= 1 && g.C % g.groups == 0, "conv: channels %d not divisible by groups %d", g.C, g.groups);
  VB_REQUIRE(g.ld >= static_cast<long long>(g.ky) * g.kx * g.C, "conv: cols pitch %lld < ky*kx*C", g.ld);
  return 0;
}

extern "C" int vb_im2col(const void* x, void* cols, int B, int H, int W, int C, int OH, int OW, int ky, int kx,
                         int sy, int sx, int pad_top, int pad_left, int groups, long long ld, cudaStream_t s) {
  ConvGeom g{B, H, W, C, OH, OW, ky, kx, sy, sx, pad_top, pad_left, groups, ld};
  if (int rc = check_geom(g)) return rc;
  const long long rows = static_cast<long long>(B) * OH * OW;
  if (rows == 0) return 0;
  const int Cg = C / groups;
  if (Cg % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0 && (reinterpret_cast<uintptr_t>(cols) & 15) == 0) {
    const long long total = rows * (ld >> 3);
    vb::launch_k(im2col_vec8_kernel, dim3(grid_cap(total, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(cols), g);
  } else {
    FastConvGeom fg{B, H, W, C, OH, OW, ky, kx, sy, sx, pad_top, pad_left, groups, ld};
    const int fr = fast_im2col(x, cols, fg, s);
    if (fr != 0) return fr < 0 ? fr : 0;
    const long long total = rows * ld;
    vb::launch_k(im2col_scalar_kernel, dim3(grid_cap(total, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(cols), g);
  }
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_col2im(const void* dcols, void* dx, int B, int H, int W, int C, int OH, int OW, int ky, int kx,
                         int sy, int sx, int pad_top, int pad_left, int groups, long long ld, const void* z,
                         int deriv, cudaStream_t s) {
  ConvGeom g{B, H, W, C, OH, OW, ky, kx, sy, sx, pad_top, pad_left, groups, ld};
  if (int rc = check_geom(g)) return rc;
  const long long total = static_cast<long long>(B) * H * W * C;
  if (total == 0) return 0;
  {
    FastConvGeom fg{B, H, W, C, OH, OW, ky, kx, sy, sx, pad_top, pad_left, groups, ld};
    const int fr = fast_col2im(dcols, dx, fg, z, deriv, s);
    if (fr != 0) return fr < 0 ? fr : 0;
  }
  vb::launch_k(col2im_kernel, dim3(grid_cap(total, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(dcols), static_cast<__nv_bfloat16*>(dx), g,
                                                      static_cast<const __nv_bfloat16*>(z), deriv);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_pool_forward(int kind, const void* x, void* y, int* offs, int dtype, int B, int H, int W, int C,
                               int OH, int OW, int ky, int kx, int sy, int sx, const void* rnd_u16, cudaStream_t s) {
  VB_REQUIRE(kind >= 0 && kind <= 4, "vb_pool_forward: bad kind %d", kind);
  VB_REQUIRE(kind < 3 || rnd_u16 != nullptr, "vb_pool_forward: stochastic pooling needs random input");
  PoolGeom g{B, H, W, C, OH, OW, ky, kx, sy, sx};
  const long long total = static_cast<long long>(B) * OH * OW * C;
  if (total == 0) return 0;
  VB_REQUIRE(static_cast<long long>(B) * H * W * C < 2147483647LL, "vb_pool_forward: input too large for int offsets");
  if (dtype == D_BF16) {
    FastPoolGeom fg{B, H, W, C, OH, OW, ky, kx, sy, sx};
    const int fr = fast_pool_forward(kind, x, y, offs, nullptr, fg, s);
    if (fr != 0) return fr < 0 ? fr : 0;
  }
  if (dtype == D_F32)
    vb::launch_k(pool_forward_kernel<float>, dim3(grid_cap(total, 256)), dim3(256), 0, s, kind, static_cast<const float*>(x), static_cast<float*>(y), offs, g, static_cast<const uint16_t*>(rnd_u16));
  else
    vb::launch_k(pool_forward_kernel<__nv_bfloat16>, dim3(grid_cap(total, 256)), dim3(256), 0, s, kind, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), offs, g, static_cast<const uint16_t*>(rnd_u16));
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_activation_backward(int act, const void* err_out, const void* x, const void* y, void* err_in,
                                      long long n, int dtype, float factor, cudaStream_t s);

// Pooling backward with the optional fused derivative of the activation that produced the pooling
// input z (err_input *= f'(z)): one pass on the fast path, generic kernel + element-wise pass otherwise.
extern "C" int vb_pool_backward_fused(int kind, const void* dy, const int* offs, void* dx, int dtype, int B, int H,
                                      int W, int C, int OH, int OW, int ky, int kx, int sy, int sx, const void* z,
                                      int deriv, cudaStream_t s);

extern "C" int vb_pool_backward(int kind, const void* dy, const int* offs, void* dx, int dtype, int B, int H, int W,
                                int C, int OH, int OW, int ky, int kx, int sy, int sx, cudaStream_t s) {
  return vb_pool_backward_fused(kind, dy, offs, dx, dtype, B, H, W, C, OH, OW, ky, kx, sy, sx, nullptr, 0, s);
}

extern "C" int vb_pool_backward_fused(int kind, const void* dy, const int* offs, void* dx, int dtype, int B, int H,
                                      int W, int C, int OH, int OW, int ky, int kx, int sy, int sx, const void* z,
                                      int deriv, cudaStream_t s) {
  VB_REQUIRE(kind == 2 || offs != nullptr, "vb_pool_backward: max-like pooling needs the offsets recorded in forward");
  PoolGeom g{B, H, W, C, OH, OW, ky, kx, sy, sx};
  const long long total = static_cast<long long>(B) * H * W * C;
  if (total == 0) return 0;
  if (dtype == D_BF16) {
    FastPoolGeom fg{B, H, W, C, OH, OW, ky, kx, sy, sx};
    const int fr = fast_pool_backward(kind, dy, offs, nullptr, dx, fg, (z && deriv) ? z : nullptr, deriv, s);
    if (fr != 0) return fr < 0 ? fr : 0;
  }
  if (dtype == D_F32)
    vb::launch_k(pool_backward_kernel<float>, dim3(grid_cap(total, 256)), dim3(256), 0, s, kind, static_cast<const float*>(dy), offs, static_cast<float*>(dx), g);
  else
    vb::launch_k(pool_backward_kernel<__nv_bfloat16>, dim3(grid_cap(total, 256)), dim3(256), 0, s, kind, static_cast<const __nv_bfloat16*>(dy), offs, static_cast<__nv_bfloat16*>(dx), g);
  VB_CHECK_CUDA(cudaGetLastError());
  if (z && deriv) return vb_activation_backward(deriv, dx, nullptr, z, dx, total, dtype, 1.0f, s);
  return 0;
}

// ---- pooling with compact argmax records (bf16, C % 8 == 0, ky*kx <= 255) ---------------------------------
// The backward pass only needs to know WHICH element of its window won, so the hot path keeps one byte per
// output (window-local index) instead of the 4-byte global offset; the Znicz-visible int32 ``input_offset``
// is expanded from it on demand.  Each returns 1 when the shape is not supported by this path (the caller
// then uses vb_pool_forward / vb_pool_backward_fused), 0 on success, negative on error.
extern "C" int vb_pool_forward_local(int kind, const void* x, void* y, int* offs, unsigned char* offs8, int B, int H,
                                     int W, int C, int OH, int OW, int ky, int kx, int sy, int sx, cudaStream_t s) {
  VB_REQUIRE(kind >= 0 && kind <= 1 && offs8 != nullptr, "vb_pool_forward_local: max / maxabs pooling with an offs8 buffer");
  FastPoolGeom fg{B, H, W, C, OH, OW, ky, kx, sy, sx};
  const int fr = fast_pool_forward(kind, x, y, offs, offs8, fg, s);
  return fr < 0 ? fr : (fr == 1 ? 0 : 1);
}

extern "C" int vb_pool_backward_local(int kind, const void* dy, const unsigned char* offs8, void* dx, int B, int H,
                                      int W, int C, int OH, int OW, int ky, int kx, int sy, int sx, const void* z,
                                      int deriv, cudaStream_t s) {
  VB_REQUIRE(kind >= 0 && kind <= 1 && offs8 != nullptr, "vb_pool_backward_local: max / maxabs pooling with an offs8 buffer");
  FastPoolGeom fg{B, H, W, C, OH, OW, ky, kx, sy, sx};
  const int fr = fast_pool_backward(kind, dy, nullptr, offs8, dx, fg, (z && deriv) ? z : nullptr, deriv, s);
  return fr < 0 ? fr : (fr == 1 ? 0 : 1);
}

extern "C" int vb_pool_expand_offsets(const unsigned char* offs8, int* offs, int B, int H, int W, int C, int OH,
                                      int OW, int ky, int kx, int sy, int sx, cudaStream_t s) {
  FastPoolGeom fg{B, H, W, C, OH, OW, ky, kx, sy, sx};
  const int fr = fast_pool_expand_offsets(offs8, offs, fg, s);
  VB_REQUIRE(fr != 0, "vb_pool_expand_offsets: tensor too large for int32 offsets");
  return fr < 0 ? fr : 0;
}

extern "C" int vb_lrn_forward(const void* x, void* y, int dtype, long long pixels, int C, int n, float alpha,
                              float beta, float k, cudaStream_t s) {
  const long long total = pixels * C;
  if (total == 0) return 0;
  if (dtype == D_BF16) {
    const int fr = fast_lrn_forward(x, y, pixels, C, n, alpha, beta, k, s);
    if (fr != 0) return fr < 0 ? fr : 0;
  }
  if (dtype == D_F32)
    vb::launch_k(lrn_forward_kernel<float>, dim3(grid_cap(total, 256)), dim3(256), 0, s, static_cast<const float*>(x), static_cast<float*>(y), pixels, C, n, alpha, beta, k);
  else
    vb::launch_k(lrn_forward_kernel<__nv_bfloat16>, dim3(grid_cap(total, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), pixels, C, n, alpha, beta, k);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// LRN backward with the optional fused derivative of the activation that produced x (dx *= f'(x)).
extern "C" int vb_lrn_backward_fused(const void* x, const void* dy, void* dx, int dtype, long long pixels, int C, int n,
                                     float alpha, float beta, float k, int deriv, cudaStream_t s);

extern "C" int vb_lrn_backward(const void* x, const void* dy, void* dx, int dtype, long long pixels, int C, int n,
                               float alpha, float beta, float k, cudaStream_t s) {
  return vb_lrn_backward_fused(x, dy, dx, dtype, pixels, C, n, alpha, beta, k, 0, s);
}

extern "C" int vb_lrn_backward_fused(const void* x, const void* dy, void* dx, int dtype, long long pixels, int C, int n,
                                     float alpha, float beta, float k, int deriv, cudaStream_t s) {
  const long long total = pixels * C;
  if (total == 0) return 0;
  if (dtype == D_BF16) {
    const int fr = fast_lrn_backward(x, dy, dx, pixels, C, n, alpha, beta, k, deriv, s);
    if (fr != 0) return fr < 0 ? fr : 0;
  }
  if (dtype == D_F32)
    vb::launch_k(lrn_backward_kernel<float>, dim3(grid_cap(total, 256)), dim3(256), 0, s, static_cast<const float*>(x), static_cast<const float*>(dy), static_cast<float*>(dx), pixels, C, n, alpha, beta, k);
  else
    vb::launch_k(lrn_backward_kernel<__nv_bfloat16>, dim3(grid_cap(total, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx), pixels, C, n, alpha, beta, k);
  VB_CHECK_CUDA(cudaGetLastError());
  if (deriv) return vb_activation_backward(deriv, dx, nullptr, x, dx, total, dtype, 1.0f, s);
  return 0;
}

// LRN backward that also leaves the per-channel sums of dx as partial rows partials[blocks][C] (fp32; *blocks_out rows,
// at most max_blocks): the bias gradient of the convolution in front of the LRN, to be folded by vb_col_sum_finish.
// bf16 only; returns 1 when the shape is not covered (run vb_lrn_backward_fused and a column sum instead), 0 on
// success.
extern "C" int vb_lrn_backward_colsum(const void* x, const void* dy, void* dx, long long pixels, int C, int n,
                                      float alpha, float beta, float k, int deriv, float* partials, int max_blocks,
                                      int* blocks_out, cudaStream_t s) {
  VB_REQUIRE(partials != nullptr && blocks_out != nullptr && max_blocks >= 1, "vb_lrn_backward_colsum: no room for partial sums");
  *blocks_out = 0;
  if (pixels * C == 0) return 0;
  const int fr = fast_lrn_backward(x, dy, dx, pixels, C, n, alpha, beta, k, deriv, s, partials, max_blocks, blocks_out);
  return fr < 0 ? fr : (fr == 0 ? 1 : 0);
}

// Second phase of the column-sum kernels on its own: grad[c] (+)= sum over `slabs` partial rows of scratch[slab][c];
// with w / vel given, the momentum GD step of the bias is applied instead and grad is left zeroed (vb_col_sum_gd).
extern "C" int vb_col_sum_finish(const float* scratch, int slabs, int cols, float* grad, int accumulate, float* w,
                                 float* vel, float lr, float wd, float l1_vs_l2, float moment, cudaStream_t s) {
  VB_REQUIRE(scratch != nullptr && grad != nullptr && slabs >= 0 && cols >= 1, "vb_col_sum_finish: bad arguments");
  VB_REQUIRE((w == nullptr) == (vel == nullptr), "vb_col_sum_finish: w and vel go together");
  ColSumGD gd{w, vel, lr, wd, l1_vs_l2, moment};
  vb::launch_k(col_sum_final_kernel, dim3(ceil_div(cols, 8)), dim3(256), 0, s, scratch, slabs, cols, grad, 1.0f, accumulate, gd);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Max-pooling backward fused into LRN backward (bf16 NHWC; 3x3 / stride 2 max pooling over the LRN output):
//   dx = LRN'(x) applied to the pooling gradient scattered by the argmax bytes, optionally times f'(x).
// x, dx: [B, H, W, C]; dpool, offs8: [B, OH, OW, C].  Returns 1 when the geometry is not supported (run the two
// kernels separately), 0 on success, negative on error.
extern "C" int vb_lrn_pool_backward(const void* x, const void* dpool, const unsigned char* offs8, void* dx, int B, int H,
                                    int W, int C, int OH, int OW, int ky, int kx, int sy, int sx, int n, float alpha,
                                    float beta, float k, int deriv, cudaStream_t s) {
  FastPoolGeom g{B, H, W, C, OH, OW, ky, kx, sy, sx};
  const int fr = fast_lrn_pool_backward(x, dpool, offs8, dx, g, n, alpha, beta, k, deriv, s);
  if (fr < 0) return fr;
  return fr == 1 ? 0 : 1;
}

// in_dtype: dtype of the pre-softmax activations s; out_dtype: dtype of y / err_output.
extern "C" int vb_softmax_eval(const void* sm_in, int in_dtype, long long lds, void* y, long long ldy, void* err,
                               long long lde, int out_dtype, const int* labels, int* max_idx, int batch, int N,
                               float err_scale, int* n_err, int* confusion, float* max_err_sum, float* loss_sum,
                               cudaStream_t s) {
  if (batch == 0) return 0;
  const int threads = 128;
  const int grid = ceil_div(batch * 32, threads);
#define VB_SM(TIN, TOUT)                                                                                      \
  vb::launch_k(softmax_eval_kernel<TIN, TOUT>, dim3(grid), dim3(threads), 0, s, static_cast<const TIN*>(sm_in), lds,                \
                                                          static_cast<TOUT*>(y), ldy, static_cast<TOUT*>(err), \
                                                          lde, labels, max_idx, batch, N, err_scale, n_err,   \
                                                          confusion, max_err_sum, loss_sum)
  if (in_dtype == D_F32 && out_dtype == D_F32) VB_SM(float, float);
  else if (in_dtype == D_F32 && out_dtype == D_BF16) VB_SM(float, __nv_bfloat16);
  else if (in_dtype == D_BF16 && out_dtype == D_BF16) VB_SM(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == D_BF16 && out_dtype == D_F32) VB_SM(__nv_bfloat16, float);
  else VB_REQUIRE(false, "vb_softmax_eval: unsupported dtypes %d/%d", in_dtype, out_dtype);
#undef VB_SM
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_mse_eval(const void* y, long long ldy, const void* target, long long ldt, void* err, long long lde,
                           int dtype, int batch, int batch_valid, int N, float err_scale, float* metrics,
                           float* mse_per_sample, cudaStream_t s) {
  if (batch == 0) return 0;
  const int threads = 128;
  const int grid = ceil_div(batch * 32, threads);
  if (dtype == D_F32)
    vb::launch_k(mse_eval_kernel<float>, dim3(grid), dim3(threads), 0, s, static_cast<const float*>(y), ldy, static_cast<const float*>(target), ldt,
                                                    static_cast<float*>(err), lde, batch, batch_valid, N, err_scale, metrics, mse_per_sample);
  else
    vb::launch_k(mse_eval_kernel<__nv_bfloat16>, dim3(grid), dim3(threads), 0, s, static_cast<const __nv_bfloat16*>(y), ldy,
                                                            static_cast<const __nv_bfloat16*>(target), ldt,
                                                            static_cast<__nv_bfloat16*>(err), lde, batch, batch_valid, N,
                                                            err_scale, metrics, mse_per_sample);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_row_sum(const void* a, int dtype, long long lda, int rows, int cols, float* out, cudaStream_t s) {
  if (rows == 0) return 0;
  const int grid = ceil_div(rows * 32, 128);
  if (dtype == D_F32) vb::launch_k(row_sum_kernel<float>, dim3(grid), dim3(128), 0, s, static_cast<const float*>(a), lda, rows, cols, out);
  else vb::launch_k(row_sum_kernel<__nv_bfloat16>, dim3(grid), dim3(128), 0, s, static_cast<const __nv_bfloat16*>(a), lda, rows, cols, out);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Number of floats of scratch vb_col_sum needs for a (rows x cols) input.
extern "C" long long vb_col_sum_scratch_elems(int rows, int cols) {
  return static_cast<long long>(ceil_div(rows > 0 ? rows : 1, COLSUM_ROWS_PER_SLAB)) * cols;
}

// out[j] (+)= alpha * sum_i a[i,j].  `scratch` must hold vb_col_sum_scratch_elems() floats.
extern "C" int vb_col_sum(const void* a, int dtype, long long lda, int rows, int cols, float* out, float alpha,
                          int accumulate, float* scratch, cudaStream_t s) {
  if (cols == 0) return 0;
  VB_REQUIRE(scratch != nullptr, "vb_col_sum: scratch buffer is required");
  return launch_col_sum(a, dtype, lda, rows, cols, out, alpha, accumulate, scratch, colsum_rows_per_slab(rows), s);
}

// Column sum whose result is consumed on the spot as the gradient of `w` (a bias vector): momentum GD step
//   g = alpha * sum_i a[i,j] (+ grad[j] when accumulate) + wd * ((1 - l1_vs_l2) * w + 0.5 * l1_vs_l2 * sign(w));
//   v = moment * v - lr * g;  w += v.   `grad` (may be null when !accumulate) is left zeroed.
extern "C" int vb_col_sum_gd(const void* a, int dtype, long long lda, int rows, int cols, float* grad, float alpha,
                             int accumulate, float* scratch, float* w, float* vel, float lr, float wd, float l1_vs_l2,
                             float moment, cudaStream_t s) {
  if (cols == 0) return 0;
  VB_REQUIRE(scratch != nullptr && w != nullptr && vel != nullptr, "vb_col_sum_gd: scratch, w and vel are required");
  VB_REQUIRE(!accumulate || grad != nullptr, "vb_col_sum_gd: accumulate needs the gradient buffer");
  ColSumGD gd{w, vel, lr, wd, l1_vs_l2, moment};
  return launch_col_sum(a, dtype, lda, rows, cols, grad, alpha, accumulate, scratch, colsum_rows_per_slab(rows), s, gd);
}

// ---------------------------------------------------------------------------
// Allocation-free column sum.
//
// Some callers of this library are not allowed to allocate device memory: a host
// application that hands us a fixed memory budget up front, code running between
// cudaStreamBeginCapture/EndCapture, or a real-time inference loop where a
// cudaMalloc (an implicit device synchronisation with unbounded latency) is off
// limits.  vb_col_sum() makes such callers supply the inter-block scratch buffer
// themselves; vb_col_sum_static() below needs nothing but inputs and outputs.  Its
// scratch lives in the fixed array declared here, which every GPU that loads the
// library reserves (4 MiB) until the process exits.
//
// The price: the array is shared by every call on a device, so two calls whose
// kernels may overlap on the GPU (different streams, no ordering between them)
// race on it and both results are undefined — keep such calls on one stream or
// order them with events.  Nothing in the array has to be zero when a call starts:
// the first pass overwrites every partial sum the second pass reads, and whatever
// the previous call left behind is never looked at.
// ---------------------------------------------------------------------------
#define VB_COLSUM_STATIC_MAX_ELEMS (1 << 20)  // 1 Mi floats = 4 MiB

static __device__ __align__(256) float g_colsum_static_scratch[VB_COLSUM_STATIC_MAX_ELEMS];

// Device address of the array above, looked up once per device.  In host code the name
// g_colsum_static_scratch is only a handle for the runtime, not a pointer a kernel could
// dereference, hence cudaGetSymbolAddress.
static float* g_colsum_static_ptr[64] = {nullptr};

static int colsum_static_scratch(float** out) {
  int dev = -1;
  VB_CHECK_CUDA(cudaGetDevice(&dev));
  VB_REQUIRE(dev >= 0 && dev < 64, "vb_col_sum_static: device ordinal %d is out of the supported range", dev);
  if (g_colsum_static_ptr[dev] == nullptr) {
    float* p = nullptr;
    // the symbol is resolved for the CURRENT device (each device has its own copy)
    VB_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_scratch));
    VB_REQUIRE(p != nullptr, "vb_col_sum_static: cudaGetSymbolAddress returned a null address");
    g_colsum_static_ptr[dev] = p;
  }
  *out = g_colsum_static_ptr[dev];
  return 0;
}

extern "C" long long vb_col_sum_static_max_elems() { return VB_COLSUM_STATIC_MAX_ELEMS; }

// out[j] (+)= alpha * sum_i a[i,j] without any caller-provided or allocated scratch.
// Limits: ceil(rows / 64) * cols <= vb_col_sum_static_max_elems() (1 Mi) and rows <= 65535 * 64
// (4 194 240), otherwise the call is rejected with an error and nothing is launched.
// The first call on each device resolves the scratch address (a host-side runtime query, no
// allocation); make that first call outside of any stream capture.
extern "C" int vb_col_sum_static(const void* a, int dtype, long long lda, int rows, int cols, float* out,
                                 float alpha, int accumulate, cudaStream_t s) {
  if (cols == 0) return 0;
  VB_REQUIRE(rows >= 0 && cols > 0, "vb_col_sum_static: bad shape %d x %d", rows, cols);
  const int slabs = ceil_div(rows > 0 ? rows : 1, COLSUM_ROWS_PER_SLAB);
  const long long need = static_cast<long long>(slabs) * cols;
  VB_REQUIRE(need <= VB_COLSUM_STATIC_MAX_ELEMS,
             "vb_col_sum_static: %d x %d needs %lld scratch floats, the built-in buffer holds %d; "
             "use vb_col_sum() with your own scratch for inputs this large",
             rows, cols, need, VB_COLSUM_STATIC_MAX_ELEMS);
  VB_REQUIRE(slabs <= 65535,
             "vb_col_sum_static: %d rows exceed the limit of %d rows (65535 slabs of %d rows); "
             "use vb_col_sum() with your own scratch for inputs this tall",
             rows, 65535 * COLSUM_ROWS_PER_SLAB, COLSUM_ROWS_PER_SLAB);
  VB_REQUIRE(dtype == D_F32 || dtype == D_BF16, "vb_col_sum_static: dtype must be f32 (0) or bf16 (1)");
  float* scratch = nullptr;
  if (int rc = colsum_static_scratch(&scratch)) return rc;
  return launch_col_sum(a, dtype, lda, rows, cols, out, alpha, accumulate, scratch, COLSUM_ROWS_PER_SLAB, s);
}
