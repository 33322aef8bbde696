// veles_b200 — bandwidth-tuned variants of the spatial kernels, selected by the entry points in
// nn_kernels.cu whenever the shapes allow (bf16, channel counts that are multiples of 8, ...).
// The generic kernels in nn_kernels.cu stay as the any-shape fallback and as the readable
// statement of the semantics.
//
// First measurement that motivated them (AlexNet, B=256, ncu launch list, profiles/alexnet_r1_*):
// col2im 17 %, im2col(C=3) 12.5 %, lrn_backward 12 %, pool_backward 8 %, col_sum_final 6 % of
// the step — all far from the HBM roofline because of per-element div/mod address math,
// 2-byte accesses and O(window^2) recomputation.
#include "common.cuh"
#include "fast_paths.cuh"

namespace vb {

__device__ __forceinline__ void unpack8(const uint4& v, float (&f)[8]) {
  f[0] = bf16_lo(v.x); f[1] = bf16_hi(v.x); f[2] = bf16_lo(v.y); f[3] = bf16_hi(v.y);
  f[4] = bf16_lo(v.z); f[5] = bf16_hi(v.z); f[6] = bf16_lo(v.w); f[7] = bf16_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 o;
  o.x = pack_bf16x2(f[0], f[1]); o.y = pack_bf16x2(f[2], f[3]);
  o.z = pack_bf16x2(f[4], f[5]); o.w = pack_bf16x2(f[6], f[7]);
  return o;
}
__device__ __forceinline__ uint4 ldg16(const void* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

static inline int grid_cap2(long long n, int block) {
  long long g = (n + block - 1) / block;
  const long long cap = static_cast<long long>(device_sm_count()) * 32;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return static_cast<int>(g);
}

// ---------------------------------------------------------------------------------------------
// im2col for ANY channel count: one thread produces 8 consecutive columns (one 16-byte store);
// the (group, tap, channel) decomposition is done once and then advanced incrementally.
// ---------------------------------------------------------------------------------------------
__global__ void im2col_any8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols,
                                   const FastConvGeom g) {
  vb::pdl_sync();
  const int Cg = g.C / g.groups;
  const int kk = g.ky * g.kx;
  const int vec_per_row = static_cast<int>(g.ld >> 3);
  const int used = g.groups * kk * Cg;
  const long long rows = static_cast<long long>(g.B) * g.OH * g.OW;
  const long long total = rows * vec_per_row;
  const unsigned short* xs = reinterpret_cast<const unsigned short*>(x);
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long row = t / vec_per_row;
    int col = static_cast<int>(t - row * vec_per_row) << 3;
    const int ow = static_cast<int>(row % g.OW);
    const long long r2 = row / g.OW;
    const int oh = static_cast<int>(r2 % g.OH);
    const int b = static_cast<int>(r2 / g.OH);
    int grp = col / (kk * Cg);
    int rem = col - grp * kk * Cg;
    int tap = rem / Cg;
    int cg = rem - tap * Cg;
    int i = tap / g.kx, j = tap - i * g.kx;
    const int ih0 = oh * g.sy - g.pt, iw0 = ow * g.sx - g.pl;
    const long long img = static_cast<long long>(b) * g.H * g.W * g.C;
    unsigned short v[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      unsigned short e = 0;
      if (col + q < used) {
        const int ih = ih0 + i, iw = iw0 + j;
        if (ih >= 0 && ih < g.H && iw >= 0 && iw < g.W)
          e = __ldg(xs + img + (static_cast<long long>(ih) * g.W + iw) * g.C + grp * Cg + cg);
      }
      v[q] = e;
      if (++cg == Cg) {
        cg = 0;
        if (++j == g.kx) {
          j = 0;
          if (++i == g.ky) {
            i = 0;
            ++grp;
          }
        }
      }
    }
    uint4 o;
    o.x = v[0] | (static_cast<uint32_t>(v[1]) << 16);
    o.y = v[2] | (static_cast<uint32_t>(v[3]) << 16);
    o.z = v[4] | (static_cast<uint32_t>(v[5]) << 16);
    o.w = v[6] | (static_cast<uint32_t>(v[7]) << 16);
    *reinterpret_cast<uint4*>(cols + row * g.ld + col) = o;
  }
}

// ---------------------------------------------------------------------------------------------
// im2col through shared memory for few-channel layers (groups == 1, e.g. the RGB first layer).
// One block per output row (b, oh): the ky input rows it touches are staged once (zero padded),
// then every thread assembles 16-byte column vectors from smem.  Within a filter row the kx*C
// elements of a tap row are contiguous in the staged row, so a vector is 8 two-byte smem reads
// instead of 8 global loads with per-element bounds checks.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
im2col_rowtile_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ cols, const FastConvGeom g,
                      int rw) {
  vb::pdl_sync();
  extern __shared__ unsigned short srows[];  // [ky][rw], rw = ((OW-1)*sx + kx) * C
  const int b = blockIdx.x / g.OH;
  const int oh = blockIdx.x - b * g.OH;
  const unsigned short* xs = reinterpret_cast<const unsigned short*>(x);
  const int wc = g.W * g.C;
  for (int idx = threadIdx.x; idx < g.ky * rw; idx += blockDim.x) {
    const int r = idx / rw;
    const int e = idx - r * rw - g.pl * g.C;
    const int h = oh * g.sy - g.pt + r;
    unsigned short v = 0;
    if (h >= 0 && h < g.H && e >= 0 && e < wc) v = __ldg(xs + (static_cast<long long>(b) * g.H + h) * wc + e);
    srows[idx] = v;
  }
  __syncthreads();
  const int kc = g.kx * g.C;
  const int used = g.ky * kc;
  const int vp = static_cast<int>(g.ld >> 3);
  __nv_bfloat16* out_row = cols + static_cast<long long>(blockIdx.x) * g.OW * g.ld;
  for (int v = threadIdx.x; v < g.OW * vp; v += blockDim.x) {
    const int ow = v / vp;
    const int j0 = (v - ow * vp) << 3;
    int r = j0 / kc;
    int t = j0 - r * kc;
    const unsigned short* src = srows + ow * g.sx * g.C;
    unsigned short e[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      e[q] = (j0 + q < used) ? src[r * rw + t] : static_cast<unsigned short>(0);
      if (++t == kc) {
        t = 0;
        ++r;
      }
    }
    uint4 o;
    o.x = e[0] | (static_cast<uint32_t>(e[1]) << 16);
    o.y = e[2] | (static_cast<uint32_t>(e[3]) << 16);
    o.z = e[4] | (static_cast<uint32_t>(e[5]) << 16);
    o.w = e[6] | (static_cast<uint32_t>(e[7]) << 16);
    *reinterpret_cast<uint4*>(out_row + static_cast<long long>(ow) * g.ld + j0) = o;
  }
}

// ---------------------------------------------------------------------------------------------
// col2im, gather form, 8 channels per thread (Cg % 8 == 0): 16-byte loads of the column matrix.
// ---------------------------------------------------------------------------------------------
__global__ void col2im_vec8_kernel(const __nv_bfloat16* __restrict__ dcols, __nv_bfloat16* __restrict__ dx,
                                   const FastConvGeom g, const __nv_bfloat16* __restrict__ z, int deriv) {
  vb::pdl_sync();
  const int Cg = g.C / g.groups;
  const int kk = g.ky * g.kx;
  const int C8 = g.C >> 3;
  const long long total = static_cast<long long>(g.B) * g.H * g.W * C8;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(t % C8) << 3;
    long long r = t / C8;
    const int iw = static_cast<int>(r % g.W);
    r /= g.W;
    const int ih = static_cast<int>(r % g.H);
    const int b = static_cast<int>(r / g.H);
    const int grp = c / Cg, cg = c - grp * Cg;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    const __nv_bfloat16* base = dcols + grp * kk * Cg + cg;
    for (int i = 0; i < g.ky; ++i) {
      const int nh = ih + g.pt - i;
      if (nh < 0) break;
      if (nh % g.sy) continue;
      const int oh = nh / g.sy;
      if (oh >= g.OH) continue;
      for (int j = 0; j < g.kx; ++j) {
        const int nw = iw + g.pl - j;
        if (nw < 0) break;
        if (nw % g.sx) continue;
        const int ow = nw / g.sx;
        if (ow >= g.OW) continue;
        const long long row = (static_cast<long long>(b) * g.OH + oh) * g.OW + ow;
        float f[8];
        unpack8(ldg16(base + row * g.ld + (i * g.kx + j) * Cg), f);
#pragma unroll
        for (int q = 0; q < 8; ++q) acc[q] += f[q];
      }
    }
    const long long o = ((static_cast<long long>(b) * g.H + ih) * g.W + iw) * g.C + c;
    if (z != nullptr) {
      float f[8];
      unpack8(ldg16(z + o), f);
#pragma unroll
      for (int q = 0; q < 8; ++q) acc[q] *= act_deriv_from_output(deriv, f[q]);
    }
    *reinterpret_cast<uint4*>(dx + o) = pack8(acc);
  }
}

// ---------------------------------------------------------------------------------------------
// Pooling, 8 channels per thread (C % 8 == 0), bf16.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128, 8)   // <= 64 registers: latency-bound gather, occupancy matters
pool_forward_vec8_kernel(int kind, const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                         int* __restrict__ offs, uint8_t* __restrict__ offs8, const FastPoolGeom g) {
  vb::pdl_sync();
  const unsigned C8 = g.C >> 3;
  const long long total = static_cast<long long>(g.B) * g.OH * g.OW * C8;   // < 2^31 (checked by the launcher)
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const unsigned tt = static_cast<unsigned>(t);   // 32-bit index arithmetic: 64-bit divisions cost ~100 instructions each
    const int c = static_cast<int>(tt % C8) << 3;
    unsigned r = tt / C8;
    const int ow = static_cast<int>(r % g.OW);
    r /= g.OW;
    const int oh = static_cast<int>(r % g.OH);
    const int b = static_cast<int>(r / g.OH);
    const int h0 = oh * g.sy, w0 = ow * g.sx;
    const int h1 = min(h0 + g.ky, g.H), w1 = min(w0 + g.kx, g.W);
    const long long img = static_cast<long long>(b) * g.H * g.W * g.C + c;
    float best[8], key[8];
    int bo[8];
    unsigned bl[8];   // window-local argmax (row * kx + col): what the backward pass really needs, 1 byte
    float sum[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      best[q] = 0.0f; key[q] = -INFINITY; sum[q] = 0.0f; bl[q] = 0;
      bo[q] = static_cast<int>(img + (static_cast<long long>(h0) * g.W + w0) * g.C) + q;
    }
    for (int h = h0; h < h1; ++h)
      for (int w = w0; w < w1; ++w) {
        const unsigned local = static_cast<unsigned>((h - h0) * g.kx + (w - w0));
        const long long o = img + (static_cast<long long>(h) * g.W + w) * g.C;
        float f[8];
        unpack8(ldg16(x + o), f);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (kind == 2) {
            sum[q] += f[q];
          } else {
            const float kq = kind == 1 ? fabsf(f[q]) : f[q];
            if (kq > key[q]) {
              key[q] = kq;
              best[q] = f[q];
              bo[q] = static_cast<int>(o) + q;
              bl[q] = local;
            }
          }
        }
      }
    const long long oo = ((static_cast<long long>(b) * g.OH + oh) * g.OW + ow) * g.C + c;
    if (kind == 2) {
      const float inv = 1.0f / static_cast<float>((h1 - h0) * (w1 - w0));
#pragma unroll
      for (int q = 0; q < 8; ++q) best[q] = sum[q] * inv;
    }
    *reinterpret_cast<uint4*>(y + oo) = pack8(best);
    if (offs != nullptr && kind != 2) {
      *reinterpret_cast<int4*>(offs + oo) = make_int4(bo[0], bo[1], bo[2], bo[3]);
      *reinterpret_cast<int4*>(offs + oo + 4) = make_int4(bo[4], bo[5], bo[6], bo[7]);
    }
    if (offs8 != nullptr && kind != 2) {
      *reinterpret_cast<uint2*>(offs8 + oo) =
          make_uint2(bl[0] | (bl[1] << 8) | (bl[2] << 16) | (bl[3] << 24), bl[4] | (bl[5] << 8) | (bl[6] << 16) | (bl[7] << 24));
    }
  }
}

// int32 global offsets (the Znicz-visible ``input_offset``) from the 1-byte window-local argmax.
__global__ void pool_expand_offsets_kernel(const uint8_t* __restrict__ offs8, int* __restrict__ offs,
                                           const FastPoolGeom g) {
  vb::pdl_sync();
  const long long total = static_cast<long long>(g.B) * g.OH * g.OW * g.C;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const unsigned tt = static_cast<unsigned>(t);
    const int c = static_cast<int>(tt % g.C);
    unsigned r = tt / g.C;
    const int ow = static_cast<int>(r % g.OW);
    r /= g.OW;
    const int oh = static_cast<int>(r % g.OH);
    const int b = static_cast<int>(r / g.OH);
    const int l = offs8[t];
    const int ih = oh * g.sy + l / g.kx, iw = ow * g.sx + l % g.kx;
    offs[t] = ((b * g.H + ih) * g.W + iw) * g.C + c;
  }
}

__global__ void pool_backward_vec8_kernel(int kind, const __nv_bfloat16* __restrict__ dy, const int* __restrict__ offs,
                                          const uint8_t* __restrict__ offs8, __nv_bfloat16* __restrict__ dx,
                                          const FastPoolGeom g, const __nv_bfloat16* __restrict__ z, int deriv) {
  vb::pdl_sync();
  const unsigned C8 = g.C >> 3;
  const long long total = static_cast<long long>(g.B) * g.H * g.W * C8;   // < 2^31 (checked by the launcher)
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const unsigned tt = static_cast<unsigned>(t);
    const int c = static_cast<int>(tt % C8) << 3;
    unsigned r = tt / C8;
    const int w = static_cast<int>(r % g.W);
    r /= g.W;
    const int h = static_cast<int>(r % g.H);
    const int b = static_cast<int>(r / g.H);
    const int oh_lo = max(0, (h - g.ky + g.sy) / g.sy), oh_hi = min(g.OH - 1, h / g.sy);
    const int ow_lo = max(0, (w - g.kx + g.sx) / g.sx), ow_hi = min(g.OW - 1, w / g.sx);
    const long long me = ((static_cast<long long>(b) * g.H + h) * g.W + w) * g.C + c;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int oh = oh_lo; oh <= oh_hi; ++oh)
      for (int ow = ow_lo; ow <= ow_hi; ++ow) {
        const long long o = ((static_cast<long long>(b) * g.OH + oh) * g.OW + ow) * g.C + c;
        float f[8];
        unpack8(ldg16(dy + o), f);
        if (kind == 2) {
          const int hh = min(oh * g.sy + g.ky, g.H) - oh * g.sy, ww = min(ow * g.sx + g.kx, g.W) - ow * g.sx;
          const float inv = 1.0f / static_cast<float>(hh * ww);
#pragma unroll
          for (int q = 0; q < 8; ++q) acc[q] += f[q] * inv;
        } else if (offs8 != nullptr) {
          // 8 one-byte window-local argmax indices: this input pixel is position (h - oh*sy, w - ow*sx)
          const uint2 l8 = __ldg(reinterpret_cast<const uint2*>(offs8 + o));
          const unsigned mine = static_cast<unsigned>((h - oh * g.sy) * g.kx + (w - ow * g.sx));
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (((l8.x >> (8 * q)) & 0xffu) == mine) acc[q] += f[q];
            if (((l8.y >> (8 * q)) & 0xffu) == mine) acc[4 + q] += f[4 + q];
          }
        } else {
          const int4 o0 = __ldg(reinterpret_cast<const int4*>(offs + o));
          const int4 o1 = __ldg(reinterpret_cast<const int4*>(offs + o + 4));
          const int of[8] = {o0.x, o0.y, o0.z, o0.w, o1.x, o1.y, o1.z, o1.w};
          const int mine = static_cast<int>(me);
#pragma unroll
          for (int q = 0; q < 8; ++q)
            if (of[q] == mine + q) acc[q] += f[q];
        }
      }
    if (z != nullptr) {
      float f[8];
      unpack8(ldg16(z + me), f);
#pragma unroll
      for (int q = 0; q < 8; ++q) acc[q] *= act_deriv_from_output(deriv, f[q]);
    }
    *reinterpret_cast<uint4*>(dx + me) = pack8(acc);
  }
}

// ---------------------------------------------------------------------------------------------
// 3x3 / stride 2 max (kind 0) and max-abs (kind 1) pooling, the AlexNet geometry, with compile-time window
// constants: all nine (forward) or four (backward) 16-byte loads of a thread are issued before any arithmetic and
// the index arithmetic has no run-time divisors besides the tensor extents.
// Backward is organised by 2x2 input blocks: block (bh, bw) = pixels (2bh+i, 2bw+j) is covered exactly by the
// windows (bh-1..bh) x (bw-1..bw), so four dy / argmax loads serve four input pixels (the per-pixel gather form
// loads up to four windows for every single pixel: 9 loads per block instead of 4).
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ __nv_bfloat162 as_bf162(unsigned u) {
  __nv_bfloat162 r;
  *reinterpret_cast<unsigned*>(&r) = u;
  return r;
}
__device__ __forceinline__ unsigned as_u32(__nv_bfloat162 v) { return *reinterpret_cast<unsigned*>(&v); }

// Forward: two channels per instruction (bf16x2 compare / max), the running argmax kept as two 16-bit lanes that are
// merged by one LOP3 per candidate; the result pairs are stored without any float conversion.
template <int KIND>
__global__ void __launch_bounds__(128)
pool_fwd_3x3s2_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, uint8_t* __restrict__ offs8,
                      const FastPoolGeom g) {
  vb::pdl_sync();
  const unsigned C8 = g.C >> 3;
  const unsigned total = static_cast<unsigned>(g.B) * g.OH * g.OW * C8;
  const unsigned NEG_INF2 = 0xff80ff80u;     // (-inf, -inf): never wins, stands for taps outside the image
  for (unsigned tt = blockIdx.x * blockDim.x + threadIdx.x; tt < total; tt += gridDim.x * blockDim.x) {
    const int c = static_cast<int>(tt % C8) << 3;
    unsigned r = tt / C8;
    const int ow = static_cast<int>(r % g.OW);
    r /= g.OW;
    const int oh = static_cast<int>(r % g.OH);
    const int b = static_cast<int>(r / g.OH);
    const int h0 = 2 * oh, w0 = 2 * ow;
    const __nv_bfloat16* base = x + ((static_cast<long long>(b) * g.H + h0) * g.W + w0) * g.C + c;
    uint4 v[9];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const bool ok = (h0 + i < g.H) && (w0 + j < g.W);
        v[i * 3 + j] = ok ? ldg16(base + (static_cast<long long>(i) * g.W + j) * g.C)
                          : make_uint4(NEG_INF2, NEG_INF2, NEG_INF2, NEG_INF2);
      }
    // tap 0 always lies inside the image: it seeds the running maximum (first maximum wins ties, like the oracle)
    unsigned best[4] = {v[0].x, v[0].y, v[0].z, v[0].w};
    unsigned idx[4] = {0u, 0u, 0u, 0u};
    __nv_bfloat162 key[4];
#pragma unroll
    for (int p2 = 0; p2 < 4; ++p2) key[p2] = KIND == 1 ? __habs2(as_bf162(best[p2])) : as_bf162(best[p2]);
#pragma unroll
    for (int l = 1; l < 9; ++l) {
      const unsigned cand[4] = {v[l].x, v[l].y, v[l].z, v[l].w};
#pragma unroll
      for (int p2 = 0; p2 < 4; ++p2) {
        const __nv_bfloat162 cv = as_bf162(cand[p2]);
        const __nv_bfloat162 ck = KIND == 1 ? __habs2(cv) : cv;
        // a tap outside the image is -inf (|-inf| = +inf would win the max-abs race: mask it out explicitly)
        unsigned m = __hgt2_mask(ck, key[p2]);
        if (KIND == 1) {
          const bool ok = (h0 + l / 3 < g.H) && (w0 + l % 3 < g.W);
          m = ok ? m : 0u;
        }
        key[p2] = as_bf162((as_u32(ck) & m) | (as_u32(key[p2]) & ~m));
        best[p2] = (cand[p2] & m) | (best[p2] & ~m);
        idx[p2] = ((static_cast<unsigned>(l) * 0x00010001u) & m) | (idx[p2] & ~m);
      }
    }
    const long long oo = ((static_cast<long long>(b) * g.OH + oh) * g.OW + ow) * g.C + c;
    *reinterpret_cast<uint4*>(y + oo) = make_uint4(best[0], best[1], best[2], best[3]);
    // idx[p2] = (index of channel 2 p2) | (index of channel 2 p2 + 1) << 16  ->  one byte per channel
    *reinterpret_cast<uint2*>(offs8 + oo) =
        make_uint2(__byte_perm(idx[0], idx[1], 0x6420), __byte_perm(idx[2], idx[3], 0x6420));
  }
}

__global__ void __launch_bounds__(128)
pool_bwd_3x3s2_kernel(const __nv_bfloat16* __restrict__ dy, const uint8_t* __restrict__ offs8,
                      __nv_bfloat16* __restrict__ dx, const FastPoolGeom g, const __nv_bfloat16* __restrict__ z,
                      int deriv) {
  vb::pdl_sync();
  const unsigned C8 = g.C >> 3;
  const int BH = (g.H + 1) >> 1, BW = (g.W + 1) >> 1;
  const unsigned total = static_cast<unsigned>(g.B) * BH * BW * C8;
  for (unsigned tt = blockIdx.x * blockDim.x + threadIdx.x; tt < total; tt += gridDim.x * blockDim.x) {
    const int c = static_cast<int>(tt % C8) << 3;
    unsigned r = tt / C8;
    const int bw = static_cast<int>(r % BW);
    r /= BW;
    const int bh = static_cast<int>(r % BH);
    const int b = static_cast<int>(r / BH);
    // windows (bh - 1 + a, bw - 1 + e), a, e in {0, 1}
    uint4 dv[4];
    uint2 lv[4];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int oh = bh - 1 + a, ow = bw - 1 + e;
        const bool ok = oh >= 0 && oh < g.OH && ow >= 0 && ow < g.OW;
        const long long o = ((static_cast<long long>(b) * g.OH + oh) * g.OW + ow) * g.C + c;
        dv[a * 2 + e] = ok ? ldg16(dy + o) : make_uint4(0, 0, 0, 0);
        // 0xff never equals a window-local index (0..8): an absent window contributes nothing
        lv[a * 2 + e] = ok ? __ldg(reinterpret_cast<const uint2*>(offs8 + o)) : make_uint2(0xffffffffu, 0xffffffffu);
      }
    float acc[4][8];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
      for (int q = 0; q < 8; ++q) acc[p][q] = 0.0f;
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float f[8];
        unpack8(dv[a * 2 + e], f);
        const uint2 l8 = lv[a * 2 + e];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const unsigned l = ((q < 4 ? l8.x : l8.y) >> (8 * (q & 3))) & 0xffu;
          // pixel (i, j) of the block sits at window-local row (a ? i : 2 [only i == 0]), column (e ? j : 2 [only j == 0])
#pragma unroll
          for (int i = 0; i < 2; ++i)
#pragma unroll
            for (int j = 0; j < 2; ++j) {
              if ((a == 0 && i == 1) || (e == 0 && j == 1)) continue;
              const unsigned want = static_cast<unsigned>((a ? i : 2) * 3 + (e ? j : 2));
              acc[i * 2 + j][q] += (l == want) ? f[q] : 0.0f;
            }
        }
      }
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int h = 2 * bh + i, w = 2 * bw + j;
        if (h >= g.H || w >= g.W) continue;
        const long long me = ((static_cast<long long>(b) * g.H + h) * g.W + w) * g.C + c;
        if (z != nullptr) {
          float f[8];
          unpack8(ldg16(z + me), f);
#pragma unroll
          for (int q = 0; q < 8; ++q) acc[i * 2 + j][q] *= act_deriv_from_output(deriv, f[q]);
        }
        *reinterpret_cast<uint4*>(dx + me) = pack8(acc[i * 2 + j]);
      }
  }
}

// ---------------------------------------------------------------------------------------------
// LRN: one warp per pixel, the pixel's channel vector staged in shared memory, so every element
// costs one pow instead of one per window member (backward was O(n^2) pows per element).
// smem per warp: 3*C floats (x, p = s^-beta, t = dy*x*p/s).
// ---------------------------------------------------------------------------------------------
__global__ void lrn_forward_row_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y,
                                       long long pixels, int C, int n, float alpha, float beta, float k) {
  vb::pdl_sync();
  extern __shared__ float sm[];
  const int warps = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* xs = sm + warp * C;
  const int half = n / 2;
  for (long long px = blockIdx.x * static_cast<long long>(warps) + warp; px < pixels;
       px += static_cast<long long>(gridDim.x) * warps) {
    const __nv_bfloat16* xr = x + px * C;
    for (int c = lane; c < C; c += 32) xs[c] = __bfloat162float(xr[c]);
    __syncwarp();
    for (int c = lane; c < C; c += 32) {
      float s = 0.0f;
      const int lo = max(0, c - half), hi = min(C - 1, c + half);
      for (int j = lo; j <= hi; ++j) s += xs[j] * xs[j];
      s = k + alpha * s;
      y[px * C + c] = __float2bfloat16(xs[c] * __powf(s, -beta));
    }
    __syncwarp();
  }
}

__global__ void lrn_backward_row_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                                        __nv_bfloat16* __restrict__ dx, long long pixels, int C, int n, float alpha,
                                        float beta, float k, int deriv) {
  vb::pdl_sync();
  extern __shared__ float sm[];
  const int warps = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* xs = sm + warp * 3 * C;
  float* ps = xs + C;
  float* ts = ps + C;
  const int half = n / 2;
  for (long long px = blockIdx.x * static_cast<long long>(warps) + warp; px < pixels;
       px += static_cast<long long>(gridDim.x) * warps) {
    const __nv_bfloat16* xr = x + px * C;
    const __nv_bfloat16* dr = dy + px * C;
    for (int c = lane; c < C; c += 32) xs[c] = __bfloat162float(xr[c]);
    __syncwarp();
    for (int c = lane; c < C; c += 32) {
      float s = 0.0f;
      const int lo = max(0, c - half), hi = min(C - 1, c + half);
      for (int j = lo; j <= hi; ++j) s += xs[j] * xs[j];
      s = k + alpha * s;
      const float p = __powf(s, -beta);
      const float d = __bfloat162float(dr[c]);
      ps[c] = d * p;                 // dy * s^-beta
      ts[c] = d * xs[c] * p / s;     // dy * x * s^(-beta-1)
    }
    __syncwarp();
    for (int c = lane; c < C; c += 32) {
      float acc = 0.0f;
      const int lo = max(0, c - half), hi = min(C - 1, c + half);
      for (int j = lo; j <= hi; ++j) acc += ts[j];
      float v = ps[c] - 2.0f * alpha * beta * xs[c] * acc;
      if (deriv) v *= act_deriv_from_output(deriv, xs[c]);
      dx[px * C + c] = __float2bfloat16(v);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------------------------
// LRN, vector form: one thread per (pixel, 8-channel group).  The thread loads its own 16-byte
// vector plus the two neighbouring ones (L1 hits: the neighbours are other threads' own vectors),
// so everything a +-HALF window needs (and, backward, the window of windows, +-2*HALF <= 8) sits
// in registers; no shared memory, no inter-thread traffic, 16-byte loads/stores throughout.
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void load24(const __nv_bfloat16* row, int g, int T, float (&v)[24]) {
  float a[8];
  if (g > 0) unpack8(ldg16(row + (g - 1) * 8), a);
  else {
#pragma unroll
    for (int q = 0; q < 8; ++q) a[q] = 0.0f;
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) v[q] = a[q];
  unpack8(ldg16(row + g * 8), a);
#pragma unroll
  for (int q = 0; q < 8; ++q) v[8 + q] = a[q];
  if (g + 1 < T) unpack8(ldg16(row + (g + 1) * 8), a);
  else {
#pragma unroll
    for (int q = 0; q < 8; ++q) a[q] = 0.0f;
  }
#pragma unroll
  for (int q = 0; q < 8; ++q) v[16 + q] = a[q];
}

// One MUFU.RSQ: s = k + alpha * sum >= k > 0 is a normal number, so the denormal fix-up code that rsqrtf() carries
// (a compare and two predicated multiplies per call) is dead weight in these instruction-bound kernels.
__device__ __forceinline__ float rsqrt_fast(float v) {
  float r;
  asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v));
  return r;
}
// s^-beta.  The AlexNet value beta = 0.75 needs no log/exp: r = rsqrt(s), s^-0.75 = r * sqrt(r), 1/s = r * r.
__device__ __forceinline__ float lrn_pow_neg(float s, float beta, bool beta34) {
  if (beta34) {
    const float r = rsqrt_fast(s);
    return r * r * rsqrt_fast(r);   // r^1.5 with two MUFU.RSQ (sqrtf would take the slow IEEE path)
  }
  return __powf(s, -beta);
}
__device__ __forceinline__ float lrn_pow_neg_inv(float s, float beta, bool beta34, float& inv_s) {
  if (beta34) {
    const float r = rsqrt_fast(s);
    inv_s = r * r;
    return inv_s * rsqrt_fast(r);
  }
  inv_s = __fdividef(1.0f, s);
  return __powf(s, -beta);
}

template <int HALF>
__global__ void __launch_bounds__(256)
lrn_forward_vec8_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long pixels, int C,
                        float alpha, float beta, float k) {
  vb::pdl_sync();
  const int T = C >> 3;
  const long long total = pixels * T;
  const bool beta34 = (beta == 0.75f);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    // 32-bit arithmetic when it fits (a 64-bit division is ~100 instructions)
    const long long px = (total <= 0xffffffffLL) ? static_cast<long long>(static_cast<unsigned>(i) / static_cast<unsigned>(T))
                                                  : i / T;
    const int g = static_cast<int>(i - px * T);
    float xv[24];
    load24(x + px * C, g, T, xv);
    float sq[24];
#pragma unroll
    for (int j = 0; j < 24; ++j) sq[j] = xv[j] * xv[j];
    float out[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      float s = 0.0f;
#pragma unroll
      for (int d = -HALF; d <= HALF; ++d) s += sq[8 + q + d];
      out[q] = xv[8 + q] * lrn_pow_neg(k + alpha * s, beta, beta34);
    }
    *reinterpret_cast<uint4*>(y + px * C + g * 8) = pack8(out);
  }
}

template <int HALF>
__global__ void __launch_bounds__(256)
lrn_backward_vec8_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                         __nv_bfloat16* __restrict__ dx, long long pixels, int C, float alpha, float beta, float k,
                         int deriv) {
  vb::pdl_sync();
  const int T = C >> 3;
  const long long total = pixels * T;
  const float coef = 2.0f * alpha * beta;
  const bool beta34 = (beta == 0.75f);
  for (long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    // 32-bit arithmetic when it fits (a 64-bit division is ~100 instructions)
    const long long px = (total <= 0xffffffffLL) ? static_cast<long long>(static_cast<unsigned>(i) / static_cast<unsigned>(T))
                                                  : i / T;
    const int g = static_cast<int>(i - px * T);
    float xv[24], dv[24];
    load24(x + px * C, g, T, xv);
    load24(dy + px * C, g, T, dv);
    float sq[24];
#pragma unroll
    for (int j = 0; j < 24; ++j) sq[j] = xv[j] * xv[j];
    // t[jj] = dy*x*s^(-beta-1) for channels c0-HALF .. c0+7+HALF;  dp[q] = dy*s^-beta for the own 8
    float t[8 + 2 * HALF], dp[8];
#pragma unroll
    for (int jj = 0; jj < 8 + 2 * HALF; ++jj) {
      const int j = 8 - HALF + jj;
      float s = 0.0f;
#pragma unroll
      for (int d = -HALF; d <= HALF; ++d) s += sq[j + d];
      s = k + alpha * s;
      float inv_s;
      const float pw = lrn_pow_neg_inv(s, beta, beta34, inv_s);
      const float dpw = dv[j] * pw;
      t[jj] = dpw * xv[j] * inv_s;
      if (jj >= HALF && jj < HALF + 8) dp[jj - HALF] = dpw;
    }
    float out[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      float acc = 0.0f;
#pragma unroll
      for (int d = 0; d <= 2 * HALF; ++d) acc += t[q + d];
      float v = dp[q] - coef * xv[8 + q] * acc;
      if (deriv) v *= act_deriv_from_output(deriv, xv[8 + q]);
      out[q] = v;
    }
    *reinterpret_cast<uint4*>(dx + px * C + g * 8) = pack8(out);
  }
}

// ---------------------------------------------------------------------------------------------
// LRN for C <= 256: the lanes of a warp hold consecutive 8-channel groups of whole pixels (32 / T pixels per warp,
// T = C / 8), so the +-HALF neighbour channels of a group are two registers of the adjacent lane: a thread loads its
// own 16 bytes only (the register-window kernels above load 48) and evaluates one power per own channel (backward:
// 8 instead of 8 + 2 HALF); 2 HALF shuffles replace the neighbour loads (backward: twice, for the squares and for
// the t terms).  Pixel borders are lanes with g == 0 / g == T - 1: their outer neighbours are zeroed.
// ---------------------------------------------------------------------------------------------
template <int HALF>
__device__ __forceinline__ void lrn_exchange(const float (&own)[8], float (&ext)[8 + 2 * HALF], bool first, bool last) {
#pragma unroll
  for (int q = 0; q < 8; ++q) ext[HALF + q] = own[q];
#pragma unroll
  for (int d = 0; d < HALF; ++d) {
    const float l = __shfl_up_sync(0xffffffffu, own[8 - HALF + d], 1);
    const float r = __shfl_down_sync(0xffffffffu, own[d], 1);
    ext[d] = first ? 0.0f : l;
    ext[HALF + 8 + d] = last ? 0.0f : r;
  }
}

// out[q] = sum of ext[q .. q + 2 HALF]: the window slides by one element per output (2 adds instead of 2 HALF)
template <int HALF>
__device__ __forceinline__ void lrn_window_sums(const float (&ext)[8 + 2 * HALF], float (&out)[8]) {
  float s = 0.0f;
#pragma unroll
  for (int d = 0; d <= 2 * HALF; ++d) s += ext[d];
  out[0] = s;
#pragma unroll
  for (int q = 1; q < 8; ++q) {
    s = s - ext[q - 1] + ext[q + 2 * HALF];
    out[q] = s;
  }
}

template <int HALF, bool B34>
__global__ void __launch_bounds__(256)
lrn_forward_shfl_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ y, long long pixels, int C,
                        float alpha, float beta, float k) {
  vb::pdl_sync();
  const int T = C >> 3;
  const int ppw = 32 / T;                       // pixels per warp
  const int lane = threadIdx.x & 31;
  const int pl = lane / T, g = lane - pl * T;
  const bool active_lane = pl < ppw;
  constexpr bool beta34 = B34;   // beta == 0.75 (checked by the launcher): two MUFU.RSQ, no log/exp
  const long long warps_total = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const long long groups = (pixels + ppw - 1) / ppw;
  for (long long w = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5; w < groups; w += warps_total) {
    const long long px = w * ppw + pl;
    const bool active = active_lane && px < pixels;
    float xv[8];
    if (active) unpack8(ldg16(x + px * C + g * 8), xv);
    else {
#pragma unroll
      for (int q = 0; q < 8; ++q) xv[q] = 0.0f;
    }
    float sq[8], ext[8 + 2 * HALF];
#pragma unroll
    for (int q = 0; q < 8; ++q) sq[q] = xv[q] * xv[q];
    lrn_exchange<HALF>(sq, ext, g == 0, g == T - 1);
    float out[8], ws[8];
    lrn_window_sums<HALF>(ext, ws);
#pragma unroll
    for (int q = 0; q < 8; ++q) out[q] = xv[q] * lrn_pow_neg(k + alpha * ws[q], beta, beta34);
    if (active) *reinterpret_cast<uint4*>(y + px * C + g * 8) = pack8(out);
  }
}

template <int HALF, bool B34, bool COLSUM>
__global__ void __launch_bounds__(256, 4)      // 4 blocks per SM with or without the 8 column accumulators
lrn_backward_shfl_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dy,
                         __nv_bfloat16* __restrict__ dx, long long pixels, int C, float alpha, float beta, float k,
                         int deriv, float* __restrict__ colsum) {
  vb::pdl_sync();
  // colsum (optional, [gridDim.x][C] fp32 partial rows, one per block): per-channel sums of dx - the bias gradient of
  // the convolution whose output this LRN normalises (dx IS that layer's err_output), so that layer does not read dx
  // again for it.  A lane keeps the same 8 channels for every pixel it visits: 8 adds per pixel in registers, one
  // shared-memory reduction per block, one plain row store per block; the second phase of the column-sum kernels
  // (vb_col_sum_finish) folds the rows.  (Global atomics from every block on the same C addresses cost more than the
  // column-sum pass they replaced: +36 us on AlexNet norm1.)
  // the running sums live in shared memory, 8 floats per thread touched with 16-byte accesses: as registers they pushed
  // the kernel from 4 to 3 resident blocks per SM (or into spills), +10-30 % on its run time
  __shared__ float bsum[256];
  __shared__ __align__(16) float bacc_s[COLSUM ? 256 * 8 : 8];
  float4* bacc = reinterpret_cast<float4*>(bacc_s + (COLSUM ? threadIdx.x * 8 : 0));
  if (COLSUM) {
    bacc[0] = make_float4(0.f, 0.f, 0.f, 0.f);
    bacc[1] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = threadIdx.x; i < C; i += blockDim.x) bsum[i] = 0.0f;
    __syncthreads();
  }
  const int T = C >> 3;
  const int ppw = 32 / T;
  const int lane = threadIdx.x & 31;
  const int pl = lane / T, g = lane - pl * T;
  const bool active_lane = pl < ppw;
  constexpr bool beta34 = B34;   // beta == 0.75 (checked by the launcher): two MUFU.RSQ, no log/exp
  const float coef = 2.0f * alpha * beta;
  const long long warps_total = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const long long groups = (pixels + ppw - 1) / ppw;
  for (long long w = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5; w < groups; w += warps_total) {
    const long long px = w * ppw + pl;
    const bool active = active_lane && px < pixels;
    float xv[8], dv[8];
    if (active) {
      const uint4 xr = ldg16(x + px * C + g * 8);
      const uint4 dr = ldg16(dy + px * C + g * 8);
      unpack8(xr, xv);
      unpack8(dr, dv);
    } else {
#pragma unroll
      for (int q = 0; q < 8; ++q) xv[q] = dv[q] = 0.0f;
    }
    float sq[8], ext[8 + 2 * HALF];
#pragma unroll
    for (int q = 0; q < 8; ++q) sq[q] = xv[q] * xv[q];
    lrn_exchange<HALF>(sq, ext, g == 0, g == T - 1);
    float t[8], dp[8], ws[8];
    lrn_window_sums<HALF>(ext, ws);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float s = k + alpha * ws[q];
      float inv_s;
      const float pw = lrn_pow_neg_inv(s, beta, beta34, inv_s);
      dp[q] = dv[q] * pw;
      t[q] = dp[q] * xv[q] * inv_s;
    }
    lrn_exchange<HALF>(t, ext, g == 0, g == T - 1);
    float out[8];
    lrn_window_sums<HALF>(ext, ws);
#pragma unroll
    for (int q = 0; q < 8; ++q) out[q] = dp[q] - coef * xv[q] * ws[q];
    if (deriv == ACT_STRICT_RELU) {          // warp-uniform: the common fused derivative costs one select per value
#pragma unroll
      for (int q = 0; q < 8; ++q) out[q] = xv[q] > 0.0f ? out[q] : 0.0f;
    } else if (deriv) {
#pragma unroll
      for (int q = 0; q < 8; ++q) out[q] *= act_deriv_from_output(deriv, xv[q]);
    }
    if (active) {
      *reinterpret_cast<uint4*>(dx + px * C + g * 8) = pack8(out);
      if (COLSUM) {
        float4 a0 = bacc[0], a1 = bacc[1];
        a0.x += out[0]; a0.y += out[1]; a0.z += out[2]; a0.w += out[3];
        a1.x += out[4]; a1.y += out[5]; a1.z += out[6]; a1.w += out[7];
        bacc[0] = a0;
        bacc[1] = a1;
      }
    }
  }
  if (COLSUM) {
    if (active_lane) {
      const float4 a0 = bacc[0], a1 = bacc[1];
      atomicAdd(&bsum[g * 8 + 0], a0.x); atomicAdd(&bsum[g * 8 + 1], a0.y);
      atomicAdd(&bsum[g * 8 + 2], a0.z); atomicAdd(&bsum[g * 8 + 3], a0.w);
      atomicAdd(&bsum[g * 8 + 4], a1.x); atomicAdd(&bsum[g * 8 + 5], a1.y);
      atomicAdd(&bsum[g * 8 + 6], a1.z); atomicAdd(&bsum[g * 8 + 7], a1.w);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < C; i += blockDim.x) colsum[static_cast<long long>(blockIdx.x) * C + i] = bsum[i];
  }
}

// ---------------------------------------------------------------------------------------------
// Max-pooling backward (3x3 / stride 2, byte argmax) FUSED into LRN backward: the thread that owns 8 channels of an
// LRN pixel gathers its dL/dy from the (up to four) pooling windows that cover the pixel - a window contributes where
// its argmax byte names this pixel - and goes straight on with the LRN backward math.  The dense dL/dy tensor of the
// LRN output (as large as the convolution output: 148 MB for AlexNet norm1) is neither written nor read.
// Same lane layout as lrn_backward_shfl_kernel (C <= 256).
// ---------------------------------------------------------------------------------------------
template <int HALF, bool B34>
__global__ void __launch_bounds__(256)
lrn_pool_backward_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ dpool,
                         const uint8_t* __restrict__ offs8, __nv_bfloat16* __restrict__ dx, const FastPoolGeom g,
                         float alpha, float beta, float k, int deriv) {
  vb::pdl_sync();
  const int C = g.C;
  const int T = C >> 3;
  const int ppw = 32 / T;
  const int lane = threadIdx.x & 31;
  const int pl = lane / T, gq = lane - pl * T;
  const bool active_lane = pl < ppw;
  constexpr bool beta34 = B34;
  const float coef = 2.0f * alpha * beta;
  const unsigned pixels = static_cast<unsigned>(g.B) * g.H * g.W;
  const long long warps_total = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const long long groups = (static_cast<long long>(pixels) + ppw - 1) / ppw;
  for (long long wv = (blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x) >> 5; wv < groups;
       wv += warps_total) {
    const unsigned px = static_cast<unsigned>(wv) * ppw + pl;
    const bool active = active_lane && px < pixels;
    float xv[8], dv[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) xv[q] = dv[q] = 0.0f;
    if (active) {
      const uint4 xr = ldg16(x + static_cast<long long>(px) * C + gq * 8);
      const unsigned w = px % g.W;
      const unsigned r = px / g.W;
      const unsigned h = r % g.H;
      const unsigned b = r / g.H;
      // windows (oh, ow) with 2 oh <= h <= 2 oh + 2: even h lies in two window rows, odd h in one
      const int oh_hi = static_cast<int>(h >> 1), ow_hi = static_cast<int>(w >> 1);
      const int oh_lo = (h & 1u) ? oh_hi : oh_hi - 1, ow_lo = (w & 1u) ? ow_hi : ow_hi - 1;
      uint4 dr[4];
      uint2 lr[4];
      unsigned want[4];
#pragma unroll
      for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int oh = a ? oh_hi : oh_lo, ow = e ? ow_hi : ow_lo;
          // (the "lo" window equals the "hi" one for odd coordinates: count it once)
          const bool ok = oh >= 0 && oh < g.OH && ow >= 0 && ow < g.OW && (a == 1 || oh_lo != oh_hi) &&
                          (e == 1 || ow_lo != ow_hi);
          const long long o = ((static_cast<long long>(b) * g.OH + (ok ? oh : 0)) * g.OW + (ok ? ow : 0)) * C + gq * 8;
          dr[a * 2 + e] = ok ? ldg16(dpool + o) : make_uint4(0, 0, 0, 0);
          lr[a * 2 + e] = ok ? __ldg(reinterpret_cast<const uint2*>(offs8 + o)) : make_uint2(0xffffffffu, 0xffffffffu);
          want[a * 2 + e] = (h - 2u * static_cast<unsigned>(ok ? oh : 0)) * 3u + (w - 2u * static_cast<unsigned>(ok ? ow : 0));
        }
      unpack8(xr, xv);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float f[8];
        unpack8(dr[i], f);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const unsigned l = ((q < 4 ? lr[i].x : lr[i].y) >> (8 * (q & 3))) & 0xffu;
          dv[q] += (l == want[i]) ? f[q] : 0.0f;
        }
      }
    }
    float sq[8], ext[8 + 2 * HALF];
#pragma unroll
    for (int q = 0; q < 8; ++q) sq[q] = xv[q] * xv[q];
    lrn_exchange<HALF>(sq, ext, gq == 0, gq == T - 1);
    float t[8], dp[8], ws[8];
    lrn_window_sums<HALF>(ext, ws);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float s = k + alpha * ws[q];
      float inv_s;
      const float pw = lrn_pow_neg_inv(s, beta, beta34, inv_s);
      dp[q] = dv[q] * pw;
      t[q] = dp[q] * xv[q] * inv_s;
    }
    lrn_exchange<HALF>(t, ext, gq == 0, gq == T - 1);
    float out[8];
    lrn_window_sums<HALF>(ext, ws);
#pragma unroll
    for (int q = 0; q < 8; ++q) out[q] = dp[q] - coef * xv[q] * ws[q];
    if (deriv == ACT_STRICT_RELU) {
#pragma unroll
      for (int q = 0; q < 8; ++q) out[q] = xv[q] > 0.0f ? out[q] : 0.0f;
    } else if (deriv) {
#pragma unroll
      for (int q = 0; q < 8; ++q) out[q] *= act_deriv_from_output(deriv, xv[q]);
    }
    if (active) *reinterpret_cast<uint4*>(dx + static_cast<long long>(px) * C + gq * 8) = pack8(out);
  }
}

// 1 = launched; 0 = geometry outside the fused kernel (3x3 / stride 2 max pooling, C % 8 == 0, C <= 256, window size
// n = 3..9, 16-byte aligned tensors)
int fast_lrn_pool_backward(const void* x, const void* dpool, const uint8_t* offs8, void* dx, const FastPoolGeom& g, int n,
                           float alpha, float beta, float k, int deriv, cudaStream_t s) {
  const int half = n / 2;
  if (g.ky != 3 || g.kx != 3 || g.sy != 2 || g.sx != 2 || g.C % 8 || g.C > 256 || half < 1 || half > 4) return 0;
  if ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(dpool) | reinterpret_cast<uintptr_t>(dx)) & 15) return 0;
  if (reinterpret_cast<uintptr_t>(offs8) & 7) return 0;
  const long long pixels = static_cast<long long>(g.B) * g.H * g.W;
  if (pixels == 0) return 1;
  if (pixels >= (1LL << 31)) return 0;
  const int ppw = 32 / (g.C >> 3);
  const int g2 = grid_cap2(((pixels + ppw - 1) / ppw) * 32, 256);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* dp = static_cast<const __nv_bfloat16*>(dpool);
  __nv_bfloat16* op = static_cast<__nv_bfloat16*>(dx);
#define VB_LRN_PB(H_)                                                                                                   \
  if (beta == 0.75f) vb::launch_k(lrn_pool_backward_kernel<H_, true>, dim3(g2), dim3(256), 0, s, xp, dp, offs8, op, g, alpha, beta, k, deriv);   \
  else vb::launch_k(lrn_pool_backward_kernel<H_, false>, dim3(g2), dim3(256), 0, s, xp, dp, offs8, op, g, alpha, beta, k, deriv)
  switch (half) {
    case 1: VB_LRN_PB(1); break;
    case 2: VB_LRN_PB(2); break;
    case 3: VB_LRN_PB(3); break;
    default: VB_LRN_PB(4); break;
  }
#undef VB_LRN_PB
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

// ---------------------------------------------------------------------------------------------
// Launchers (return 1 when the fast path was taken, 0 when the caller must use the generic one,
// negative on error)
// ---------------------------------------------------------------------------------------------
int fast_im2col(const void* x, void* cols, const FastConvGeom& g, cudaStream_t s) {
  if (g.ld % 8 || (reinterpret_cast<uintptr_t>(cols) & 15)) return 0;
  const long long total = static_cast<long long>(g.B) * g.OH * g.OW * (g.ld >> 3);
  if (total == 0) return 1;
  if (g.groups == 1) {
    const int rw = ((g.OW - 1) * g.sx + g.kx) * g.C;
    const size_t smem = static_cast<size_t>(g.ky) * rw * sizeof(unsigned short);
    if (smem <= 96 * 1024 && static_cast<long long>(g.B) * g.OH < 2147483647LL) {
      static bool attr_set = false;
      if (!attr_set) {
        VB_CHECK_CUDA(cudaFuncSetAttribute(im2col_rowtile_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        attr_set = true;
      }
      vb::launch_k(im2col_rowtile_kernel, dim3(g.B * g.OH), dim3(256), smem, s, static_cast<const __nv_bfloat16*>(x),
                                                          static_cast<__nv_bfloat16*>(cols), g, rw);
      VB_CHECK_CUDA(cudaGetLastError());
      return 1;
    }
  }
  vb::launch_k(im2col_any8_kernel, dim3(grid_cap2(total, 256)), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x),
                                                           static_cast<__nv_bfloat16*>(cols), g);
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

int fast_col2im(const void* dcols, void* dx, const FastConvGeom& g, const void* z, int deriv, cudaStream_t s) {
  const int Cg = g.C / g.groups;
  if (Cg % 8 || g.ld % 8 || (reinterpret_cast<uintptr_t>(dcols) & 15) || (reinterpret_cast<uintptr_t>(dx) & 15) ||
      (z && (reinterpret_cast<uintptr_t>(z) & 15)))
    return 0;
  const long long total = static_cast<long long>(g.B) * g.H * g.W * (g.C >> 3);
  if (total == 0) return 1;
  vb::launch_k(col2im_vec8_kernel, dim3(grid_cap2(total, 128)), dim3(128), 0, s, static_cast<const __nv_bfloat16*>(dcols),
                                                           static_cast<__nv_bfloat16*>(dx), g,
                                                           static_cast<const __nv_bfloat16*>(z), deriv);
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

int fast_pool_forward(int kind, const void* x, void* y, int* offs, uint8_t* offs8, const FastPoolGeom& g,
                      cudaStream_t s) {
  if (static_cast<long long>(g.B) * g.H * g.W * g.C >= 2147483647LL || g.kx * g.ky > 255) return 0;
  if (offs8 && (reinterpret_cast<uintptr_t>(offs8) & 7)) return 0;
  if (kind > 2 || g.C % 8 || (reinterpret_cast<uintptr_t>(x) & 15) || (reinterpret_cast<uintptr_t>(y) & 15) ||
      (offs && (reinterpret_cast<uintptr_t>(offs) & 15)))
    return 0;
  const long long total = static_cast<long long>(g.B) * g.OH * g.OW * (g.C >> 3);
  if (total == 0) return 1;
  if (offs == nullptr && offs8 != nullptr && kind != 2 && g.kx == 3 && g.ky == 3 && g.sx == 2 && g.sy == 2) {
    if (kind == 1)
      vb::launch_k(pool_fwd_3x3s2_kernel<1>, dim3(grid_cap2(total, 128)), dim3(128), 0, s, static_cast<const __nv_bfloat16*>(x),
                                                                    static_cast<__nv_bfloat16*>(y), offs8, g);
    else
      vb::launch_k(pool_fwd_3x3s2_kernel<0>, dim3(grid_cap2(total, 128)), dim3(128), 0, s, static_cast<const __nv_bfloat16*>(x),
                                                                    static_cast<__nv_bfloat16*>(y), offs8, g);
    VB_CHECK_CUDA(cudaGetLastError());
    return 1;
  }
  vb::launch_k(pool_forward_vec8_kernel, dim3(grid_cap2(total, 128)), dim3(128), 0, s, kind, static_cast<const __nv_bfloat16*>(x),
                                                                 static_cast<__nv_bfloat16*>(y), offs, offs8, g);
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

int fast_pool_backward(int kind, const void* dy, const int* offs, const uint8_t* offs8, void* dx,
                       const FastPoolGeom& g, const void* z, int deriv, cudaStream_t s) {
  if (static_cast<long long>(g.B) * g.H * g.W * g.C >= 2147483647LL) return 0;
  if (offs8 && (reinterpret_cast<uintptr_t>(offs8) & 7)) return 0;
  if (g.C % 8 || (reinterpret_cast<uintptr_t>(dy) & 15) || (reinterpret_cast<uintptr_t>(dx) & 15) ||
      (offs && (reinterpret_cast<uintptr_t>(offs) & 15)) || (z && (reinterpret_cast<uintptr_t>(z) & 15)))
    return 0;
  const long long total = static_cast<long long>(g.B) * g.H * g.W * (g.C >> 3);
  if (total == 0) return 1;
  if (offs8 != nullptr && kind != 2 && g.kx == 3 && g.ky == 3 && g.sx == 2 && g.sy == 2) {
    const long long blocks4 = static_cast<long long>(g.B) * ((g.H + 1) / 2) * ((g.W + 1) / 2) * (g.C >> 3);
    vb::launch_k(pool_bwd_3x3s2_kernel, dim3(grid_cap2(blocks4, 128)), dim3(128), 0, s, static_cast<const __nv_bfloat16*>(dy), offs8,
                                                                 static_cast<__nv_bfloat16*>(dx), g,
                                                                 static_cast<const __nv_bfloat16*>(z), deriv);
    VB_CHECK_CUDA(cudaGetLastError());
    return 1;
  }
  vb::launch_k(pool_backward_vec8_kernel, dim3(grid_cap2(total, 128)), dim3(128), 0, s, kind, static_cast<const __nv_bfloat16*>(dy), offs, offs8,
                                                                  static_cast<__nv_bfloat16*>(dx), g,
                                                                  static_cast<const __nv_bfloat16*>(z), deriv);
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

int fast_pool_expand_offsets(const uint8_t* offs8, int* offs, const FastPoolGeom& g, cudaStream_t s) {
  const long long total = static_cast<long long>(g.B) * g.OH * g.OW * g.C;
  if (total == 0) return 1;
  if (static_cast<long long>(g.B) * g.H * g.W * g.C >= 2147483647LL) return 0;
  vb::launch_k(pool_expand_offsets_kernel, dim3(grid_cap2(total, 256)), dim3(256), 0, s, offs8, offs, g);
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

int fast_lrn_forward(const void* x, void* y, long long pixels, int C, int n, float alpha, float beta, float k,
                     cudaStream_t s) {
  const int half = n / 2;
  if (C % 8 == 0 && half >= 1 && half <= 4 && !(reinterpret_cast<uintptr_t>(x) & 15) &&
      !(reinterpret_cast<uintptr_t>(y) & 15)) {
    const long long total = pixels * (C >> 3);
    if (total == 0) return 1;
    const int grid = grid_cap2(total, 256);
    const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
    __nv_bfloat16* yp = static_cast<__nv_bfloat16*>(y);
    if (C <= 256) {      // whole pixels per warp: neighbours by shuffle
      const int ppw = 32 / (C >> 3);
      const int g2 = grid_cap2(((pixels + ppw - 1) / ppw) * 32, 256);
#define VB_LRN_F(H_)                                                                                              \
  if (beta == 0.75f) vb::launch_k(lrn_forward_shfl_kernel<H_, true>, dim3(g2), dim3(256), 0, s, xp, yp, pixels, C, alpha, beta, k);        \
  else vb::launch_k(lrn_forward_shfl_kernel<H_, false>, dim3(g2), dim3(256), 0, s, xp, yp, pixels, C, alpha, beta, k)
      switch (half) {
        case 1: VB_LRN_F(1); break;
        case 2: VB_LRN_F(2); break;
        case 3: VB_LRN_F(3); break;
        default: VB_LRN_F(4); break;
      }
#undef VB_LRN_F
      VB_CHECK_CUDA(cudaGetLastError());
      return 1;
    }
    switch (half) {
      case 1: vb::launch_k(lrn_forward_vec8_kernel<1>, dim3(grid), dim3(256), 0, s, xp, yp, pixels, C, alpha, beta, k); break;
      case 2: vb::launch_k(lrn_forward_vec8_kernel<2>, dim3(grid), dim3(256), 0, s, xp, yp, pixels, C, alpha, beta, k); break;
      case 3: vb::launch_k(lrn_forward_vec8_kernel<3>, dim3(grid), dim3(256), 0, s, xp, yp, pixels, C, alpha, beta, k); break;
      default: vb::launch_k(lrn_forward_vec8_kernel<4>, dim3(grid), dim3(256), 0, s, xp, yp, pixels, C, alpha, beta, k); break;
    }
    VB_CHECK_CUDA(cudaGetLastError());
    return 1;
  }
  if (C > 1024) return 0;
  const int warps = 8;
  const size_t smem = static_cast<size_t>(warps) * C * sizeof(float);
  long long blocks = (pixels + warps - 1) / warps;
  const long long cap = static_cast<long long>(device_sm_count()) * 8;
  if (blocks > cap) blocks = cap;
  vb::launch_k(lrn_forward_row_kernel, dim3(static_cast<int>(blocks)), dim3(warps * 32), smem, s, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(y), pixels, C, n, alpha, beta, k);
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

int fast_lrn_backward(const void* x, const void* dy, void* dx, long long pixels, int C, int n, float alpha, float beta,
                      float k, int deriv, cudaStream_t s, float* colsum, int max_blocks, int* blocks_out) {
  const int half = n / 2;
  if (colsum != nullptr && !(C % 8 == 0 && C <= 256 && half >= 1 && half <= 4)) return 0;   // only the lane-per-8-channels form sums
  if (C % 8 == 0 && half >= 1 && half <= 4 && !(reinterpret_cast<uintptr_t>(x) & 15) &&
      !(reinterpret_cast<uintptr_t>(dy) & 15) && !(reinterpret_cast<uintptr_t>(dx) & 15)) {
    const long long total = pixels * (C >> 3);
    if (total == 0) return 1;
    const int grid = grid_cap2(total, 256);
    const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
    const __nv_bfloat16* dp = static_cast<const __nv_bfloat16*>(dy);
    __nv_bfloat16* op = static_cast<__nv_bfloat16*>(dx);
    if (C <= 256) {
      const int ppw = 32 / (C >> 3);
      int g2 = grid_cap2(((pixels + ppw - 1) / ppw) * 32, 256);
      if (colsum != nullptr) {                       // one partial row per block
        if (g2 > device_sm_count() * 4) g2 = device_sm_count() * 4;        // the four resident blocks per SM: one wave
        if (g2 > max_blocks) g2 = max_blocks;
        if (g2 < 1) return 0;
        if (blocks_out) *blocks_out = g2;
      }
#define VB_LRN_B2(H_, B_)                                                                                                \
  if (colsum != nullptr) vb::launch_k(lrn_backward_shfl_kernel<H_, B_, true>, dim3(g2), dim3(256), 0, s, xp, dp, op, pixels, C, alpha, beta, k, deriv, colsum);   \
  else vb::launch_k(lrn_backward_shfl_kernel<H_, B_, false>, dim3(g2), dim3(256), 0, s, xp, dp, op, pixels, C, alpha, beta, k, deriv, colsum)
#define VB_LRN_B(H_)                                                                                                     \
  if (beta == 0.75f) { VB_LRN_B2(H_, true); } else { VB_LRN_B2(H_, false); }
      switch (half) {
        case 1: VB_LRN_B(1); break;
        case 2: VB_LRN_B(2); break;
        case 3: VB_LRN_B(3); break;
        default: VB_LRN_B(4); break;
      }
#undef VB_LRN_B
#undef VB_LRN_B2
      VB_CHECK_CUDA(cudaGetLastError());
      return 1;
    }
    switch (half) {
      case 1: vb::launch_k(lrn_backward_vec8_kernel<1>, dim3(grid), dim3(256), 0, s, xp, dp, op, pixels, C, alpha, beta, k, deriv); break;
      case 2: vb::launch_k(lrn_backward_vec8_kernel<2>, dim3(grid), dim3(256), 0, s, xp, dp, op, pixels, C, alpha, beta, k, deriv); break;
      case 3: vb::launch_k(lrn_backward_vec8_kernel<3>, dim3(grid), dim3(256), 0, s, xp, dp, op, pixels, C, alpha, beta, k, deriv); break;
      default: vb::launch_k(lrn_backward_vec8_kernel<4>, dim3(grid), dim3(256), 0, s, xp, dp, op, pixels, C, alpha, beta, k, deriv); break;
    }
    VB_CHECK_CUDA(cudaGetLastError());
    return 1;
  }
  if (C > 1024) return 0;
  const int warps = 4;
  const size_t smem = static_cast<size_t>(warps) * 3 * C * sizeof(float);
  long long blocks = (pixels + warps - 1) / warps;
  const long long cap = static_cast<long long>(device_sm_count()) * 12;
  if (blocks > cap) blocks = cap;
  vb::launch_k(lrn_backward_row_kernel, dim3(static_cast<int>(blocks)), dim3(warps * 32), smem, s, static_cast<const __nv_bfloat16*>(x), static_cast<const __nv_bfloat16*>(dy), static_cast<__nv_bfloat16*>(dx),
      pixels, C, n, alpha, beta, k, deriv);
  VB_CHECK_CUDA(cudaGetLastError());
  return 1;
}

// ---------------------------------------------------------------------------------------------
// Space-to-depth for strided few-channel convolutions (the 11x11 stride-4 RGB first layer):
//   xs[b, h', w', (dy*sx + dx)*C + c] = x[b, h'*sy + dy, w'*sx + dx, c]   (zero outside the image)
// turns the stride-s convolution into a unit-stride one with ceil(k/s)^2 taps over sy*sx*C
// channels, which the TMA im2col-mode implicit GEMM can walk (channel vectors of >= 16 bytes).
// One thread = 8 output channels (one 16-byte store).
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void space_to_depth_kernel(const T* __restrict__ x, __nv_bfloat16* __restrict__ xs, int B, int H, int W,
                                      int C, int sy, int sx, int H2, int W2) {
  vb::pdl_sync();
  const int C2 = sy * sx * C;
  const int V = C2 >> 3;
  const long long total = static_cast<long long>(B) * H2 * W2 * V;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int v = static_cast<int>(t % V);
    long long r = t / V;
    const int w2 = static_cast<int>(r % W2);
    r /= W2;
    const int h2 = static_cast<int>(r % H2);
    const int b = static_cast<int>(r / H2);
    int q = v << 3;
    int blk = q / C;
    int c = q - blk * C;
    int dy = blk / sx, dx = blk - dy * sx;
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int h = h2 * sy + dy, w = w2 * sx + dx;
      float val = 0.0f;
      if (h < H && w < W) val = static_cast<float>(x[((static_cast<long long>(b) * H + h) * W + w) * C + c]);
      f[e] = val;
      if (++c == C) {
        c = 0;
        if (++dx == sx) {
          dx = 0;
          ++dy;
        }
      }
    }
    *reinterpret_cast<uint4*>(xs + t * 8) = pack8(f);
  }
}

// Row-segment form of the same transform for bf16 inputs with (sx * C) % 4 == 0: the sx * C elements that one output
// pixel takes from input row 4 h' + dy are CONTIGUOUS in the input, so a thread copies one such segment (scalar 2-byte
// loads - an NHWC row of 3-channel pixels is only 2-byte aligned - packed into 8-byte stores) with no per-element
// index arithmetic.  The AlexNet first layer (sx * C = 12) moves 24 bytes per thread.
__global__ void __launch_bounds__(256)
space_to_depth_rows_kernel(const unsigned short* __restrict__ x, unsigned short* __restrict__ xs, int B, int H, int W,
                           int C, int sy, int sx, int H2, int W2) {
  vb::pdl_sync();
  const int n = sx * C;                 // segment length (elements), multiple of 4
  const int C2 = sy * n;
  const long long total = static_cast<long long>(B) * H2 * W2 * sy;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const unsigned tt = static_cast<unsigned>(t);       // < 2^32 (checked by the launcher)
    const int dy = static_cast<int>(tt % sy);
    unsigned r = tt / sy;
    const int w2 = static_cast<int>(r % W2);
    r /= W2;
    const int h2 = static_cast<int>(r % H2);
    const int b = static_cast<int>(r / H2);
    const int h = h2 * sy + dy;
    const int w0 = w2 * sx;
    int avail = 0;                       // elements of the segment that lie inside the image
    if (h < H) avail = min(n, (W - w0) * C);
    const unsigned short* src = x + ((static_cast<long long>(b) * H + h) * W + w0) * C;
    unsigned short* dst = xs + (static_cast<long long>(tt / sy) * C2 + dy * n);
    for (int k = 0; k < n; k += 4) {
      unsigned short e[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) e[j] = (k + j < avail) ? __ldg(src + k + j) : static_cast<unsigned short>(0);
      *reinterpret_cast<uint2*>(dst + k) = make_uint2(e[0] | (static_cast<unsigned>(e[1]) << 16),
                                                      e[2] | (static_cast<unsigned>(e[3]) << 16));
    }
  }
}

// dst[r, j] (+)= src[r, map[j]] (0 where map[j] < 0); optional zeroing of the consumed source row
// range afterwards is left to the caller.  Used to move weights / weight gradients between the
// Znicz column order and the space-to-depth column order (tiny matrices).
template <typename TS, typename TD>
__global__ void gather_cols_kernel(const TS* __restrict__ src, long long ld_src, TD* __restrict__ dst,
                                   long long ld_dst, const int* __restrict__ map, int rows, int cols, int accumulate) {
  vb::pdl_sync();
  const long long total = static_cast<long long>(rows) * cols;
  for (long long t = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; t < total;
       t += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int j = static_cast<int>(t % cols);
    const long long r = t / cols;
    const int m = map[j];
    float v = m >= 0 ? static_cast<float>(src[r * ld_src + m]) : 0.0f;
    if (accumulate) v += static_cast<float>(dst[r * ld_dst + j]);
    dst[r * ld_dst + j] = static_cast<TD>(v);
  }
}

}  // namespace vb

using namespace vb;

// x: [B,H,W,C] (dtype 0 = f32, 1 = bf16) -> xs: bf16 [B, ceil(H/sy), ceil(W/sx), sy*sx*C]; sy*sx*C % 8 == 0.
extern "C" int vb_space_to_depth(const void* x, int dtype, void* xs, int B, int H, int W, int C, int sy, int sx,
                                 cudaStream_t s) {
  VB_REQUIRE(sy >= 1 && sx >= 1 && C >= 1 && (sy * sx * C) % 8 == 0, "vb_space_to_depth: sy*sx*C must be a multiple of 8");
  VB_REQUIRE((reinterpret_cast<uintptr_t>(xs) & 15) == 0, "vb_space_to_depth: output must be 16-byte aligned");
  const int H2 = (H + sy - 1) / sy, W2 = (W + sx - 1) / sx;
  const long long total = static_cast<long long>(B) * H2 * W2 * ((sy * sx * C) >> 3);
  if (total == 0) return 0;
  const long long segs = static_cast<long long>(B) * H2 * W2 * sy;
  if (dtype == 1 && (sx * C) % 4 == 0 && segs < 0xffffffffLL) {
    vb::launch_k(space_to_depth_rows_kernel, dim3(grid_cap2(segs, 256)), dim3(256), 0, s, static_cast<const unsigned short*>(x),
                                                                   static_cast<unsigned short*>(xs), B, H, W, C, sy, sx,
                                                                   H2, W2);
    VB_CHECK_CUDA(cudaGetLastError());
    return 0;
  }
  const int grid = grid_cap2(total, 256);
  if (dtype == 0)
    vb::launch_k(space_to_depth_kernel<float>, dim3(grid), dim3(256), 0, s, static_cast<const float*>(x), static_cast<__nv_bfloat16*>(xs), B, H, W, C, sy, sx, H2, W2);
  else if (dtype == 1)
    vb::launch_k(space_to_depth_kernel<__nv_bfloat16>, dim3(grid), dim3(256), 0, s, static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(xs), B, H, W, C, sy, sx, H2, W2);
  else
    VB_REQUIRE(false, "vb_space_to_depth: dtype must be f32 (0) or bf16 (1)");
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// dst[r, j] (+)= src[r, map[j]]; dtypes 0 = f32, 1 = bf16 (any pair).
extern "C" int vb_gather_cols(const void* src, int src_dtype, long long ld_src, void* dst, int dst_dtype,
                              long long ld_dst, const int* map, int rows, int cols, int accumulate, cudaStream_t s) {
  const long long total = static_cast<long long>(rows) * cols;
  if (total == 0) return 0;
  const int grid = grid_cap2(total, 256);
#define VB_GC(TS, TD) vb::launch_k(gather_cols_kernel<TS, TD>, dim3(grid), dim3(256), 0, s, static_cast<const TS*>(src), ld_src, static_cast<TD*>(dst), ld_dst, map, rows, cols, accumulate)
  if (src_dtype == 0 && dst_dtype == 0) VB_GC(float, float);
  else if (src_dtype == 0 && dst_dtype == 1) VB_GC(float, __nv_bfloat16);
  else if (src_dtype == 1 && dst_dtype == 1) VB_GC(__nv_bfloat16, __nv_bfloat16);
  else if (src_dtype == 1 && dst_dtype == 0) VB_GC(__nv_bfloat16, float);
  else VB_REQUIRE(false, "vb_gather_cols: unsupported dtype pair %d -> %d", src_dtype, dst_dtype);
#undef VB_GC
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
