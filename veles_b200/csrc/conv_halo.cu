// veles_b200 — "halo tile" implicit-GEMM convolution for unit-stride layers with many taps.
//
// The im2col-mode kernel (gemm_tcgen05.cu) re-fetches the activation operand once per filter tap:
// for a 5x5 layer every input element enters an SM 25 times, and with few output channels per
// tile (AlexNet conv2: 48 / 128 per group) that operand ingest, not the tensor core, bounds the
// kernel.  Here a CTA loads, per 64-channel block, ONE spatial tile of the input with its halo
//     (R + ky - 1) image rows  x  P pixels  x  64 channels        (P = 16/32/64 >= OW + kx - 1)
// through a 4-D tiled TMA box (negative / out-of-range coordinates zero-fill = the padding), and
// issues the MMAs of all ky*kx taps from that same shared-memory tile: tap (r, s) is simply the
// 128-row window that starts (r*P + s) rows further down, described by a shifted UMMA descriptor.
// The M rows of the MMA are R image rows of P lanes each (lane w' < OW is a real output pixel,
// the rest is discarded), so a 32-lane epilogue warp covers whole image-row segments and writes
// them with rectangular 4-D TMA store boxes that the tensor map clips to the image.
// Only the (small) weight tiles stream per tap.  Modes: forward, and err_input (the same walk
// over dY with flipped taps).  Replaces Znicz conv / gd_conv err_input (absent).
#include "common.cuh"

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <type_traits>

namespace vb {

constexpr int HALO_THREADS = 320;   // 8 epilogue warps, producer warp, MMA warp (roles in the highest warps: see
                                    // gemm_tcgen05.cu)
constexpr int PRODUCER_WARP = 8;
constexpr int MMA_WARP = 9;
constexpr int HALO_A_SLOT = 32768;  // <= 256 rows x 128 B
constexpr int HALO_EPI_BYTES = 8 * 4096;

struct HaloParams {
  int B, OH, OW;         // output grid (err_input mode: the layer INPUT grid)
  int P, R;              // lanes per image row, image rows per 128-row sub-tile (R = 128 / P)
  int kx, ky;            // taps
  int low_w, low_h;      // first input coordinate of output 0 = -(padding)   (err_input: -(k-1-padding))
  int flip;              // 1: weight tap of window offset (r, s) is (ky-1-r, kx-1-s)    (err_input)
  int cblocks, cg;       // 64-channel blocks of the walked tensor; weight-matrix columns per tap
  int ksteps_last;       // 16-channel MMA steps that hold data in the LAST channel block (1..4): padding is not multiplied
  int N, n_blocks;       // output channels, N tiles
  int tiles_per_img;     // ceil(OH / (R * MT))
  int a_rows;            // (R + ky - 1) * P rows of 128 B per activation tile
  int resident;          // 1: the layer's taps*cblocks weight tiles are loaded ONCE and stay in shared memory
  const float* bias;
  int act;
  const __nv_bfloat16* Z;  // optional [pixels, N] (pixel pitch ldz): D *= f'(Z)
  long long ldz;
  int deriv;
  long long* trace;        // profiling (vb_conv_halo_set_trace): CTA 0 records clock64 stamps of its first tiles here
};

// Trace layout: TRACE_TILES tiles x 8 stamps: [0] MMA thread starts waiting for the accumulator, [1] has it,
// [2] activation tile present, [3] all MMAs of the tile issued, [4] epilogue warp 0 starts waiting for the tile,
// [5] tile complete in tensor memory, [6] epilogue warp 0 done, [7] producer issued the tile's activation load.
constexpr int TRACE_TILES = 24;

template <int ACT>
__device__ __forceinline__ void halo_apply_act(float (&f)[32]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = act_forward_fast<ACT>(f[j]);
}
template <int ACT>
__device__ __forceinline__ void halo_apply_deriv(float (&f)[32], const float (&y)[32]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] *= act_deriv_from_output(ACT, y[j]);
}

// MT = 128-row sub-tiles per CTA tile (consecutive groups of R image rows): every weight tile that streams in is
// used by MT accumulators, so the weight bytes per output halve with MT = 2.
template <int BN, int MT>
struct HaloSmem {
  static constexpr int B_BYTES = BN * 128;
  static constexpr int A_SLOTS = 2 * MT;   // double-buffered over channel blocks / tiles
  static constexpr int BUDGET = 227 * 1024 - A_SLOTS * HALO_A_SLOT - HALO_EPI_BYTES - 2048;
  static constexpr int B_STAGES = (BUDGET / B_BYTES) > 16 ? 16 : (BUDGET / B_BYTES);
  static constexpr int TOTAL = A_SLOTS * HALO_A_SLOT + B_STAGES * B_BYTES + HALO_EPI_BYTES + 2048;
};

// B_MN = false: weights K-major [N, taps*cg] (forward).  B_MN = true: weights [K = kernels, taps*cg] read as
// MN-major [k, n = input channel] tiles (err_input).
template <int BN, bool B_MN, int MT>
__global__ void __launch_bounds__(HALO_THREADS, 1)
conv_halo_kernel(const __grid_constant__ CUtensorMap tma_x, const __grid_constant__ CUtensorMap tma_w,
                 const __grid_constant__ CUtensorMap tma_d, const HaloParams p) {
  using S = HaloSmem<BN, MT>;
  constexpr int B_STAGES = S::B_STAGES;
  constexpr int A_SLOTS = S::A_SLOTS;
  constexpr int ACC_STRIDE = (BN <= 64) ? 64 : (BN <= 128 ? 128 : 256);   // TMEM columns per accumulator
  // Accumulation chains (see gemm_tcgen05.cu): a dependent tcgen05.mma waits ~170 cycles for its predecessor, so the
  // MMAs of a sub-tile rotate over KS accumulators which the epilogue adds, and consecutive MMAs alternate between
  // sub-tiles: MT * KS independent chains, MT * KS * ACC_STRIDE = 256 columns in flight.
  constexpr int KS = 256 / (MT * ACC_STRIDE) < 1 ? 1 : 256 / (MT * ACC_STRIDE);
  constexpr int TMEM_NEED = 2 * MT * KS * ACC_STRIDE;
  constexpr int TMEM_COLS = TMEM_NEED <= 64 ? 64 : (TMEM_NEED <= 128 ? 128 : (TMEM_NEED <= 256 ? 256 : 512));
  static_assert(TMEM_NEED <= 512, "accumulators exceed tensor memory");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* a_smem = smem;                                   // activation slots
  uint8_t* b_smem = smem + A_SLOTS * HALO_A_SLOT;           // weight ring (or the resident weights)
  uint8_t* epi_smem = b_smem + B_STAGES * S::B_BYTES;       // 8 x 4 KB staging slabs (1024-aligned)
  uint64_t* a_full = reinterpret_cast<uint64_t*>(epi_smem + HALO_EPI_BYTES);
  uint64_t* a_empty = a_full + A_SLOTS;
  uint64_t* b_full = a_empty + A_SLOTS;
  uint64_t* b_empty = b_full + B_STAGES;
  uint64_t* tmem_full = b_empty + B_STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == PRODUCER_WARP && elect_one()) {
    tma_prefetch_desc(&tma_x);
    tma_prefetch_desc(&tma_w);
    tma_prefetch_desc(&tma_d);
    for (int s = 0; s < A_SLOTS; ++s) {
      mbar_init(&a_full[s], 1);
      mbar_init(&a_empty[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], 8);
    }
    for (int s = 0; s < B_STAGES; ++s) {
      mbar_init(&b_full[s], 1);
      mbar_init(&b_empty[s], 1);
    }
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == MMA_WARP) tmem_alloc<TMEM_COLS>(tmem_ptr);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  pdl_sync();   // the set-up above overlaps the previous kernel's tail; global memory is touched from here on

  const int taps = p.kx * p.ky;
  const int num_tiles = p.B * p.tiles_per_img * p.n_blocks;

  // The two single-thread roles are bound by their own instruction streams (ncu source view): no divisions and no
  // descriptor construction inside the tap loops - tap coordinates and descriptors advance by increments.
  if (warp == PRODUCER_WARP) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int b_stage = 0, a_par = 0;
      uint32_t b_phase = 0, a_phase_bits = 0;   // one phase bit per activation slot (sub-tiles can be skipped)
      const int wtap_first = p.flip ? taps - 1 : 0;   // err_input walks the filter taps backwards
      const int wtap_step = p.flip ? -1 : 1;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const bool first = (tile == static_cast<int>(blockIdx.x));
        const int nb = tile % p.n_blocks;
        const int sp = tile / p.n_blocks;
        const int img = sp / p.tiles_per_img;
        const int oh0 = (sp - img * p.tiles_per_img) * (p.R * MT);
        for (int cb = 0; cb < p.cblocks; ++cb) {
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            if (oh0 + mt * p.R >= p.OH) continue;   // this sub-tile lies below the image
            const int slot = a_par * MT + mt;
            mbar_wait(&a_empty[slot], ((a_phase_bits >> slot) & 1u) ^ 1u);
            a_phase_bits ^= 1u << slot;
            mbar_expect_tx(&a_full[slot], static_cast<uint32_t>(p.a_rows) * 128u);
            tma_load_4d(a_smem + slot * HALO_A_SLOT, &tma_x, &a_full[slot], cb * 64, p.low_w,
                        oh0 + mt * p.R + p.low_h, img);
          }
          a_par ^= 1;
          if (p.trace != nullptr && blockIdx.x == 0 && cb == 0) {
            const int ti = (tile - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x);
            if (ti < TRACE_TILES) p.trace[ti * 8 + 7] = clock64();
          }
          if (p.resident && !first) continue;   // the weights are already in shared memory
          // weight column of window offset (r, s): tap (r, s), or the mirrored tap (ky-1-r, kx-1-s) = taps-1-tap
          int wcol = wtap_first * p.cg + (B_MN ? nb * BN : cb * 64);
          const int wcol_step = wtap_step * p.cg;
          for (int tap = 0; tap < taps; ++tap, wcol += wcol_step) {
            if (!p.resident) mbar_wait(&b_empty[b_stage], b_phase ^ 1);
            uint8_t* sb = b_smem + b_stage * S::B_BYTES;
            mbar_expect_tx(&b_full[b_stage], S::B_BYTES);
            if (!B_MN) {
              tma_load_2d(sb, &tma_w, &b_full[b_stage], wcol, nb * BN);
            } else {
#pragma unroll
              for (int c = 0; c < BN / 64; ++c)
                tma_load_2d(sb + c * 8192, &tma_w, &b_full[b_stage], wcol + c * 64, cb * 64);
            }
            if (++b_stage == B_STAGES) {
              b_stage = 0;
              b_phase ^= 1;
            }
          }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ===================== MMA issuer: ONE thread runs the whole loop =====================
    if (elect_one()) {
      int b_stage = 0, acc = 0, a_par = 0;
      uint32_t b_phase = 0, acc_phase = 0, a_phase_bits = 0;
      // Descriptors are carried as low words (the 14-bit shared-memory address field never carries out: addresses
      // stay below 2^18); the high words are constants.
      constexpr uint32_t B_STEP = static_cast<uint32_t>(S::B_BYTES >> 4);
      constexpr uint32_t A_SLOT_STEP = static_cast<uint32_t>(HALO_A_SLOT >> 4);
      constexpr uint32_t B_KSTEP = static_cast<uint32_t>((B_MN ? 2048 : 32) >> 4);
      const uint64_t b_desc_s0 = B_MN ? make_smem_desc_sw128(smem_u32(b_smem), 8192, 1024)
                                      : make_smem_desc_sw128(smem_u32(b_smem), 16, 1024);
      // activation windows: plain SWIZZLE_128B descriptors (base-offset field 0) - the swizzle is a function of the
      // absolute shared-memory address, so a window shifted by whole 128-byte rows reads exactly what TMA wrote
      // (measured: setting the base-offset field to the row phase gives wrong products)
      const uint64_t a_desc_s0 = make_smem_desc_sw128(smem_u32(a_smem), 16, 1024);
      const uint32_t a_hi = static_cast<uint32_t>(a_desc_s0 >> 32), b_hi = static_cast<uint32_t>(b_desc_s0 >> 32);
      const uint32_t a_lo_s0 = static_cast<uint32_t>(a_desc_s0), b_lo_s0 = static_cast<uint32_t>(b_desc_s0);
      const uint32_t line_step = static_cast<uint32_t>((p.P - p.kx) * 8);   // from the last tap of a filter row to the
                                                                            // first tap of the next one (16-byte units)
      const int ky = p.ky, kx = p.kx;
      uint32_t b_lo = b_lo_s0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const bool first = (tile == static_cast<int>(blockIdx.x));
        const int nb = tile % p.n_blocks;
        const int sp = tile / p.n_blocks;
        const int img = sp / p.tiles_per_img;
        const int oh0 = (sp - img * p.tiles_per_img) * (p.R * MT);
        bool mt_ok[MT];
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) mt_ok[mt] = oh0 + mt * p.R < p.OH;
        // only the columns that exist are multiplied (N = 96 of a 128-wide tile: a quarter of the MMA time saved)
        int n_mma = p.N - nb * BN;
        n_mma = (B_MN || n_mma >= BN) ? BN : ((n_mma + 15) & ~15);
        const uint32_t idesc = make_idesc(/*BF16*/ 1, false, B_MN, 128, n_mma);
        const int ti = (tile - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x);
        const bool tr = p.trace != nullptr && blockIdx.x == 0 && ti < TRACE_TILES;
        if (tr) p.trace[ti * 8 + 0] = clock64();
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        if (tr) p.trace[ti * 8 + 1] = clock64();
        const bool stream = !(p.resident && !first);   // weight tiles arrive through the ring (else: already resident)
        if (p.resident) {
          b_stage = 0;
          b_lo = b_lo_s0;
        }
        const uint32_t d_tile = tmem_base + static_cast<uint32_t>(acc * MT * KS * ACC_STRIDE);
        for (int cb = 0; cb < p.cblocks; ++cb) {
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            if (!mt_ok[mt]) continue;
            const int slot = a_par * MT + mt;
            mbar_wait(&a_full[slot], (a_phase_bits >> slot) & 1u);
            a_phase_bits ^= 1u << slot;
          }
          tc_fence_after();
          if (tr && cb == 0) p.trace[ti * 8 + 2] = clock64();
          // All taps of one channel block.  KSTEPS (16-channel MMA steps per tap) and STREAM are compile-time in the
          // body: the loop is the pacing instruction stream of the kernel (one thread, ~30 cycles per MMA plus
          // ~350 per iteration measured with the generic form), so it carries no run-time branches.
          auto run_cb = [&](auto ksteps_c, auto stream_c) {
            constexpr int KSTEPS = decltype(ksteps_c)::value;
            constexpr bool STREAM = decltype(stream_c)::value;
            uint32_t a_lo = a_lo_s0 + static_cast<uint32_t>(a_par * MT) * A_SLOT_STEP;   // window of tap (0, 0)
            uint32_t fresh = (cb == 0) ? 0u : 1u;     // 0: the chains' first MMAs of the tile overwrite
            uint32_t rot = (cb == 0) ? 0u : static_cast<uint32_t>(KS);   // KSTEPS < KS: MMAs issued so far (capped)
            for (int r = 0; r < ky; ++r) {
              for (int sx = 0; sx < kx; ++sx) {
                if (STREAM) {
                  mbar_wait(&b_full[b_stage], b_phase);
                  tc_fence_after();
                }
#pragma unroll
                for (int k = 0; k < KSTEPS; ++k) {
#pragma unroll
                  for (int mt = 0; mt < MT; ++mt) {
                    if (!mt_ok[mt]) continue;
                    if (KSTEPS >= KS) {
                      // k-step k feeds chain k % KS: the same assignment at every tap, all compile-time
                      umma_f16_lohi(d_tile + static_cast<uint32_t>((mt * KS + (k % KS)) * ACC_STRIDE),
                                    a_lo + mt * A_SLOT_STEP + 2 * k, a_hi, b_lo + k * B_KSTEP, b_hi, idesc,
                                    k < KS ? fresh : 1u);
                    } else {
                      // fewer k-steps than chains (16-channel layers): rotate over the chains from tap to tap
                      const uint32_t n = rot + k;
                      umma_f16_lohi(d_tile + static_cast<uint32_t>(mt * KS * ACC_STRIDE) + (n & (KS - 1)) * ACC_STRIDE,
                                    a_lo + mt * A_SLOT_STEP + 2 * k, a_hi, b_lo + k * B_KSTEP, b_hi, idesc,
                                    n >= KS ? 1u : 0u);
                    }
                  }
                }
                fresh = 1u;
                if (KSTEPS < KS) rot = rot + KSTEPS >= 2 * KS ? rot + KSTEPS - KS : rot + KSTEPS;   // keeps n >= KS true
                if (STREAM) {
                  if (!p.resident) umma_commit(&b_empty[b_stage]);
                }
                a_lo += 8;                // next tap: one pixel (128 B) to the right
                b_lo += B_STEP;
                if (++b_stage == B_STAGES) {
                  b_stage = 0;
                  b_phase ^= 1;
                  b_lo = b_lo_s0;
                }
              }
              a_lo += line_step;          // start of the next filter row
            }
          };
          const int ksteps = (cb == p.cblocks - 1) ? p.ksteps_last : 4;
          using std::integral_constant;
          if (stream) {
            switch (ksteps) {
              case 1: run_cb(integral_constant<int, 1>{}, integral_constant<bool, true>{}); break;
              case 2: run_cb(integral_constant<int, 2>{}, integral_constant<bool, true>{}); break;
              case 3: run_cb(integral_constant<int, 3>{}, integral_constant<bool, true>{}); break;
              default: run_cb(integral_constant<int, 4>{}, integral_constant<bool, true>{}); break;
            }
          } else {
            switch (ksteps) {
              case 1: run_cb(integral_constant<int, 1>{}, integral_constant<bool, false>{}); break;
              case 2: run_cb(integral_constant<int, 2>{}, integral_constant<bool, false>{}); break;
              case 3: run_cb(integral_constant<int, 3>{}, integral_constant<bool, false>{}); break;
              default: run_cb(integral_constant<int, 4>{}, integral_constant<bool, false>{}); break;
            }
          }
#pragma unroll
          for (int mt = 0; mt < MT; ++mt)
            if (mt_ok[mt]) umma_commit(&a_empty[a_par * MT + mt]);
          a_par ^= 1;
        }
        umma_commit(&tmem_full[acc]);
        if (tr) p.trace[ti * 8 + 3] = clock64();
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== Epilogue (warps 0..7) =====================
    const int quarter = warp & 3;        // TMEM lanes [32*quarter, 32*quarter + 32)
    const int ehalf = warp >> 2;   // which 64-column slabs this warp takes
    uint8_t* buf = epi_smem + warp * 4096;
    uint8_t* srow = buf + lane * 128;
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int nb = tile % p.n_blocks;
      const int sp = tile / p.n_blocks;
      const int img = sp / p.tiles_per_img;
      const int oh0 = (sp - img * p.tiles_per_img) * (p.R * MT);
      const int ti = (tile - static_cast<int>(blockIdx.x)) / static_cast<int>(gridDim.x);
      const bool tr = p.trace != nullptr && blockIdx.x == 0 && warp == 0 && lane == 0 && ti < TRACE_TILES;
      if (tr) p.trace[ti * 8 + 4] = clock64();
      mbar_wait_relaxed(&tmem_full[acc], acc_phase);
      tc_fence_after();
      if (tr) p.trace[ti * 8 + 5] = clock64();
#pragma unroll 1
      for (int mt = 0; mt < MT; ++mt) {
        const int ohs = oh0 + mt * p.R;
        if (ohs >= p.OH) break;
        // this lane's output pixel
        const int m = quarter * 32 + lane;
        const int r_loc = m / p.P;
        const int w = m - r_loc * p.P;
        const int oh = ohs + r_loc;
        const bool px_ok = (oh < p.OH) && (w < p.OW);
        const long long pixel = (static_cast<long long>(img) * p.OH + oh) * p.OW + w;
        // the warp's store box: lanes 0..31 = rows (w fastest, then image rows) starting here
        const int box_w0 = (quarter * 32) % p.P;
        const int box_h0 = ohs + (quarter * 32) / p.P;
        if (box_h0 >= p.OH) continue;   // warp-uniform: all of this warp's rows lie below the image
#pragma unroll 1
        for (int c0 = ehalf * 64; c0 < BN; c0 += 128) {
          const int col0 = nb * BN + c0;
          if (col0 >= p.N) break;
          if (lane == 0) tma_store_wait_read<0>();
          __syncwarp();
#pragma unroll 1
          for (int half = 0; half < 2; ++half) {
            const int colh = col0 + half * 32;
            const int nvalid = p.N - colh;   // may be <= 0 for the second half of the last slab
            if (nvalid <= 0) break;          // never stored: the box is clipped at N
            const uint32_t t0 = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                                static_cast<uint32_t>((acc * MT + mt) * KS * ACC_STRIDE + c0 + half * 32);
            uint32_t v[32];
            float f[32];
            tmem_ld_32x32(t0, v);
            if (KS == 1) {
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
            } else {
              uint32_t v2[32];   // the sub-tile is the sum of its KS chains
              tmem_ld_32x32(t0 + ACC_STRIDE, v2);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) + __uint_as_float(v2[j]);
              if (KS == 4) {
                tmem_ld_32x32(t0 + 2 * ACC_STRIDE, v);
                tmem_ld_32x32(t0 + 3 * ACC_STRIDE, v2);
                tmem_ld_wait();
#pragma unroll
                for (int j = 0; j < 32; ++j) f[j] += __uint_as_float(v[j]) + __uint_as_float(v2[j]);
              }
            }
            if (p.bias != nullptr) {
              const float* bp = p.bias + colh;
              if (nvalid >= 32 && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
                for (int j = 0; j < 32; j += 4) {
                  const float4 b4 = __ldg(reinterpret_cast<const float4*>(bp + j));
                  f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
                }
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (j < nvalid) f[j] += __ldg(bp + j);
              }
            }
            switch (p.act) {
              case ACT_TANH: halo_apply_act<ACT_TANH>(f); break;
              case ACT_SIGMOID: halo_apply_act<ACT_SIGMOID>(f); break;
              case ACT_RELU: halo_apply_act<ACT_RELU>(f); break;
              case ACT_STRICT_RELU: halo_apply_act<ACT_STRICT_RELU>(f); break;
              default: break;
            }
            if (p.Z != nullptr && px_ok) {
              const __nv_bfloat16* zrow = p.Z + pixel * p.ldz + colh;
              float d[32];
              if (nvalid >= 32 && ((p.ldz & 7) == 0) && ((reinterpret_cast<uintptr_t>(zrow) & 15) == 0)) {
#pragma unroll
                for (int j = 0; j < 32; j += 8) {
                  const float4 z = *reinterpret_cast<const float4*>(zrow + j);
                  const uint32_t* zu = reinterpret_cast<const uint32_t*>(&z);
#pragma unroll
                  for (int q = 0; q < 4; ++q) {
                    d[j + 2 * q] = bf16_lo(zu[q]);
                    d[j + 2 * q + 1] = bf16_hi(zu[q]);
                  }
                }
              } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) d[j] = (j < nvalid) ? __bfloat162float(zrow[j]) : 0.0f;
              }
              switch (p.deriv) {
                case ACT_TANH: halo_apply_deriv<ACT_TANH>(f, d); break;
                case ACT_SIGMOID: halo_apply_deriv<ACT_SIGMOID>(f, d); break;
                case ACT_RELU: halo_apply_deriv<ACT_RELU>(f, d); break;
                case ACT_STRICT_RELU: halo_apply_deriv<ACT_STRICT_RELU>(f, d); break;
                default: break;
              }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 o;
              o.x = pack_bf16x2(f[8 * q], f[8 * q + 1]);
              o.y = pack_bf16x2(f[8 * q + 2], f[8 * q + 3]);
              o.z = pack_bf16x2(f[8 * q + 4], f[8 * q + 5]);
              o.w = pack_bf16x2(f[8 * q + 6], f[8 * q + 7]);
              *reinterpret_cast<uint4*>(srow + (((half * 4 + q) ^ (lane & 7)) << 4)) = o;
            }
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            // {64 channels, min(P,32) pixels, 32/min(P,32) image rows, 1 image}: clipped to N, OW, OH by the map
            tma_store_4d(&tma_d, buf, col0, box_w0, box_h0, img);
            tma_store_commit();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty[acc]);
      if (tr) p.trace[ti * 8 + 6] = clock64();
      if (++acc == 2) {
        acc = 0;
        acc_phase ^= 1;
      }
    }
    if (lane == 0) tma_store_wait_all<0>();
  }

  tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

// ---------------------------------------------------------------------------------------------------
// Host side
// ---------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn halo_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(f);
  });
  // the driver call needs this thread to have a current context: bind the primary one once per thread
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(nullptr);
    ctx_bound = true;
  }
  return fn;
}

// bf16 NHWC tensor {C, W, H, B} (pixel pitch in elements), 128B-swizzled boxes {bc, bw, bh, 1}
static int halo_tmap_4d(CUtensorMap* out, const void* base, int C, int W, int H, int B, long long pitch, int bc, int bw,
                        int bh) {
  EncodeTiledFn fn = halo_encode_fn();
  VB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  cuuint64_t gdim[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                        static_cast<cuuint64_t>(B)};
  cuuint64_t gstride[3] = {static_cast<cuuint64_t>(pitch) * 2, static_cast<cuuint64_t>(pitch) * 2 * W,
                           static_cast<cuuint64_t>(pitch) * 2 * W * H};
  cuuint32_t box[4] = {static_cast<cuuint32_t>(bc), static_cast<cuuint32_t>(bw), static_cast<cuuint32_t>(bh), 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  VB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(4D) failed with CUresult %d (C=%d W=%d H=%d B=%d box %dx%dx%d)",
             static_cast<int>(r), C, W, H, B, bc, bw, bh);
  return 0;
}

int make_tmap_bf16_2d(CUtensorMap* out, const void* base, long long inner, long long outer, long long pitch_elems,
                      int box_inner, int box_outer);   // gemm_tcgen05.cu

template <int BN, bool B_MN, int MT>
static int launch_halo(const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& td, const HaloParams& p,
                       cudaStream_t s) {
  using S = HaloSmem<BN, MT>;
  auto kern = conv_halo_kernel<BN, B_MN, MT>;
  static bool attr_set = false;
  if (!attr_set) {
    VB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
    attr_set = true;
  }
  const long long tiles = static_cast<long long>(p.B) * p.tiles_per_img * p.n_blocks;
  const int sms = gemm_sm_count();
  const int grid = static_cast<int>(tiles < sms ? tiles : sms);
  VB_CHECK_CUDA(launch_k(kern, dim3(grid), dim3(HALO_THREADS), S::TOTAL, s, tx, tw, td, p));
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int BN, bool B_MN>
static int launch_halo_mt(int mt, const CUtensorMap& tx, const CUtensorMap& tw, const CUtensorMap& td,
                          const HaloParams& p, cudaStream_t s) {
  if (mt == 2) return launch_halo<BN, B_MN, 2>(tx, tw, td, p, s);
  return launch_halo<BN, B_MN, 1>(tx, tw, td, p, s);
}

}  // namespace vb

using namespace vb;

// Tuning / test hook: sub-tiles per CTA tile (0 = choose, 1, 2) and whether small weight sets may stay resident in
// shared memory (-1 = choose, 0 = never, 1 = whenever they fit).
static int g_halo_mt = [] {
  const char* e = getenv("VB_HALO_MT");
  return e ? atoi(e) : 0;
}();
static int g_halo_resident = [] {
  const char* e = getenv("VB_HALO_RESIDENT");
  return e ? atoi(e) : -1;
}();
// Profiling hook: a device buffer of 24 x 8 long longs that CTA 0 fills with clock64 stamps (see TRACE_TILES); null = off.
static long long* g_halo_trace = nullptr;
extern "C" int vb_conv_halo_set_trace(long long* buf) {
  g_halo_trace = buf;
  return 0;
}

extern "C" int vb_conv_halo_set_options(int mt, int resident) {
  g_halo_mt = (mt == 1 || mt == 2) ? mt : 0;
  g_halo_resident = (resident == 0 || resident == 1) ? resident : -1;
  return 0;
}

// mode 1 forward:   D[B,oH,oW,N]  = act(conv(T, W) + bias),  W = [N, ky*kx*cg] (row pitch ldw), T = [tB,tH,tW,t_ch]
// mode 2 err_input: D[B,oH,oW,cg] = convT(T = dY, W) * f'(Z), W = [w_rows = t_ch kernels, ky*kx*cg]
// (low_h, low_w) = -padding (mode 2: -(k-1-padding)); unit stride only.  T and D may be channel slices of wider
// NHWC tensors (pitches in elements).  Returns 1 when the geometry is not supported by this kernel (use
// vb_conv_gemm_bf16), 0 on success, negative on error.
extern "C" int vb_conv_halo_bf16(int mode, const void* T, int tB, int tH, int tW, long long t_pitch, int t_ch, int ky,
                                 int kx, int low_h, int low_w, int oH, int oW, const void* W, long long ldw,
                                 int w_rows, int cg, void* D, long long d_pitch, int N, const float* bias, int act,
                                 const void* Z, long long ldz, int deriv, cudaStream_t stream) {
  VB_REQUIRE(mode == 1 || mode == 2, "vb_conv_halo_bf16: mode must be 1 (forward) or 2 (err_input)");
  if (oW + kx - 1 > 64 || t_ch % 8 || t_pitch % 8 || d_pitch % 8 || ldw % 8 || kx * ky < 2) return 1;
  if ((reinterpret_cast<uintptr_t>(T) | reinterpret_cast<uintptr_t>(D) | reinterpret_cast<uintptr_t>(W)) & 15) return 1;
  int P = 16;
  while (P < oW + kx - 1) P <<= 1;
  const int R = 128 / P;
  const int a_rows = (R + ky - 1) * P;
  if (a_rows > 256 || R + ky - 1 > 256) return 1;
  HaloParams p;
  memset(&p, 0, sizeof(p));
  p.B = tB; p.OH = oH; p.OW = oW; p.P = P; p.R = R; p.kx = kx; p.ky = ky; p.low_w = low_w; p.low_h = low_h;
  p.flip = (mode == 2) ? 1 : 0;
  p.cblocks = ceil_div(t_ch, 64); p.cg = cg; p.N = N;
  p.ksteps_last = ceil_div(t_ch - (p.cblocks - 1) * 64, 16);
  p.a_rows = a_rows;
  p.bias = bias; p.act = act; p.Z = static_cast<const __nv_bfloat16*>(Z); p.ldz = ldz; p.deriv = deriv;
  p.trace = g_halo_trace;
  // tile width: the narrowest of 64 / 96 / 128 that holds N in one tile, else 128-wide tiles
  const int BN = (N <= 64) ? 64 : ((N <= 96 && mode == 1) ? 96 : 128);
  p.n_blocks = ceil_div(N, BN);
  // A layer whose whole weight set fits beside two activation slots keeps it in shared memory (AlexNet conv1 after
  // space-to-depth: 9 taps x 96 kernels x 128 B = 108 KB): only activations stream.  Everything else streams weight
  // tiles and shares each of them between two sub-tiles.
  const int taps = kx * ky;
  const int stages_mt1 = (227 * 1024 - 2 * HALO_A_SLOT - HALO_EPI_BYTES - 2048) / (BN * 128);
  const bool fits = p.n_blocks == 1 && taps * p.cblocks <= (stages_mt1 > 16 ? 16 : stages_mt1);
  p.resident = (g_halo_resident < 0 ? fits : (g_halo_resident && fits)) ? 1 : 0;
  int mt = g_halo_mt ? g_halo_mt : (p.resident ? 1 : 2);
  if (p.resident || oH <= R || BN == 96) mt = 1;   // 96-wide tiles exist for single sub-tiles only
  p.tiles_per_img = ceil_div(oH, R * mt);
  {
    // every accumulation chain of a sub-tile must receive at least one MMA (see KS in the kernel)
    const int stride = BN <= 64 ? 64 : 128;
    const int ks = 256 / (mt * stride) < 1 ? 1 : 256 / (mt * stride);
    // every accumulation chain of a sub-tile must receive an MMA: a single channel block with fewer k-steps than
    // chains rotates over them tap by tap
    if (p.cblocks == 1 && p.ksteps_last < ks && taps * p.ksteps_last < ks) return 1;
  }
  CUtensorMap tx, tw, td;
  int rc = halo_tmap_4d(&tx, T, t_ch, tW, tH, tB, t_pitch, 64, P, R + ky - 1);
  if (rc) return rc;
  const int pw = P < 32 ? P : 32;
  rc = halo_tmap_4d(&td, D, N, oW, oH, tB, d_pitch, 64, pw, 32 / pw);
  if (rc) return rc;
  const long long kk = static_cast<long long>(kx) * ky * cg;
  if (mode == 1) rc = make_tmap_bf16_2d(&tw, W, kk, w_rows, ldw, 64, BN);
  else rc = make_tmap_bf16_2d(&tw, W, kk, w_rows, ldw, 64, 64);
  if (rc) return rc;
  if (mode == 1) {
    if (BN == 64) return launch_halo_mt<64, false>(mt, tx, tw, td, p, stream);
    if (BN == 96) return launch_halo<96, false, 1>(tx, tw, td, p, stream);   // resident / single sub-tile only
    return launch_halo_mt<128, false>(mt, tx, tw, td, p, stream);
  }
  if (BN == 64) return launch_halo_mt<64, true>(mt, tx, tw, td, p, stream);
  return launch_halo_mt<128, true>(mt, tx, tw, td, p, stream);
}
