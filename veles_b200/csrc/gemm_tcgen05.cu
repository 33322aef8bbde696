// veles_b200 — tcgen05/TMEM/TMA GEMM family for sm_100a.
//
//   D[M,N] = epilogue( sum_k A(m,k) * B(n,k) )
//
// A and B are bf16 in global memory in either "K-major" (row-major [M,K] /
// [N,K]) or "MN-major" (row-major [K,M] / [K,N]) storage, which covers the
// three products a Znicz layer needs without ever materialising a transpose:
//   forward      Y  = X  * W^T   (A K-major,  B K-major )   ref: znicz all2all (absent)
//   backward-dX  dX = dY * W     (A K-major,  B MN-major)   ref: znicz gd err_input
//   backward-dW  dW = dY^T * X   (A MN-major, B MN-major)   ref: znicz gd weights gradient
// replacing the reference's cuBLAS sgemm/dgemm call sites
// (veles/accelerated_units.py:761-766, veles/backends.py:829-836) and the
// OpenCL tiled GEMM with its STORE_OUTPUT epilogue include
// (ocl/matrix_multiplication_precise.cl:43-45).
//
// Kernel anatomy (one persistent CTA per SM, 320 threads):
//   warps 0-7: epilogue       (tcgen05.ld -> bias/activation/derivative -> smem -> TMA store)
//   warp 8   : TMA producer   (cp.async.bulk.tensor -> 128B-swizzled smem ring), one elected thread
//   warp 9   : TMEM allocator + single-thread tcgen05.mma issuer
// The fp32 accumulator lives in TMEM, double-buffered (2 x BN columns) so the
// epilogue of tile i overlaps the mainloop of tile i+1.
#include "common.cuh"

#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <tuple>

namespace vb {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;
constexpr int GEMM_THREADS = 320;  // 8 epilogue warps, producer warp, MMA warp
constexpr int EPI_WARPS = 8;
// The two single-thread roles sit in the HIGHEST warps: the warp scheduler prefers the highest eligible warp id of a
// sub-partition, and the epilogue warps that share it spend most of their time polling mbarriers - with the roles in
// warps 0 and 1 their (pacing) instruction streams kept losing the issue slot to those polls.
constexpr int PRODUCER_WARP = 8;
constexpr int MMA_WARP = 9;

// Implicit-GEMM convolution: one operand is fetched by TMA in im2col mode straight from the NHWC
// activation tensor (no column matrix is ever materialised).  mode:
//   0 plain GEMM
//   1 forward   A = im2col(X)  [pixels x (tap, c)]      B = W  K-major  [n, tap*cg + c]
//   2 backward  A = im2col(dY) with flipped taps         B = W  MN-major [k = n, n = c] per tap   (stride 1)
//   3 weights   A = dY MN-major [k = pixel, m = n]       B = im2col(X) MN-major [k = pixel, n = (tap, c)]
struct ConvParams {
  int mode;
  int kx, ky;        // filter taps
  int OW, OH;        // base positions per row / rows per image of the im2col traversal
  int sx, sy;        // traversal strides
  int low_w, low_h;  // bounding-box lower corner (= -pad; -(k-1-pad) for mode 2)
  int cblocks;       // 64-channel blocks per tap along the im2col tensor's channel dimension
  int cg;            // weight-matrix columns per tap
  int q64, r64;      // 64 / OW and 64 % OW: how the traversal position advances per 64-pixel k-block (mode 3)
  int nbpg;          // grouped convolution in one launch: N tiles per group (0 = not grouped); tile nb belongs to
                     // group nb / nbpg, whose im2col channels start at group * cblocks * 64
};

// Row-sharded B operand ("peer-TMA weight delivery"): rank r of a data-parallel group owns rows
// [r*b_shard_rows, (r+1)*b_shard_rows) of the [N, K] weight matrix (the slice its update kernel
// writes); every other rank's copy of those rows is stale.  The producer therefore loads each B
// tile through the tensor map of the OWNER's replica: a peer address over NVLink/NVSwitch, so
// the all-gather of updated weights rides inside the GEMM's operand pipeline.
struct PeerMaps {
  CUtensorMap m[8];
};

struct GemmParams {
  int M, N, K;
  int b_shard_rows;  // 0: B is read from tma_b only
  ConvParams cv;
  void* D;          // bf16 or fp32, row-major, leading dimension ldd (elements)
  long long ldd;
  const float* bias;  // [N] or null
  int act;            // vb::Act applied to (acc*alpha + bias)
  int out_fp32;       // 1: D is float, 0: D is bf16
  int accumulate;     // 1: atomically add into fp32 D (split-K / gradient accumulation)
  const __nv_bfloat16* Z;  // optional [M,N] saved activation output: D *= f'(Z)
  long long ldz;
  int deriv;          // vb::Act code of the derivative taken through Z
  int z_fp32;         // 1: Z is float (TF32 mode), 0: bf16
  float alpha;
  int split_k;
  int dyn_sched;      // 1: tiles are handed out by cluster launch control (grid = one CTA / pair per tile)
  int tma_store;      // 1: epilogue goes through smem + TMA store / reduce-add (tma_d valid)
  int m_blocks, n_blocks, k_blocks;
  // Fused gradient-descent epilogue (gd_w != null): the accumulator tile IS the weight gradient; instead of storing
  // it, the epilogue reads the fp32 master weights and momentum at the tile's position, applies the Znicz update and
  // writes weights, momentum and the bf16 shadow - the gradient never reaches HBM.
  float* gd_w;
  float* gd_vel;
  __nv_bfloat16* gd_shadow;   // may be null
  long long gd_ld, gd_ld_shadow;
  float gd_lr, gd_wd, gd_l12, gd_moment;
};

__device__ __forceinline__ float gd_step_momentum(const GemmParams& p, float w, float g, float& v) {
  const float sgn = (w > 0.0f) ? 1.0f : ((w < 0.0f) ? -1.0f : 0.0f);
  g = g + p.gd_wd * ((1.0f - p.gd_l12) * w + 0.5f * p.gd_l12 * sgn);
  v = p.gd_moment * v - p.gd_lr * g;
  return w + v;
}

template <int BN, bool PAIR>
struct SmemLayout {
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;
  static constexpr int B_ROWS = PAIR ? BN / 2 : BN;  // a CTA of a pair stages half of the B tile
  static constexpr int B_BYTES = B_ROWS * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int BAR_BYTES = 1024;
  static constexpr int EPI_BYTES = EPI_WARPS * 4096;  // one 32-row x 128 B slab per epilogue warp
  static constexpr int BUDGET = 227 * 1024 - EPI_BYTES - BAR_BYTES - 1024;
  static constexpr int STAGES = (BUDGET / STAGE_BYTES) > 8 ? 8 : (BUDGET / STAGE_BYTES);
  static constexpr int TOTAL = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024 /*align slack*/;
};
constexpr int EPI_BUF_BYTES = 4096;

__device__ __forceinline__ float4 ld_z8(const __nv_bfloat16* p) {
  return *reinterpret_cast<const float4*>(p);
}

// Tile rasterisation: N-blocks run fastest inside bands of up to 8 N-blocks, so the tiles in
// flight at any moment share their A row-strips (A is the big operand of the layer GEMMs: it is
// then fetched from HBM once and hit in L2 by the other N-blocks) while a band's B strips stay
// L2-resident across the whole sweep over M.
constexpr int RASTER_BAND = 8;
__device__ __forceinline__ void decode_tile(int mn, int m_blocks, int n_blocks, int& mb, int& nb) {
  const int band_tiles = RASTER_BAND * m_blocks;
  const int band = mn / band_tiles;
  const int within = mn - band * band_tiles;
  const int n0 = band * RASTER_BAND;
  const int w = min(RASTER_BAND, n_blocks - n0);
  mb = within / w;
  nb = n0 + (within - mb * w);
}

// Epilogue math on a 32-column register slab with the activation fixed at compile time.
template <int ACT>
__device__ __forceinline__ void apply_act(float (&f)[32]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = act_forward_fast<ACT>(f[j]);
}
template <int ACT>
__device__ __forceinline__ void apply_deriv(float (&f)[32], const float (&y)[32]) {
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] *= act_deriv_from_output(ACT, y[j]);
}

// PAIR = true: the kernel runs as clusters of two CTAs (one TPC) driving tcgen05.mma.cta_group::2.
// A pair owns a 256 x BN tile: CTA r stages A rows [r*128, r*128+128) and B rows
// [r*BN/2, (r+1)*BN/2) of the tile, so every SM ingests 128 + BN/2 operand rows per k-block
// instead of 128 + BN; the leader's MMA reads both halves (own smem + the peer's over the pair's
// shared-memory path) and each CTA receives its 128 accumulator rows in its own TMEM.
// TF32 = true: fp32 operands in global memory, `kind::tf32` MMAs (10-bit mantissa products, fp32 accumulate) -
// the "float" precision mode.  A 128-byte swizzle row then holds 32 elements, so a k-block is 32 elements deep and
// an MN-major chunk 32 elements wide; stage sizes, the four MMAs per k-block and everything downstream are unchanged.
template <int BN, bool A_MN, bool B_MN, bool PAIR, bool TF32>
__global__ void __launch_bounds__(GEMM_THREADS, 1)
gemm_bf16_tcgen05_kernel(const __grid_constant__ CUtensorMap tma_a,
                         const __grid_constant__ CUtensorMap tma_b,
                         const __grid_constant__ CUtensorMap tma_d, const __grid_constant__ PeerMaps tma_bp,
                         const GemmParams p) {
  using L = SmemLayout<BN, PAIR>;
  constexpr int STAGES = L::STAGES;
  constexpr int TILE_M = PAIR ? 2 * BLOCK_M : BLOCK_M;
  constexpr int KE = TF32 ? 32 : 64;        // elements per k-block (one 128-byte row)
  constexpr int MNC = TF32 ? 32 : 64;       // elements per MN-major chunk (128 bytes along M / N)
  constexpr int CHUNK_BYTES = KE * 128;     // an MN-major chunk: KE k-rows of 128 bytes
  constexpr int UK = TF32 ? 8 : 16;         // k extent of one MMA
  constexpr int BN_LOAD = L::B_ROWS;  // B rows / columns this CTA stages per k-block
  const uint32_t cta_rank = PAIR ? cluster_ctarank() : 0u;
  const bool leader = (cta_rank == 0);
  // Accumulation chains.  A tcgen05.mma that accumulates into the columns the previous one wrote waits for it: measured
  // ~170 cycles from one dependent MMA to the next whatever N is (8192^3 at N = 256: 190; conv2 at N = 128: 153;
  // conv1 halo at N = 96: 190), so a single chain keeps the tensor pipe busy for (N / 2) / 170 of the time - 75 % at
  // N = 256, 38 % at N = 128, 19 % at N = 64.  Narrow tiles therefore split the k-steps of a k-block over KS
  // independent accumulators (k-step k -> chain k % KS), which the epilogue adds: KS * BN = 256 columns in flight.
  constexpr int KS = BN <= 64 ? 4 : (BN <= 128 ? 2 : 1);
  constexpr int TMEM_NEED = 2 * KS * BN;
  constexpr int TMEM_COLS = TMEM_NEED <= 32 ? 32 : (TMEM_NEED <= 64 ? 64 : (TMEM_NEED <= 128 ? 128 : (TMEM_NEED <= 256 ? 256 : 512)));
  static_assert(TMEM_NEED <= 512, "accumulators exceed tensor memory");

  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* epi_smem = smem + STAGES * L::STAGE_BYTES;  // 1024-aligned: swizzle atoms line up
  uint8_t* bar_base = epi_smem + L::EPI_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bar_base);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  // dynamic tile scheduler state: two 16-byte response slots + full/empty barriers
  uint64_t* clc_resp = reinterpret_cast<uint64_t*>(bar_base + 512);   // [2][2]
  uint64_t* clc_full = clc_resp + 4;
  uint64_t* clc_empty = clc_full + 2;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == PRODUCER_WARP && elect_one()) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
    if (p.tma_store) tma_prefetch_desc(&tma_d);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tmem_full[s], 1);
      mbar_init(&tmem_empty[s], PAIR ? 2 * EPI_WARPS : EPI_WARPS);  // one arrive per epilogue warp (of both CTAs)
      mbar_init(&clc_full[s], 1);
      // readers of a response: producer thread + MMA warp + epilogue warps (the peer CTA has no MMA loop)
      mbar_init(&clc_empty[s], PAIR ? 2 * (EPI_WARPS + 1) + 1 : EPI_WARPS + 2);
    }
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == MMA_WARP) {
    if (PAIR) tmem_alloc_pair<TMEM_COLS>(tmem_ptr);
    else tmem_alloc<TMEM_COLS>(tmem_ptr);
  }
  tc_fence_before();
  if (PAIR) cluster_sync_all();  // the peer's barriers exist before anything signals them
  else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  // barriers, tensor memory and descriptor prefetches above overlap the previous kernel's tail; global memory is
  // touched only from here on
  pdl_sync();

  const int tiles_mn = p.m_blocks * p.n_blocks;
  const int num_tiles = tiles_mn * p.split_k;
  const int kb_per_split = (p.k_blocks + p.split_k - 1) / p.split_k;
  // persistent schedule: one work unit = one CTA (or one CTA pair).  Static: tiles unit, unit + n_units, ...
  // Dynamic (p.dyn_sched): the grid has one unit per tile; a running unit starts with its own tile and then
  // keeps cancelling pending units and taking their tiles (cluster launch control), so SMs held by another
  // kernel (the data-parallel update kernel on its side stream) merely shrink the pool instead of delaying
  // the tiles statically assigned to them.
  const int unit = PAIR ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int n_units = PAIR ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  // request #it is issued while tile #it is being worked on; its answer names tile #it+1
  auto sched_request = [&](int it) {   // producer thread of every CTA
    const int slot = it & 1;
    if (leader && it >= 2) mbar_wait(&clc_empty[slot], static_cast<uint32_t>(((it >> 1) & 1) ^ 1));
    mbar_expect_tx(&clc_full[slot], 16);
    if (leader) clc_try_cancel(clc_resp + 2 * slot, &clc_full[slot], PAIR);
  };
  auto sched_next = [&](int tile, int it, bool arrive, bool whole_warp) -> int {
    if (!p.dyn_sched) return tile + n_units;
    const int slot = it & 1;
    mbar_wait(&clc_full[slot], static_cast<uint32_t>((it >> 1) & 1));
    uint32_t x;
    const bool ok = clc_query(clc_resp + 2 * slot, x);
    if (whole_warp) __syncwarp();   // every lane has read the slot before lane 0 releases it
    if (arrive) {
      if (PAIR) mbar_arrive_cluster(mapa_u32(&clc_empty[slot], 0));
      else mbar_arrive(&clc_empty[slot]);
    }
    return ok ? static_cast<int>(PAIR ? (x >> 1) : x) : num_tiles;
  };

  // Both single-thread roles below are bound by their own instruction stream, not by the units they feed (ncu
  // source view, AlexNet conv2 forward: the producer thread spent its samples on the integer divisions that turned
  // a k-block index into (tap, channel block, filter row) - ~115 dependent instructions per k-block = ~800 cycles
  // for 256 cycles of tensor work).  So every per-k-block quantity is carried incrementally, and descriptors are
  // built once and advanced by constants.
  if (warp == PRODUCER_WARP) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      const int mode = p.cv.mode;
      const int per_img = p.cv.OH * p.cv.OW;
      for (int tile = unit; tile < num_tiles; tile = sched_next(tile, it, true, false), ++it) {
        if (p.dyn_sched) sched_request(it);
        const int ks = tile / tiles_mn;
        const int mn = tile - ks * tiles_mn;
        int mb, nb;
        decode_tile(mn, p.m_blocks, p.n_blocks, mb, nb);
        const int kb0 = ks * kb_per_split;
        const int kb1 = min(p.k_blocks, kb0 + kb_per_split);
        const int m_base = mb * TILE_M + static_cast<int>(cta_rank) * BLOCK_M;   // first A row staged here
        const int n_base = nb * BN + static_cast<int>(cta_rank) * BN_LOAD;       // first B row/column staged here
        // ---- per-tile set-up of the incremental walk (the only divisions of the producer) ----
        int a_w = 0, a_h = 0, a_n = 0;       // modes 1, 2: first pixel of this CTA's rows in traversal coordinates
        int cb = 0, tap_r = 0, tap_s = 0;    // modes 1, 2: channel block / filter row / filter column of k-block kb
        int grp_c0 = 0, n_local = 0;
        int p_img = 0, p_oh = 0, p_ow = 0;   // mode 3: image / row / column of the first pixel of k-block kb
        int c_ch[4], c_wo[4], c_ho[4];       // mode 3: channel start and tap of each 64-column chunk staged here
        if (mode == 1 || mode == 2) {
          a_n = m_base / per_img;
          const int rem = m_base - a_n * per_img;
          const int oh = rem / p.cv.OW;
          a_h = oh * p.cv.sy + p.cv.low_h;
          a_w = (rem - oh * p.cv.OW) * p.cv.sx + p.cv.low_w;
          const int tap = kb0 / p.cv.cblocks;
          cb = kb0 - tap * p.cv.cblocks;
          tap_r = tap / p.cv.kx;
          tap_s = tap - tap_r * p.cv.kx;
          // grouped layers in one launch: this N tile's group selects the channel range of the walked tensor
          const int grp = p.cv.nbpg > 0 ? nb / p.cv.nbpg : 0;
          grp_c0 = grp * p.cv.cblocks * 64;
          n_local = n_base - grp * p.cv.nbpg * BN;   // column inside the group's [kernels, taps*cg] slab
        } else if (mode == 3) {
          const int px0 = kb0 * BLOCK_K;
          p_img = px0 / per_img;
          const int rem = px0 - p_img * per_img;
          p_oh = rem / p.cv.OW;
          p_ow = rem - p_oh * p.cv.OW;
          const int nblk = p.cv.kx * p.cv.ky * p.cv.cblocks;
#pragma unroll
          for (int c = 0; c < BN_LOAD / 64; ++c) {
            const int blk = (n_base >> 6) + c;
            if (blk < nblk) {
              const int tap = blk / p.cv.cblocks;
              const int r = tap / p.cv.kx;
              c_ch[c] = (blk - tap * p.cv.cblocks) * 64;
              c_wo[c] = tap - r * p.cv.kx;
              c_ho[c] = r;
            } else {  // past the last tap: channel coordinate out of range -> a box of zeros
              c_ch[c] = p.cv.cblocks * 64;
              c_wo[c] = 0;
              c_ho[c] = 0;
            }
          }
        }
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * L::STAGE_BYTES;
          uint8_t* sb = sa + L::A_BYTES;
          // completion lands on the leader's barrier, armed for the bytes of both CTAs
          const uint32_t fb = PAIR ? mapa_u32(&full_bar[stage], 0) : smem_u32(&full_bar[stage]);
          if (leader) mbar_expect_tx(&full_bar[stage], (PAIR ? 2 : 1) * L::STAGE_BYTES);
          auto ld2 = [&](void* dst, const CUtensorMap* m, int c0, int c1) {
            if (PAIR) tma_load_2d_pair(dst, m, fb, c0, c1);
            else tma_load_2d(dst, m, &full_bar[stage], c0, c1);
          };
          auto ldi = [&](void* dst, const CUtensorMap* m, int c, int w, int h, int n, int wo, int ho) {
            if (PAIR) tma_load_im2col_4d_pair(dst, m, fb, c, w, h, n, static_cast<uint16_t>(wo), static_cast<uint16_t>(ho));
            else tma_load_im2col_4d(dst, m, &full_bar[stage], c, w, h, n, static_cast<uint16_t>(wo), static_cast<uint16_t>(ho));
          };
          if (mode == 0) {
            if (!A_MN) {
              ld2(sa, &tma_a, kb * KE, m_base);
            } else {
#pragma unroll
              for (int c = 0; c < BLOCK_M / MNC; ++c)
                ld2(sa + c * CHUNK_BYTES, &tma_a, m_base + c * MNC, kb * KE);
            }
            // weight rows of this tile: N index for K-major B, K index for MN-major B (dX = dY * W)
            const CUtensorMap* mb_ptr = &tma_b;
            if (p.b_shard_rows > 0) mb_ptr = &tma_bp.m[(B_MN ? kb * KE : n_base) / p.b_shard_rows];
            if (!B_MN) {
              ld2(sb, mb_ptr, kb * KE, n_base);
            } else {
#pragma unroll
              for (int c = 0; c < BN_LOAD / MNC; ++c)
                ld2(sb + c * CHUNK_BYTES, mb_ptr, n_base + c * MNC, kb * KE);
            }
          } else if (mode == 3) {
            // A: dY, MN-major tiled.  B: one im2col box (64 pixels x 64 channels) per 64-column chunk.
#pragma unroll
            for (int c = 0; c < BLOCK_M / 64; ++c)
              ld2(sa + c * (BLOCK_K * 128), &tma_a, m_base + c * 64, kb * BLOCK_K);
            const int bh = p_oh * p.cv.sy + p.cv.low_h;
            const int bw = p_ow * p.cv.sx + p.cv.low_w;
#pragma unroll
            for (int c = 0; c < BN_LOAD / 64; ++c)
              ldi(sb + c * (BLOCK_K * 128), &tma_b, c_ch[c], bw, bh, p_img, c_wo[c], c_ho[c]);
            // the next k-block starts 64 pixels further along the traversal
            p_ow += p.cv.r64;
            p_oh += p.cv.q64;
            if (p_ow >= p.cv.OW) {
              p_ow -= p.cv.OW;
              ++p_oh;
            }
            while (p_oh >= p.cv.OH) {
              p_oh -= p.cv.OH;
              ++p_img;
            }
          } else {
            const int ch0 = grp_c0 + cb * 64;
            ldi(sa, &tma_a, ch0, a_w, a_h, a_n, tap_s, tap_r);
            if (mode == 1) {
              ld2(sb, &tma_b, (tap_r * p.cv.kx + tap_s) * p.cv.cg + cb * 64, n_base);
            } else {
              // dX: the dY pixel reached through offset (s', r') pairs with filter tap (ky-1-r', kx-1-s')
              const int wtap = (p.cv.ky - 1 - tap_r) * p.cv.kx + (p.cv.kx - 1 - tap_s);
#pragma unroll
              for (int c = 0; c < BN_LOAD / 64; ++c)
                ld2(sb + c * (BLOCK_K * 128), &tma_b, wtap * p.cv.cg + n_local + c * 64, ch0);
            }
            if (++cb == p.cv.cblocks) {
              cb = 0;
              if (++tap_s == p.cv.kx) {
                tap_s = 0;
                ++tap_r;
              }
            }
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == MMA_WARP) {
    // ===================== MMA issuer =====================
    // ONE thread of the leader CTA runs the whole loop (waits included): no per-k-block election / reconvergence.
    if (leader && elect_one()) {
      constexpr uint32_t idesc = make_idesc(/*BF16 = 1, TF32 = 2*/ TF32 ? 2 : 1, A_MN, B_MN, TILE_M, BN);
      // K-major : rows of 128 B, 8-row swizzle atoms 1024 B apart (SBO); k-step = +32 B.
      // MN-major: 64-element column chunks BLOCK_K*128 B apart (LBO), 8 k-rows per
      //           1024 B atom (SBO); k-step (16 rows) = +2048 B.
      // MN-major tf32 operands use the 32-byte-atom flavour of the 128-byte swizzle: 4 k-rows per pattern
      constexpr uint32_t MN_LAYOUT = TF32 ? 1u : 2u;
      constexpr uint32_t MN_SBO = TF32 ? 512u : 1024u;
      constexpr uint64_t STAGE_STEP = static_cast<uint64_t>(L::STAGE_BYTES >> 4);   // descriptor address units
      constexpr uint64_t A_KSTEP = static_cast<uint64_t>((A_MN ? UK * 128 : 32) >> 4);
      constexpr uint64_t B_KSTEP = static_cast<uint64_t>((B_MN ? UK * 128 : 32) >> 4);
      const uint32_t sa0 = smem_u32(smem);
      const uint64_t a_desc_s0 =
          A_MN ? make_smem_desc(sa0, CHUNK_BYTES, MN_SBO, MN_LAYOUT) : make_smem_desc_sw128(sa0, 16, 1024);
      const uint64_t b_desc_s0 = B_MN ? make_smem_desc(sa0 + L::A_BYTES, CHUNK_BYTES, MN_SBO, MN_LAYOUT)
                                      : make_smem_desc_sw128(sa0 + L::A_BYTES, 16, 1024);
      // (shared-memory addresses stay below 2^18, so adding stage offsets never carries out of the 14-bit field)
      uint64_t a_desc = a_desc_s0, b_desc = b_desc_s0;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      int it = 0;
      for (int tile = unit; tile < num_tiles; tile = sched_next(tile, it, true, false), ++it) {
        const int ks = tile / tiles_mn;
        const int kb0 = ks * kb_per_split;
        const int kb1 = min(p.k_blocks, kb0 + kb_per_split);
        if (kb0 >= kb1) continue;  // empty split: the epilogue writes zeros without TMEM
        mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem0 = tmem_base + static_cast<uint32_t>(acc * KS * BN);
        uint32_t accf = 0u;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&full_bar[stage], phase);
          tc_fence_after();
#pragma unroll
          for (int k = 0; k < KE / UK; ++k) {   // 4 MMAs per k-block: +32 B along a K-major row, +UK rows MN-major
            const uint64_t ad = a_desc + k * A_KSTEP;
            const uint64_t bd = b_desc + k * B_KSTEP;
            const uint32_t d_tmem = d_tmem0 + static_cast<uint32_t>((k % KS) * BN);   // consecutive MMAs: other chains
            const uint32_t af = k < KS ? accf : 1u;   // a chain's first MMA of the tile overwrites
            if (TF32) {
              if (PAIR) umma_tf32_pair(d_tmem, ad, bd, idesc, af);
              else umma_tf32(d_tmem, ad, bd, idesc, af);
            } else {
              if (PAIR) umma_f16_pair(d_tmem, ad, bd, idesc, af);
              else umma_f16(d_tmem, ad, bd, idesc, af);
            }
          }
          accf = 1u;
          // frees the smem slot (in both CTAs of a pair) once these MMAs retire
          if (PAIR) umma_commit_pair(&empty_bar[stage]);
          else umma_commit(&empty_bar[stage]);
          a_desc += STAGE_STEP;
          b_desc += STAGE_STEP;
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
            a_desc = a_desc_s0;
            b_desc = b_desc_s0;
          }
        }
        if (PAIR) umma_commit_pair(&tmem_full[acc]);
        else umma_commit(&tmem_full[acc]);
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== Epilogue (warps 0..7) =====================
    // Two warps per TMEM lane quarter (a warp may only touch lanes 32*(warp%4)..+31): they take
    // alternate column slabs, so every SM sub-partition has two epilogue warps to interleave and
    // the dependent tcgen05.ld -> math -> st.shared chain of one hides behind the other.
    const int quarter = warp & 3;
    const int ehalf = warp >> 2;
    int acc = 0;
    uint32_t acc_phase = 0;
    int it = 0;
    for (int tile = unit; tile < num_tiles; tile = sched_next(tile, it, lane == 0, true), ++it) {
      const int ks = tile / tiles_mn;
      const int mn = tile - ks * tiles_mn;
      int mb, nb;
      decode_tile(mn, p.m_blocks, p.n_blocks, mb, nb);
      const int kb0 = ks * kb_per_split;
      const bool has_k = kb0 < p.k_blocks;
      const bool add_bias = (p.bias != nullptr) && (ks == 0);
      if (has_k) {
        mbar_wait_relaxed(&tmem_full[acc], acc_phase);
        tc_fence_after();
      }
      const int row0 = mb * TILE_M + static_cast<int>(cta_rank) * BLOCK_M + quarter * 32;
      const int row = row0 + lane;
      const bool row_ok = row < p.M;
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 32) {
        const int col0 = nb * BN + c0;
        if (col0 >= p.N) break;  // warp-uniform
        // slab ownership: 32-column fp32 slabs / 64-column bf16 slab pairs alternate between the two warps
        if ((((p.out_fp32 && p.tma_store) ? (c0 >> 5) : (c0 >> 6)) & 1) != ehalf) continue;
        int dcol0 = col0;          // first column of this slab in D
        int nvalid = p.N - col0;   // columns of the slab that exist in D
        if (p.cv.mode == 3) {      // virtual (tap, 64-channel block) columns -> weight-matrix columns
          const int blk = col0 >> 6;
          const int tap = blk / p.cv.cblocks;
          const int cpos = (blk - tap * p.cv.cblocks) * 64 + (col0 & 63);
          if (cpos >= p.cv.cg) continue;  // channel padding only
          dcol0 = tap * p.cv.cg + cpos;
          nvalid = p.cv.cg - cpos;
        }
        float f[32];
        if (has_k) {
          const uint32_t t0 = tmem_base + (static_cast<uint32_t>(quarter * 32) << 16) +
                              static_cast<uint32_t>(acc * KS * BN + c0);
          uint32_t v[32];
          tmem_ld_32x32(t0, v);
          if (KS == 1) {
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]);
          } else {
            // the tile is the sum of its KS accumulation chains (two loads in flight at a time)
            uint32_t v2[32];
            tmem_ld_32x32(t0 + BN, v2);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = __uint_as_float(v[j]) + __uint_as_float(v2[j]);
            if (KS == 4) {
              tmem_ld_32x32(t0 + 2 * BN, v);
              tmem_ld_32x32(t0 + 3 * BN, v2);
              tmem_ld_wait();
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] += __uint_as_float(v[j]) + __uint_as_float(v2[j]);
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = 0.0f;
        }
        if (p.alpha != 1.0f) {
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] *= p.alpha;
        }
        const bool full = nvalid >= 32;
        if (p.gd_w != nullptr) {
          // The accumulator slab has one weight ROW per lane; touching w / momentum that way would make every
          // load and store instruction hit 32 different cache lines.  So the slab is transposed through this warp's
          // (swizzled) staging buffer, and the update runs with 8 lanes x 16 bytes along each row: four full
          // 128-byte row segments per instruction, for w, momentum and the bf16 shadow alike.
          uint8_t* buf = epi_smem + warp * EPI_BUF_BYTES;
          uint8_t* srow = buf + lane * 128;
          __syncwarp();
#pragma unroll
          for (int q = 0; q < 8; ++q)
            *reinterpret_cast<float4*>(srow + ((q ^ (lane & 7)) << 4)) =
                make_float4(f[4 * q], f[4 * q + 1], f[4 * q + 2], f[4 * q + 3]);
          __syncwarp();
          const int q = lane & 7;                 // 16-byte chunk (4 columns) of a row
          const int cbase = dcol0 + 4 * q;
          const int cvalid = nvalid - 4 * q;      // columns of this chunk that exist
          const bool vec_ok = ((p.gd_ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(p.gd_w) & 15) == 0) &&
                              ((reinterpret_cast<uintptr_t>(p.gd_vel) & 15) == 0);
          const bool svec_ok = p.gd_shadow != nullptr && ((p.gd_ld_shadow & 3) == 0) &&
                               ((reinterpret_cast<uintptr_t>(p.gd_shadow) & 7) == 0);
          float4 g4[8], w4[8], v4[8];
          bool ok[8];
#pragma unroll
          for (int i2 = 0; i2 < 8; ++i2) {         // all loads of the slab first
            const int r = i2 * 4 + (lane >> 3);
            const int grow = row0 + r;
            ok[i2] = grow < p.M && cvalid > 0;
            g4[i2] = *reinterpret_cast<const float4*>(buf + r * 128 + ((q ^ (r & 7)) << 4));
            if (ok[i2]) {
              const float* wr = p.gd_w + static_cast<long long>(grow) * p.gd_ld + cbase;
              const float* vr = p.gd_vel + static_cast<long long>(grow) * p.gd_ld + cbase;
              if (vec_ok && cvalid >= 4) {
                w4[i2] = *reinterpret_cast<const float4*>(wr);
                v4[i2] = *reinterpret_cast<const float4*>(vr);
              } else {
                w4[i2].x = wr[0]; v4[i2].x = vr[0];
                w4[i2].y = cvalid > 1 ? wr[1] : 0.0f; v4[i2].y = cvalid > 1 ? vr[1] : 0.0f;
                w4[i2].z = cvalid > 2 ? wr[2] : 0.0f; v4[i2].z = cvalid > 2 ? vr[2] : 0.0f;
                w4[i2].w = cvalid > 3 ? wr[3] : 0.0f; v4[i2].w = cvalid > 3 ? vr[3] : 0.0f;
              }
            }
          }
#pragma unroll
          for (int i2 = 0; i2 < 8; ++i2) {
            if (!ok[i2]) continue;
            const int grow = row0 + i2 * 4 + (lane >> 3);
            w4[i2].x = gd_step_momentum(p, w4[i2].x, g4[i2].x, v4[i2].x);
            w4[i2].y = gd_step_momentum(p, w4[i2].y, g4[i2].y, v4[i2].y);
            w4[i2].z = gd_step_momentum(p, w4[i2].z, g4[i2].z, v4[i2].z);
            w4[i2].w = gd_step_momentum(p, w4[i2].w, g4[i2].w, v4[i2].w);
            float* wr = p.gd_w + static_cast<long long>(grow) * p.gd_ld + cbase;
            float* vr = p.gd_vel + static_cast<long long>(grow) * p.gd_ld + cbase;
            __nv_bfloat16* sr = p.gd_shadow ? p.gd_shadow + static_cast<long long>(grow) * p.gd_ld_shadow + cbase : nullptr;
            if (vec_ok && cvalid >= 4) {
              *reinterpret_cast<float4*>(wr) = w4[i2];
              *reinterpret_cast<float4*>(vr) = v4[i2];
            } else {
              wr[0] = w4[i2].x; vr[0] = v4[i2].x;
              if (cvalid > 1) { wr[1] = w4[i2].y; vr[1] = v4[i2].y; }
              if (cvalid > 2) { wr[2] = w4[i2].z; vr[2] = v4[i2].z; }
              if (cvalid > 3) { wr[3] = w4[i2].w; vr[3] = v4[i2].w; }
            }
            if (sr != nullptr) {
              if (svec_ok && cvalid >= 4) {
                *reinterpret_cast<uint2*>(sr) = make_uint2(pack_bf16x2(w4[i2].x, w4[i2].y), pack_bf16x2(w4[i2].z, w4[i2].w));
              } else {
                sr[0] = __float2bfloat16(w4[i2].x);
                if (cvalid > 1) sr[1] = __float2bfloat16(w4[i2].y);
                if (cvalid > 2) sr[2] = __float2bfloat16(w4[i2].z);
                if (cvalid > 3) sr[3] = __float2bfloat16(w4[i2].w);
              }
            }
          }
          continue;
        }
        if (add_bias) {
          const float* bp = p.bias + dcol0;
          if (full && (reinterpret_cast<uintptr_t>(bp) & 15) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 b4 = __ldg(reinterpret_cast<const float4*>(bp + j));
              f[j] += b4.x; f[j + 1] += b4.y; f[j + 2] += b4.z; f[j + 3] += b4.w;
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (full || j < nvalid) f[j] += __ldg(bp + j);
          }
        }
        // The activation code is warp-uniform: branch once, keep f[] in registers.
        switch (p.act) {
          case ACT_TANH: apply_act<ACT_TANH>(f); break;
          case ACT_SIGMOID: apply_act<ACT_SIGMOID>(f); break;
          case ACT_RELU: apply_act<ACT_RELU>(f); break;
          case ACT_STRICT_RELU: apply_act<ACT_STRICT_RELU>(f); break;
          default: break;
        }
        if (p.Z != nullptr && row_ok) {
          const __nv_bfloat16* zrow = p.Z + static_cast<long long>(row) * p.ldz + dcol0;
          float d[32];
          if (p.z_fp32) {
            const float* zf = reinterpret_cast<const float*>(p.Z) + static_cast<long long>(row) * p.ldz + dcol0;
#pragma unroll
            for (int j = 0; j < 32; ++j) d[j] = (j < nvalid) ? zf[j] : 0.0f;
          } else if (full && ((p.ldz & 7) == 0)) {
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              float4 z = ld_z8(zrow + j);
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
            case ACT_TANH: apply_deriv<ACT_TANH>(f, d); break;
            case ACT_SIGMOID: apply_deriv<ACT_SIGMOID>(f, d); break;
            case ACT_RELU: apply_deriv<ACT_RELU>(f, d); break;
            case ACT_STRICT_RELU: apply_deriv<ACT_STRICT_RELU>(f, d); break;
            default: break;
          }
        }
        if (p.tma_store) {
          // Stage the 32-row slab in 128B-swizzled smem (conflict-free 16-byte writes: the eight
          // lanes of a quarter-warp phase hit eight different chunks), then one TMA store /
          // L2 reduce-add per slab: full-line writes, M/N tails clipped by the tensor map.
          uint8_t* buf = epi_smem + warp * EPI_BUF_BYTES;
          uint8_t* srow = buf + lane * 128;
          if (p.out_fp32) {
            if (lane == 0) tma_store_wait_read<0>();  // the previous slab has left this warp's buffer
            __syncwarp();
#pragma unroll
            for (int q = 0; q < 8; ++q)
              *reinterpret_cast<float4*>(srow + ((q ^ (lane & 7)) << 4)) =
                  make_float4(f[4 * q], f[4 * q + 1], f[4 * q + 2], f[4 * q + 3]);
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
              if (p.accumulate) tma_reduce_add_2d(&tma_d, buf, dcol0, row0);
              else tma_store_2d(&tma_d, buf, dcol0, row0);
              tma_store_commit();
            }
          } else {
            const int half = (c0 >> 5) & 1;
            if (half == 0) {
              if (lane == 0) tma_store_wait_read<0>();
              __syncwarp();
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
            if (half == 1 || col0 + 32 >= p.N || c0 + 32 >= BN) {
              fence_proxy_async_smem();
              __syncwarp();
              if (lane == 0) {
                tma_store_2d(&tma_d, buf, col0 - half * 32, row0);
                tma_store_commit();
              }
            }
          }
        } else if (row_ok) {
          if (p.out_fp32) {
            float* drow = reinterpret_cast<float*>(p.D) + static_cast<long long>(row) * p.ldd + dcol0;
            if (p.accumulate) {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (full || j < nvalid) atomicAdd(drow + j, f[j]);
            } else if (full && ((p.ldd & 3) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(drow + j) = make_float4(f[j], f[j + 1], f[j + 2], f[j + 3]);
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nvalid) drow[j] = f[j];
            }
          } else {
            __nv_bfloat16* drow =
                reinterpret_cast<__nv_bfloat16*>(p.D) + static_cast<long long>(row) * p.ldd + dcol0;
            if (full && ((p.ldd & 7) == 0)) {
#pragma unroll
              for (int j = 0; j < 32; j += 8) {
                uint4 o;
                o.x = pack_bf16x2(f[j], f[j + 1]);
                o.y = pack_bf16x2(f[j + 2], f[j + 3]);
                o.z = pack_bf16x2(f[j + 4], f[j + 5]);
                o.w = pack_bf16x2(f[j + 6], f[j + 7]);
                *reinterpret_cast<uint4*>(drow + j) = o;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (j < nvalid) drow[j] = __float2bfloat16(f[j]);
            }
          }
        }
      }
      if (has_k) {
        tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (PAIR) mbar_arrive_cluster(mapa_u32(&tmem_empty[acc], 0));  // the leader's MMA warp waits on it
          else mbar_arrive(&tmem_empty[acc]);
        }
        if (++acc == 2) {
          acc = 0;
          acc_phase ^= 1;
        }
      }
    }
    if (p.tma_store && lane == 0) tma_store_wait_all<0>();  // smem must outlive the bulk reads
  }

  tc_fence_before();
  if (PAIR) cluster_sync_all();  // nobody may still signal barriers or read operands of a CTA that exits
  else __syncthreads();
  if (warp == MMA_WARP) {
    tc_fence_after();
    if (PAIR) tmem_dealloc_pair<TMEM_COLS>(tmem_base);
    else tmem_dealloc<TMEM_COLS>(tmem_base);
  }
}

// ----------------------------------------------------------------------------
// Host side: tensor-map construction (driver entry point fetched at run time so
// the library never links libcuda) and launch.
// ----------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
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

// 2-D bf16 tensor map with 128B swizzle: dims {inner, outer}, row pitch in elements.
static int make_tmap_2d(CUtensorMap* out, const void* base, long long inner, long long outer,
                        long long pitch_elems, int box_inner, int box_outer, bool fp32, bool atom32 = false);

int make_tmap_bf16_2d(CUtensorMap* out, const void* base, long long inner, long long outer,
                      long long pitch_elems, int box_inner, int box_outer) {
  return make_tmap_2d(out, base, inner, outer, pitch_elems, box_inner, box_outer, false);
}

static int make_tmap_2d(CUtensorMap* out, const void* base, long long inner, long long outer,
                        long long pitch_elems, int box_inner, int box_outer, bool fp32, bool atom32) {
  const int es = fp32 ? 4 : 2;
  EncodeTiledFn fn = get_encode_fn();
  VB_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled entry point unavailable");
  VB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base pointer must be 16-byte aligned");
  VB_REQUIRE((pitch_elems * es) % 16 == 0, "TMA row pitch must be a multiple of 16 bytes, got %lld elements of %d bytes", pitch_elems, es);
  cuuint64_t gdim[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t gstride[1] = {static_cast<cuuint64_t>(pitch_elems) * es};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, fp32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstride, box,
                  estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  VB_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (inner=%lld outer=%lld pitch=%lld box=%dx%d)",
             static_cast<int>(r), inner, outer, pitch_elems, box_inner, box_outer);
  return 0;
}

// Peer tensor maps of the launch being assembled on this host thread (all-zero when B is not
// sharded): set by vb_gemm_bf16_sharded_b around its call, read by launch_gemm.
static PeerMaps g_zero_peer_maps;
static thread_local const PeerMaps* g_peer_maps = &g_zero_peer_maps;

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const int*, const int*, cuuint32_t, cuuint32_t,
                                   const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeIm2colFn get_encode_im2col_fn() {
  static EncodeIm2colFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeIm2colFn>(f);
  });
  // the driver call needs this thread to have a current context: bind the primary one once per thread
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) {
    cudaFree(nullptr);
    ctx_bound = true;
  }
  return fn;
}

// NHWC bf16 tensor walked in im2col mode: `n_w x n_h` base positions per image starting at
// (low_w, low_h) with traversal strides (sx, sy); a box is `pixels` consecutive base positions
// (wrapping rows and images) x `channels` channels, 128B-swizzled; out-of-image taps read zeros.
static int make_tmap_im2col_bf16(CUtensorMap* out, const void* base, int C, int W, int H, int B, long long pitch,
                                 int low_w, int low_h, int n_w, int n_h, int sx, int sy, int channels, int pixels) {
  EncodeIm2colFn fn = get_encode_im2col_fn();
  VB_REQUIRE(fn != nullptr, "cuTensorMapEncodeIm2col entry point unavailable");
  VB_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "im2col TMA base pointer must be 16-byte aligned");
  cuuint64_t gdim[4] = {static_cast<cuuint64_t>(C), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H),
                        static_cast<cuuint64_t>(B)};
  cuuint64_t gstride[3] = {static_cast<cuuint64_t>(pitch) * 2, static_cast<cuuint64_t>(pitch) * 2 * W,
                           static_cast<cuuint64_t>(pitch) * 2 * W * H};
  // upper corner chosen so that exactly n_w x n_h base positions exist
  int lower[2] = {low_w, low_h};
  int upper[2] = {low_w + (n_w - 1) * sx - (W - 1), low_h + (n_h - 1) * sy - (H - 1)};
  cuuint32_t estr[4] = {1, static_cast<cuuint32_t>(sx), static_cast<cuuint32_t>(sy), 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), gdim, gstride, lower, upper,
                  static_cast<cuuint32_t>(channels), static_cast<cuuint32_t>(pixels), estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  VB_REQUIRE(r == CUDA_SUCCESS,
             "cuTensorMapEncodeIm2col failed with CUresult %d (C=%d W=%d H=%d B=%d pitch=%lld lower=%d,%d upper=%d,%d)",
             static_cast<int>(r), C, W, H, B, pitch, lower[0], lower[1], upper[0], upper[1]);
  return 0;
}

template <int BN, bool A_MN, bool B_MN, bool PAIR, bool TF32 = false>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& td,
                       const GemmParams& p, int grid, cudaStream_t stream) {
  const PeerMaps& tp = *g_peer_maps;
  using L = SmemLayout<BN, PAIR>;
  auto kern = gemm_bf16_tcgen05_kernel<BN, A_MN, B_MN, PAIR, TF32>;
  static bool attr_set = false;
  if (!attr_set) {
    VB_CHECK_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::TOTAL));
    attr_set = true;
  }
  if (PAIR) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(GEMM_THREADS);
    cfg.dynamicSmemBytes = L::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = g_pdl_enabled ? 2 : 1;
    VB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, td, tp, p));
  } else {
    VB_CHECK_CUDA(launch_k(kern, dim3(grid), dim3(GEMM_THREADS), L::TOTAL, stream, ta, tb, td, tp, p));
  }
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

template <int BN, bool PAIR, bool TF32 = false>
static int dispatch_major(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb,
                          const CUtensorMap& td, const GemmParams& p, int grid, cudaStream_t s) {
  if (!a_mn && !b_mn) return launch_gemm<BN, false, false, PAIR, TF32>(ta, tb, td, p, grid, s);
  if (!a_mn && b_mn) return launch_gemm<BN, false, true, PAIR, TF32>(ta, tb, td, p, grid, s);
  if (a_mn && !b_mn) return launch_gemm<BN, true, false, PAIR, TF32>(ta, tb, td, p, grid, s);
  return launch_gemm<BN, true, true, PAIR, TF32>(ta, tb, td, p, grid, s);
}

}  // namespace vb

using namespace vb;

static int g_disable_tma_store = 0;

// Test hook: 1 forces the per-thread store epilogue so both paths stay covered.
extern "C" int vb_gemm_set_direct_epilogue(int on) {
  g_disable_tma_store = on ? 1 : 0;
  return 0;
}

// Tile width and split-K selection.  The mainloop is bound by the per-SM operand ingest
// (128 + BN rows of 128 B per k-block), so the cost of a candidate is
//   waves * (k-blocks per split + per-tile overhead) * (128 + BN)
// and the widest tile wins unless N-tail waste or a ragged last wave says otherwise.
static int g_dyn_sched = [] {
  const char* e = getenv("VB_GEMM_DYNAMIC");
  return e ? atoi(e) : 0;
}();   // 1: tiles handed out by cluster launch control (see the kernel); VB_GEMM_DYNAMIC overrides the default
static int g_pair_mode_override = 0;  // 1: this call may not pair (sharded-B retry)
static int g_pair_mode = 0;  // 0 = choose by cost, 1 = never pair, 2 = pair whenever the shape allows (tests)

static void choose_config(int M, int N, int k_blocks, bool can_split, bool auto_split, bool b_mn, int* bn_io,
                          int* split_io, int* pair_out) {
  const int sms = gemm_sm_count();
  static const int cands[4] = {64, 128, 192, 256};
  double best = 1e300;
  int best_bn = 128, best_split = 1, best_pair = 0;
  for (int pair = 0; pair < 2; ++pair) {
    if (pair && (g_pair_mode == 1 || g_pair_mode_override || M <= BLOCK_M)) continue;
    const int tile_m = pair ? 2 * BLOCK_M : BLOCK_M;
    const int units = pair ? sms / 2 : sms;
    const int m_blocks = ceil_div(M, tile_m);
    for (int ci = 0; ci < 4; ++ci) {
      const int bn = cands[ci];
      if (*bn_io != 0 && bn != *bn_io) continue;
      if (*bn_io == 0 && bn > 64 && bn - 64 >= N) continue;  // a narrower candidate already covers N
      if (pair && (bn < 128 || (b_mn && (bn / 2) % 64 != 0))) continue;
      const long long mn_tiles = static_cast<long long>(m_blocks) * ceil_div(N, bn);
      int split = *split_io;
      if (auto_split) {
        split = 1;
        if (can_split && mn_tiles * 2 <= units && k_blocks >= 8) {
          long long sp = units / mn_tiles;
          const long long cap = k_blocks / 4;
          if (sp > cap) sp = cap;
          if (sp > 1) split = static_cast<int>(sp);
        }
      }
      if (split > k_blocks) split = k_blocks;
      const int per = ceil_div(k_blocks, split);
      split = ceil_div(k_blocks, per);  // every split non-empty
      const long long waves = (mn_tiles * split + units - 1) / units;
      double cost = static_cast<double>(waves) * (per + 2.0) * (128 + (pair ? bn / 2 : bn));
      if (pair && g_pair_mode == 2) cost *= 1e-6;
      if (cost < best) {
        best = cost;
        best_bn = bn;
        best_split = split;
        best_pair = pair;
      }
    }
  }
  *bn_io = best_bn;
  *split_io = best_split;
  *pair_out = best_pair;
}

struct GdRequest {
  float* w;
  float* vel;
  void* shadow;
  long long ld, ld_shadow;
  float lr, wd, l12, moment;
};
static thread_local const GdRequest* g_gd_request = nullptr;

// Builds the output tensor map (when D qualifies) and launches the right instantiation.
static int finish_and_launch(GemmParams& p, const CUtensorMap& ta, const CUtensorMap& tb, bool a_mn, bool b_mn,
                             int BN, int pair, cudaStream_t stream, int d_cols = -1, bool tf32 = false) {
  const int sms = gemm_sm_count();
  // Output through TMA whenever D satisfies the tensor-map alignment rules; otherwise the
  // epilogue falls back to per-thread global stores / atomics.
  CUtensorMap td;
  memset(&td, 0, sizeof(td));
  const int es = p.out_fp32 ? 4 : 2;
  p.tma_store = ((reinterpret_cast<uintptr_t>(p.D) & 15) == 0 && (p.ldd * es) % 16 == 0 && !g_disable_tma_store) ? 1 : 0;
  if (g_gd_request != nullptr) {
    const GdRequest& r = *g_gd_request;
    p.gd_w = r.w; p.gd_vel = r.vel; p.gd_shadow = static_cast<__nv_bfloat16*>(r.shadow);
    p.gd_ld = r.ld; p.gd_ld_shadow = r.ld_shadow;
    p.gd_lr = r.lr; p.gd_wd = r.wd; p.gd_l12 = r.l12; p.gd_moment = r.moment;
    p.tma_store = 0;       // nothing is stored through D
  }
  if (p.tma_store) {
    int rc = make_tmap_2d(&td, p.D, d_cols > 0 ? d_cols : p.N, p.M, p.ldd, p.out_fp32 ? 32 : 64, 32, p.out_fp32 != 0);
    if (rc) return rc;
  }
  const long long tiles = static_cast<long long>(p.m_blocks) * p.n_blocks * p.split_k;
  // dynamic (cluster-launch-control) scheduling: one unit per tile in the grid, running units steal the rest
  p.dyn_sched = (g_dyn_sched && tiles < (1LL << 30)) ? 1 : 0;
  if (pair) {
    const int units = sms / 2;
    const int grid = 2 * static_cast<int>((p.dyn_sched || tiles < units) ? tiles : units);
    if (tf32) {
      switch (BN) {
        case 128: return dispatch_major<128, true, true>(a_mn, b_mn, ta, tb, td, p, grid, stream);
        case 256: return dispatch_major<256, true, true>(a_mn, b_mn, ta, tb, td, p, grid, stream);
        default: VB_REQUIRE(false, "tf32 CTA-pair tiles are 128 or 256 wide");
      }
    }
    switch (BN) {
      case 128: return dispatch_major<128, true>(a_mn, b_mn, ta, tb, td, p, grid, stream);
      case 192:
        VB_REQUIRE(!b_mn, "paired 192-wide tiles need a K-major B operand");
        if (a_mn) return launch_gemm<192, true, false, true>(ta, tb, td, p, grid, stream);
        return launch_gemm<192, false, false, true>(ta, tb, td, p, grid, stream);
      default: return dispatch_major<256, true>(a_mn, b_mn, ta, tb, td, p, grid, stream);
    }
  }
  const int grid = static_cast<int>((p.dyn_sched || tiles < sms) ? tiles : sms);
  if (tf32) {
    switch (BN) {
      case 64: return dispatch_major<64, false, true>(a_mn, b_mn, ta, tb, td, p, grid, stream);
      case 128: return dispatch_major<128, false, true>(a_mn, b_mn, ta, tb, td, p, grid, stream);
      case 256: return dispatch_major<256, false, true>(a_mn, b_mn, ta, tb, td, p, grid, stream);
      default: VB_REQUIRE(false, "tf32 tiles are 64, 128 or 256 wide");
    }
  }
  switch (BN) {
    case 64: return dispatch_major<64, false>(a_mn, b_mn, ta, tb, td, p, grid, stream);
    case 128: return dispatch_major<128, false>(a_mn, b_mn, ta, tb, td, p, grid, stream);
    case 192: return dispatch_major<192, false>(a_mn, b_mn, ta, tb, td, p, grid, stream);
    default: return dispatch_major<256, false>(a_mn, b_mn, ta, tb, td, p, grid, stream);
  }
}

// 1: dynamic (cluster launch control) tile scheduling, 0: static round-robin over a persistent grid.
extern "C" int vb_gemm_set_dynamic_schedule(int on) {
  g_dyn_sched = on ? 1 : 0;
  return 0;
}

// Test hook: 0 = cost model decides, 1 = single-CTA tiles only, 2 = CTA pairs whenever the shape allows.
extern "C" int vb_gemm_set_pair_mode(int mode) {
  g_pair_mode = (mode == 1 || mode == 2) ? mode : 0;
  return 0;
}

static int gemm_impl(const void* A, long long lda, int a_mn_major, const void* B, long long ldb,
                     int b_mn_major, void* D, long long ldd, int M, int N, int K,
                     const float* bias, int act, int out_fp32, int accumulate, const void* Z,
                     long long ldz, int deriv, float alpha, int split_k, int block_n,
                     const void* const* b_peers, int world, int shard_rows, cudaStream_t stream) {
  VB_REQUIRE(M > 0 && N > 0 && K > 0, "vb_gemm_bf16: empty problem %dx%dx%d", M, N, K);
  VB_REQUIRE(!(accumulate && !out_fp32), "vb_gemm_bf16: accumulate requires fp32 output");
  VB_REQUIRE(split_k >= 0, "vb_gemm_bf16: split_k must be >= 0 (0 = choose automatically)");
  const bool auto_split = (split_k == 0);
  if (auto_split) split_k = 1;
  VB_REQUIRE(split_k == 1 || (accumulate && act == 0 && Z == nullptr),
             "vb_gemm_bf16: split-K needs accumulate=1 and a linear epilogue");
  VB_REQUIRE(block_n == 0 || block_n == 64 || block_n == 128 || block_n == 192 || block_n == 256,
             "vb_gemm_bf16: block_n must be 0/64/128/192/256");
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N; p.K = K;
  p.D = D; p.ldd = ldd; p.bias = bias; p.act = act; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.Z = reinterpret_cast<const __nv_bfloat16*>(Z); p.ldz = ldz; p.deriv = deriv; p.alpha = alpha;
  p.k_blocks = ceil_div(K, BLOCK_K);
  int BN = block_n, pair = 0;
  {
    // a row-sharded K-major B needs every staged B tile inside one owner's shard: narrow the tile until it does
    const int forced[4] = {block_n, 128, 64, 64};
    const int split_req = split_k;
    for (int attempt = 0; attempt < 4; ++attempt) {
      BN = forced[attempt];
      split_k = split_req;
      // peer replicas are read by plain single-CTA TMA loads only: every CTA fetches its own tile
      if (attempt == 3 || b_peers != nullptr) g_pair_mode_override = 1;
      choose_config(M, N, p.k_blocks, accumulate && act == 0 && Z == nullptr, auto_split, b_mn_major != 0, &BN,
                    &split_k, &pair);
      g_pair_mode_override = 0;
      if (b_peers == nullptr || b_mn_major || shard_rows % (pair ? BN / 2 : BN) == 0) break;
      if (block_n != 0 && attempt == 0) break;  // the caller fixed the tile: report the mismatch below
    }
  }
  p.m_blocks = ceil_div(M, pair ? 2 * BLOCK_M : BLOCK_M);
  p.n_blocks = ceil_div(N, BN);
  p.split_k = split_k;
  const int bn_load = pair ? BN / 2 : BN;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap_bf16_2d(&ta, A, K, M, lda, BLOCK_K, BLOCK_M);
  else rc = make_tmap_bf16_2d(&ta, A, M, K, lda, 64, BLOCK_K);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_bf16_2d(&tb, B, K, N, ldb, BLOCK_K, bn_load);
  else rc = make_tmap_bf16_2d(&tb, B, N, K, ldb, 64, BLOCK_K);
  if (rc) return rc;
  PeerMaps peers;
  if (b_peers != nullptr) {
    // every replica has the same shape and pitch; only the base address differs
    VB_REQUIRE(world >= 1 && world <= 8 && shard_rows > 0, "sharded B: bad world %d / shard_rows %d", world, shard_rows);
    VB_REQUIRE(!pair, "sharded B: peer replicas are read by single-CTA tiles only");
    const int w_rows = b_mn_major ? K : N;   // the weight-matrix row index the shard is cut along
    VB_REQUIRE(static_cast<long long>(shard_rows) * world >= w_rows, "sharded B: %d ranks x %d rows do not cover %d rows",
               world, shard_rows, w_rows);
    VB_REQUIRE(shard_rows % (b_mn_major ? BLOCK_K : bn_load) == 0,
               "sharded B: shard of %d rows is not a multiple of the %d-row tile", shard_rows,
               b_mn_major ? BLOCK_K : bn_load);
    memset(&peers, 0, sizeof(peers));
    for (int r = 0; r < world; ++r) {
      if (!b_mn_major) rc = make_tmap_bf16_2d(&peers.m[r], b_peers[r], K, N, ldb, BLOCK_K, bn_load);
      else rc = make_tmap_bf16_2d(&peers.m[r], b_peers[r], N, K, ldb, 64, BLOCK_K);
      if (rc) return rc;
    }
    p.b_shard_rows = shard_rows;
    g_peer_maps = &peers;
  }
  rc = finish_and_launch(p, ta, tb, a_mn_major != 0, b_mn_major != 0, BN, pair, stream);
  g_peer_maps = &g_zero_peer_maps;
  return rc;
}

// What the scheduler would pick for a plain GEMM: tile width, split-K factor, CTA pairing.  Callers use it to
// decide whether a gradient buffer needs accumulate (split-K) semantics or can be plainly stored.
extern "C" int vb_gemm_plan(int M, int N, int K, int b_mn_major, int can_split, int block_n, int* bn_out,
                            int* split_out, int* pair_out) {
  VB_REQUIRE(M > 0 && N > 0 && K > 0, "vb_gemm_plan: empty problem");
  int bn = block_n, split = 1, pair = 0;
  choose_config(M, N, ceil_div(K, BLOCK_K), can_split != 0, true, b_mn_major != 0, &bn, &split, &pair);
  if (bn_out) *bn_out = bn;
  if (split_out) *split_out = split;
  if (pair_out) *pair_out = pair;
  return 0;
}

extern "C" int vb_gemm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb,
                            int b_mn_major, void* D, long long ldd, int M, int N, int K,
                            const float* bias, int act, int out_fp32, int accumulate, const void* Z,
                            long long ldz, int deriv, float alpha, int split_k, int block_n,
                            cudaStream_t stream) {
  return gemm_impl(A, lda, a_mn_major, B, ldb, b_mn_major, D, ldd, M, N, K, bias, act, out_fp32, accumulate, Z, ldz,
                   deriv, alpha, split_k, block_n, nullptr, 0, 0, stream);
}

// Weight-gradient GEMM with the gradient-descent step in its epilogue (single GPU, momentum solver, no split-K):
//   g = alpha * sum_k A(m,k) B(n,k) + wd * ((1 - l1_vs_l2) * w + 0.5 * l1_vs_l2 * sign(w));  v = moment * v - lr * g;  w += v
// applied to the fp32 master weights `w` / momentum `vel` ([M, N], pitch ldw) and mirrored into the bf16 `shadow`
// ([M, N], pitch ld_shadow, may be null).  The gradient itself is never written.
extern "C" int vb_gemm_bf16_fused_gd(const void* A, long long lda, int a_mn_major, const void* B, long long ldb,
                                     int b_mn_major, int M, int N, int K, float* w, long long ldw, float* vel,
                                     void* shadow, long long ld_shadow, float lr, float wd, float l1_vs_l2,
                                     float moment, float alpha, int block_n, cudaStream_t stream) {
  VB_REQUIRE(w != nullptr && vel != nullptr, "vb_gemm_bf16_fused_gd: weights and momentum are required");
  VB_REQUIRE(ldw >= N && (shadow == nullptr || ld_shadow >= N), "vb_gemm_bf16_fused_gd: pitch smaller than N");
  GdRequest req{w, vel, shadow, ldw, ld_shadow, lr, wd, l1_vs_l2, moment};
  g_gd_request = &req;
  // D is only a formal argument here (fp32, pitch ldw): the epilogue takes the GD branch before any store
  const int rc = gemm_impl(A, lda, a_mn_major, B, ldb, b_mn_major, w, ldw, M, N, K, nullptr, 0, 1, 0, nullptr, 0, 0,
                           alpha, 1, block_n, nullptr, 0, 0, stream);
  g_gd_request = nullptr;
  return rc;
}

// Same product with the B (weight) operand row-sharded over `world` replicas: rows
// [r*shard_rows, (r+1)*shard_rows) are read from b_peers[r] (device pointers to the same matrix in
// each rank's symmetric memory; b_peers[own rank] is the local replica).  Fuses the all-gather of
// freshly updated weight shards into the GEMM's TMA operand loads (see PeerMaps).
extern "C" int vb_gemm_bf16_sharded_b(const void* A, long long lda, int a_mn_major, const void* const* b_peers,
                                      int world, int shard_rows, int own_rank, long long ldb, int b_mn_major, void* D,
                                      long long ldd, int M, int N, int K, const float* bias, int act, int out_fp32,
                                      int accumulate, const void* Z, long long ldz, int deriv, float alpha,
                                      int split_k, int block_n, cudaStream_t stream) {
  VB_REQUIRE(b_peers != nullptr && world >= 1 && world <= 8 && own_rank >= 0 && own_rank < world,
             "vb_gemm_bf16_sharded_b: bad peer list");
  return gemm_impl(A, lda, a_mn_major, b_peers[own_rank], ldb, b_mn_major, D, ldd, M, N, K, bias, act, out_fp32,
                   accumulate, Z, ldz, deriv, alpha, split_k, block_n, b_peers, world, shard_rows, stream);
}

// ----------------------------------------------------------------------------
// Implicit-GEMM convolution entry point (see ConvParams).  `T` is the NHWC tensor walked in
// im2col mode (X for modes 1/3, dY for mode 2) with any group channel offset already applied:
// tB x tH x tW pixels, `t_pitch` elements between pixels, `t_ch` channels visible.  `other` is
// the weight matrix [rows, taps*cg] (modes 1, 2) or dY [pixels, M] (mode 3).
// ----------------------------------------------------------------------------
static int conv_gemm_impl(int mode, const void* T, int tB, int tH, int tW, long long t_pitch, int t_ch,
                          int ky, int kx, int sy, int sx, int low_h, int low_w, int oH, int oW,
                          const void* other, long long ld_other, int other_rows, int cg, void* D,
                          long long ldd, int M, int N, const float* bias, int act, int out_fp32,
                          int accumulate, const void* Z, long long ldz, int deriv, float alpha,
                          int split_k, int block_n, int groups, cudaStream_t stream) {
  VB_REQUIRE(mode >= 1 && mode <= 3, "vb_conv_gemm_bf16: mode must be 1 (forward), 2 (err_input) or 3 (weights)");
  VB_REQUIRE(tB > 0 && tH > 0 && tW > 0 && oH > 0 && oW > 0 && M > 0 && N > 0, "vb_conv_gemm_bf16: empty problem");
  VB_REQUIRE(t_ch % 8 == 0 && t_pitch % 8 == 0, "vb_conv_gemm_bf16: channels (%d) and pixel pitch (%lld) must be multiples of 8", t_ch, t_pitch);
  VB_REQUIRE(kx >= 1 && ky >= 1 && kx <= 32 && ky <= 32, "vb_conv_gemm_bf16: filter %dx%d out of range", ky, kx);
  VB_REQUIRE(!(accumulate && !out_fp32), "vb_conv_gemm_bf16: accumulate requires fp32 output");
  VB_REQUIRE(mode != 2 || (sx == 1 && sy == 1), "vb_conv_gemm_bf16: err_input mode needs unit stride");
  const long long pixels = static_cast<long long>(tB) * oH * oW;
  VB_REQUIRE(pixels < (1LL << 31), "vb_conv_gemm_bf16: too many pixels");
  const int taps = kx * ky;
  VB_REQUIRE(groups >= 1, "vb_conv_gemm_bf16: groups must be >= 1");
  if (groups > 1) {
    // all groups in one launch: T / other / D are the WHOLE tensors, N the total output width
    VB_REQUIRE(mode != 3, "vb_conv_gemm_bf16: grouped launches cover forward and err_input");
    VB_REQUIRE(t_ch % groups == 0 && (t_ch / groups) % 64 == 0 && N % groups == 0 && (N / groups) % 64 == 0,
               "vb_conv_gemm_bf16: grouped launch needs per-group channel counts that are multiples of 64");
  }
  const int cblocks = ceil_div(t_ch / groups, 64);
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.cv.mode = mode; p.cv.kx = kx; p.cv.ky = ky; p.cv.OW = oW; p.cv.OH = oH; p.cv.sx = sx; p.cv.sy = sy;
  p.cv.low_w = low_w; p.cv.low_h = low_h; p.cv.cblocks = cblocks; p.cv.cg = cg;
  p.cv.q64 = BLOCK_K / oW; p.cv.r64 = BLOCK_K % oW;
  p.D = D; p.ldd = ldd; p.bias = bias; p.act = act; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.Z = reinterpret_cast<const __nv_bfloat16*>(Z); p.ldz = ldz; p.deriv = deriv; p.alpha = alpha;
  const int sms = gemm_sm_count();
  int n_virtual = N;
  if (mode == 3) {
    VB_REQUIRE(out_fp32, "vb_conv_gemm_bf16: the weight gradient is fp32");
    VB_REQUIRE(N == taps * cg && cg <= t_ch, "vb_conv_gemm_bf16: N must be taps*cg");
    VB_REQUIRE(accumulate || cg % 32 == 0, "vb_conv_gemm_bf16: cg %% 32 != 0 needs accumulate=1");
    VB_REQUIRE(bias == nullptr && act == 0 && Z == nullptr, "vb_conv_gemm_bf16: weights mode has a linear epilogue");
    n_virtual = taps * cblocks * 64;
  }
  VB_REQUIRE(block_n == 0 || block_n == 64 || block_n == 128 || block_n == 192 || block_n == 256,
             "vb_conv_gemm_bf16: block_n must be 0/64/128/192/256");
  const bool can_split = accumulate && act == 0 && Z == nullptr;
  VB_REQUIRE(split_k >= 0 && (split_k <= 1 || can_split), "vb_conv_gemm_bf16: split-K needs accumulate=1 and a linear epilogue");
  const int kb_total = (mode == 3) ? ceil_div(static_cast<int>(pixels), BLOCK_K) : taps * cblocks;
  int BN = block_n, pair = 0;
  {
    const bool auto_split = (split_k == 0);
    if (auto_split) split_k = 1;
    choose_config(M, n_virtual, kb_total, can_split, auto_split, mode != 1, &BN, &split_k, &pair);
    if (groups > 1 && (N / groups) % BN != 0) {
      // a tile may not straddle two groups: widest tile that divides the per-group width
      const int npg = N / groups;
      static const int widths[4] = {256, 192, 128, 64};
      for (int i = 0; i < 4; ++i) {
        if (npg % widths[i] != 0) continue;
        BN = widths[i];
        int sk = auto_split ? 1 : split_k;
        choose_config(M, n_virtual, kb_total, can_split, auto_split, mode != 1, &BN, &sk, &pair);
        split_k = sk;
        break;
      }
      if (pair && (mode != 1) && (BN / 2) % 64 != 0) pair = 0;
    }
  }
  if (groups > 1) p.cv.nbpg = (N / groups) / BN;
  p.M = M; p.N = n_virtual;
  p.m_blocks = ceil_div(M, pair ? 2 * BLOCK_M : BLOCK_M);
  p.n_blocks = ceil_div(n_virtual, BN);
  int rc;
  CUtensorMap ta, tb;
  bool a_mn = false, b_mn = false;
  if (mode == 1 || mode == 2) {
    VB_REQUIRE(static_cast<long long>(M) == pixels, "vb_conv_gemm_bf16: M must equal tB*oH*oW");
    p.k_blocks = taps * cblocks;
    p.K = p.k_blocks * BLOCK_K;
    rc = make_tmap_im2col_bf16(&ta, T, t_ch, tW, tH, tB, t_pitch, low_w, low_h, oW, oH, sx, sy, 64, BLOCK_M);
    if (rc) return rc;
    if (mode == 1) {
      rc = make_tmap_bf16_2d(&tb, other, static_cast<long long>(taps) * cg, other_rows, ld_other, BLOCK_K, pair ? BN / 2 : BN);
    } else {
      b_mn = true;
      rc = make_tmap_bf16_2d(&tb, other, static_cast<long long>(taps) * cg, other_rows, ld_other, 64, BLOCK_K);
    }
    if (rc) return rc;
  } else {
    a_mn = true; b_mn = true;
    p.K = static_cast<int>(pixels);
    p.k_blocks = ceil_div(p.K, BLOCK_K);
    rc = make_tmap_bf16_2d(&ta, other, M, pixels, ld_other, 64, BLOCK_K);
    if (rc) return rc;
    rc = make_tmap_im2col_bf16(&tb, T, t_ch, tW, tH, tB, t_pitch, low_w, low_h, oW, oH, sx, sy, 64, BLOCK_K);
    if (rc) return rc;
  }
  p.split_k = split_k;
  // D is described with its real extent so stores of padded / neighbouring-tap columns are clipped
  p.N = n_virtual;
  return finish_and_launch(p, ta, tb, a_mn, b_mn, BN, pair, stream, mode == 3 ? N : -1);
}

extern "C" int vb_conv_gemm_bf16(int mode, const void* T, int tB, int tH, int tW, long long t_pitch, int t_ch,
                                 int ky, int kx, int sy, int sx, int low_h, int low_w, int oH, int oW,
                                 const void* other, long long ld_other, int other_rows, int cg, void* D,
                                 long long ldd, int M, int N, const float* bias, int act, int out_fp32,
                                 int accumulate, const void* Z, long long ldz, int deriv, float alpha,
                                 int split_k, int block_n, cudaStream_t stream) {
  return conv_gemm_impl(mode, T, tB, tH, tW, t_pitch, t_ch, ky, kx, sy, sx, low_h, low_w, oH, oW, other, ld_other,
                        other_rows, cg, D, ldd, M, N, bias, act, out_fp32, accumulate, Z, ldz, deriv, alpha, split_k,
                        block_n, 1, stream);
}

// All `groups` of a grouped convolution in ONE launch (modes 1 and 2): T, other (the whole [n_kernels, taps*cg] weight
// matrix) and D are the full tensors, t_ch / N the total channel counts; tile -> group by its N index.  Needs
// per-group channel and kernel counts that are multiples of 64 (AlexNet conv4 / conv5).
extern "C" int vb_conv_gemm_grouped_bf16(int mode, int groups, const void* T, int tB, int tH, int tW, long long t_pitch,
                                         int t_ch, int ky, int kx, int sy, int sx, int low_h, int low_w, int oH,
                                         int oW, const void* other, long long ld_other, int other_rows, int cg,
                                         void* D, long long ldd, int M, int N, const float* bias, int act,
                                         int out_fp32, const void* Z, long long ldz, int deriv, int block_n,
                                         cudaStream_t stream) {
  return conv_gemm_impl(mode, T, tB, tH, tW, t_pitch, t_ch, ky, kx, sy, sx, low_h, low_w, oH, oW, other, ld_other,
                        other_rows, cg, D, ldd, M, N, bias, act, out_fp32, 0, Z, ldz, deriv, 1.0f, 1, block_n, groups,
                        stream);
}

// D[M,N] = epilogue(sum_k A(m,k) B(n,k)) with fp32 operands multiplied as TF32 (`tcgen05.mma.kind::tf32`, fp32
// accumulate): the "float" precision mode for layers that must track an fp32 reference more closely than bf16 does.
// Same conventions as vb_gemm_bf16; A, B (and Z) are float, leading dimensions multiples of 4 elements.
extern "C" int vb_gemm_tf32(const float* A, long long lda, int a_mn_major, const float* B, long long ldb,
                            int b_mn_major, void* D, long long ldd, int M, int N, int K, const float* bias, int act,
                            int out_fp32, int accumulate, const float* Z, long long ldz, int deriv, float alpha,
                            int split_k, int block_n, cudaStream_t stream) {
  VB_REQUIRE(M > 0 && N > 0 && K > 0, "vb_gemm_tf32: empty problem %dx%dx%d", M, N, K);
  VB_REQUIRE(!(accumulate && !out_fp32), "vb_gemm_tf32: accumulate requires fp32 output");
  VB_REQUIRE(split_k >= 0, "vb_gemm_tf32: split_k must be >= 0 (0 = choose automatically)");
  VB_REQUIRE(block_n == 0 || block_n == 64 || block_n == 128 || block_n == 256, "vb_gemm_tf32: block_n must be 0/64/128/256");
  const bool auto_split = (split_k == 0);
  if (auto_split) split_k = 1;
  const bool can_split = accumulate && act == 0 && Z == nullptr;
  VB_REQUIRE(split_k == 1 || can_split, "vb_gemm_tf32: split-K needs accumulate=1 and a linear epilogue");
  constexpr int KE = 32;   // fp32 elements per 128-byte k-block row
  GemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = M; p.N = N; p.K = K;
  p.D = D; p.ldd = ldd; p.bias = bias; p.act = act; p.out_fp32 = out_fp32; p.accumulate = accumulate;
  p.Z = reinterpret_cast<const __nv_bfloat16*>(Z); p.ldz = ldz; p.deriv = deriv; p.z_fp32 = 1; p.alpha = alpha;
  p.k_blocks = ceil_div(K, KE);
  int BN = block_n, pair = 0;
  const int split_req = split_k;
  for (int attempt = 0; attempt < 2; ++attempt) {   // 192-wide tiles are not instantiated for tf32: fall back to 128
    BN = attempt == 0 ? block_n : 128;
    split_k = split_req;
    choose_config(M, N, p.k_blocks, can_split, auto_split, b_mn_major != 0, &BN, &split_k, &pair);
    if (BN != 192) break;
  }
  p.m_blocks = ceil_div(M, pair ? 2 * BLOCK_M : BLOCK_M);
  p.n_blocks = ceil_div(N, BN);
  p.split_k = split_k;
  const int bn_load = pair ? BN / 2 : BN;
  CUtensorMap ta, tb;
  int rc;
  if (!a_mn_major) rc = make_tmap_2d(&ta, A, K, M, lda, KE, BLOCK_M, true);
  else rc = make_tmap_2d(&ta, A, M, K, lda, 32, KE, true, true);
  if (rc) return rc;
  if (!b_mn_major) rc = make_tmap_2d(&tb, B, K, N, ldb, KE, bn_load, true);
  else rc = make_tmap_2d(&tb, B, N, K, ldb, 32, KE, true, true);
  if (rc) return rc;
  return finish_and_launch(p, ta, tb, a_mn_major != 0, b_mn_major != 0, BN, pair, stream, -1, true);
}
