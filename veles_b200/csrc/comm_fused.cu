// veles_b200 — C1: per-layer gradient all-reduce FUSED with the Znicz gradient-descent update
// and the updated-weight broadcast, in ONE kernel over NVLink 5 / NVSwitch.
//
// It replaces the reference's whole master/slave exchange for data-parallel training:
//   slave: Array.map_read (D2H) -> pickle -> ZeroMQ/shm "update"      veles/client.py:138-160, 333-342
//   master: apply_data_from_slave under a lock, then pickle the new weights into the next "job"
//                                                                      veles/server.py:141-169, 401-430
//   slave: unpickle -> Array.map_write/unmap (H2D)                     veles/workflow.py:513-529
// and the conventional NCCL all-reduce + separate optimizer kernel (kept only as the baseline).
//
// Every rank owns 1/world of each parameter segment.  For its slice the kernel
//   1. sums the gradient over all ranks:  multimem.ld_reduce.add.v4.f32 on the NVLS multicast
//      address (the switch reduces in-fabric), or P2P loads from every peer when there is no
//      multicast mapping;
//   2. applies weight decay (L1/L2 mix), momentum and the learning rate in registers, updating
//      the local fp32 master weights and velocity;
//   3. publishes the new fp32 weights and the bf16 shadow (what the next forward GEMM's TMA
//      loads read) to ALL ranks with multimem.st (or P2P stores).
// Synchronisation is device-side only: monotonically increasing epoch words in symmetric
// memory (st.release.sys / ld.acquire.sys), so the kernel is CUDA-graph capturable and needs
// no host round trip.  Gradients are double-buffered by step parity; the kernel zeroes the
// other parity's local buffer, which every peer is known to have finished reading (they could
// not have entered this epoch otherwise).
#include "common.cuh"

#include <cstring>

namespace vb {

constexpr int MAX_WORLD = 8;
constexpr int MAX_SEGS = 16;
constexpr int FUSED_THREADS = 512;   // x <= 128 registers = one SM's register file per CTA
constexpr int SEG_ZERO_OTHER = 1;    // this segment's gradient GEMMs accumulate: clear the other parity's buffer
constexpr int SEG_BCAST_MASTER = 2;  // the forward pass reads this segment's fp32 master (bias; tf32 layers): always
                                     // publish the updated fp32 values to every rank, even with a sharded master

struct FusedSeg {
  long long offset;         // element offset into the flat fp32 slabs
  long long numel;
  long long shadow_offset;  // element offset into the bf16 shadow slab, -1 = no shadow
  int row_len, row_pitch;   // shadow rows: element i -> shadow[(i / row_len) * row_pitch + i % row_len]
  int flags;                // SEG_* bits
  float lr, wd, l12, moment;
};

struct FusedParams {
  int world, rank, nseg, channel;
  FusedSeg seg[MAX_SEGS];
  float* grads_local;        // this step's parity
  float* grads_mc;           // multicast alias of the same slab or null
  float* grads_peer[MAX_WORLD];
  float* grads_other_local;  // other parity: zeroed here (pull); null in push mode
  float* stage_local;        // push mode: region r of a segment holds rank r's contribution to MY slice
  float* stage_peer[MAX_WORLD];
  int push;                  // 1: gradients travel by remote STORES into the owners' staging slabs
  float* weights_local;
  float* weights_mc;
  float* weights_peer[MAX_WORLD];
  float* vel_local;
  __nv_bfloat16* shadow_local;
  __nv_bfloat16* shadow_mc;
  __nv_bfloat16* shadow_peer[MAX_WORLD];
  uint32_t* flags_local;     // [channels][2][MAX_WORLD] epoch words written by peers
  uint32_t* flags_peer[MAX_WORLD];
  uint32_t* epochs;          // local [channels] last completed epoch
  unsigned int* cta_counter; // local [channels]
  int* status;               // local: set non-zero on a synchronisation timeout
  float grad_scale;
  long long timeout_cycles;
  int phase;         // 0: handshake + work in one launch; 1: ready handshake only (1 CTA); 2: work only
  int bcast_master;  // 0: the updated fp32 master slice stays on its owner (gathered on demand at epoch ends)
  int bcast_shadow;  // 0: the bf16 shadow slice stays on its owner; GEMMs fetch it through peer tensor maps
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 mc_ld_reduce_f32x4(const float* p) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p)
               : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f32x4(float* p, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_st_f32x2(float* p, float a, float b) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}

__device__ __forceinline__ float gd_apply(float w, float g, float& v, const FusedSeg& s, float grad_scale) {
  const float sgn = (w > 0.0f) ? 1.0f : ((w < 0.0f) ? -1.0f : 0.0f);
  g = g * grad_scale + s.wd * ((1.0f - s.l12) * w + 0.5f * s.l12 * sgn);
  v = s.moment * v - s.lr * g;
  return w + v;
}

// Wait until flags_local[slot + r] >= epoch for every rank r != me; false on timeout.
__device__ __forceinline__ bool wait_all(const FusedParams& p, int slot, uint32_t epoch) {
  const long long t0 = clock64();
  for (int r = 0; r < p.world; ++r) {
    if (r == p.rank) continue;
    const uint32_t* f = p.flags_local + slot + r;
    while (static_cast<int32_t>(ld_acquire_sys(f) - epoch) < 0) {
      if (clock64() - t0 > p.timeout_cycles) return false;
      __nanosleep(64);
    }
  }
  return true;
}

// Footprint: FEW CTAs (4-8, launched as 2-CTA clusters so that each pair takes exactly one TPC) of 512 threads x <= 128
// registers x 8 independent 16-byte chains, no shared memory.  While a launch of this kernel is pending the GEMMs of
// the backward pass size their persistent grids to the remaining SMs (vb_set_sm_reserve), so the exchange runs
// BESIDE the backward GEMMs instead of taking turns with them: a tcgen05 GEMM CTA owns its SM (227 KB of shared
// memory, 54 K registers), and a full-width launch of either kernel simply waits for the other's CTAs to drain.
// Bytes in flight per SM are bounded by the register file (64 KB of loads = 512 x 8 x 16 B), i.e. ~25 GB/s per CTA
// at the ~2.5 us NVLink round trip of multimem.ld_reduce: enough to hide AlexNet's 244 MB / world behind the
// convolution backward pass.  (Round 1 used 64-96 CTAs of 256 threads per LAYER: 0.4-0.5 ms per step exposed.)
template <bool MULTIMEM>
__device__ __forceinline__ void fused_allreduce_gd_body(const FusedParams& p) {
  const int ready_slot = (p.channel * 2 + 0) * MAX_WORLD;
  const int done_slot = (p.channel * 2 + 1) * MAX_WORLD;
  // identical in every CTA: epochs[] is bumped by the LAST CTA to finish, i.e. after every CTA has read it here
  const uint32_t epoch = p.epochs[p.channel] + 1;
  int my_ok = 1;
  if (p.phase == 2) {
    my_ok = (*reinterpret_cast<volatile int*>(p.status) == 0) ? 1 : 0;   // the handshake launch may have timed out
  } else if (threadIdx.x == 0) {
    if (blockIdx.x == 0) {
      // my gradients for this segment are complete (stream order): tell every peer
      __threadfence_system();
      for (int r = 0; r < p.world; ++r)
        if (r != p.rank) st_release_sys(p.flags_peer[r] + ready_slot + p.rank, epoch);
    }
    my_ok = wait_all(p, ready_slot, epoch) ? 1 : 0;
  }
  const bool ok = __syncthreads_and(my_ok) != 0;   // barrier + predicate broadcast without shared memory
  if (p.phase == 1) {   // handshake-only launch: one small CTA waits for the peers so that the worker CTAs never
    if (!ok && threadIdx.x == 0) atomicExch(p.status, 1 + p.channel);   // hold SMs while ranks are skewed
    return;
  }

  if (ok) {
    const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
    for (int si = 0; si < p.nseg; ++si) {
      const FusedSeg s = p.seg[si];
      // zero the other parity's local gradients of this segment (everybody is done reading them); only segments
      // whose gradient kernels accumulate (split-K reduce-add, column sums) need it - plainly stored ones do not
      if (p.grads_other_local != nullptr && (s.flags & SEG_ZERO_OTHER)) {
        float* z = p.grads_other_local + s.offset;
        const long long n4 = s.numel >> 2;
        for (long long i = tid; i < n4; i += nthreads) reinterpret_cast<float4*>(z)[i] = make_float4(0, 0, 0, 0);
        for (long long i = (n4 << 2) + tid; i < s.numel; i += nthreads) z[i] = 0.0f;
      }
      // my slice of the segment
      long long chunk = (s.numel + p.world - 1) / p.world;
      chunk = (chunk + 3) & ~3LL;
      const long long lo = min(s.numel, chunk * p.rank);
      const long long hi = min(s.numel, lo + chunk);
      const long long n4 = (hi - lo) >> 2;  // lo is a multiple of 4 and offset is 64-element aligned
      const bool flat_shadow = s.shadow_offset >= 0 && s.row_len == s.row_pitch;
      const bool bcast_master = p.bcast_master || (s.flags & SEG_BCAST_MASTER);
      // UNROLL independent 16-byte chains per thread, and the NEXT batch of reduced gradients is requested before
      // the current one is consumed: the NVLink round trip of multimem.ld_reduce is several microseconds, so an SM
      // earns bandwidth only while its loads are in flight (without the prefetch they were in flight a third of
      // the time: 61 GB/s from 8 SMs at 2 ranks).
      constexpr int UNROLL = 4;
      auto load_g = [&](long long b4, float4 (&g)[UNROLL]) {
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const long long i4 = b4 + u * nthreads;
          if (i4 < n4) {
            const long long gi = s.offset + lo + (i4 << 2);
            if (MULTIMEM) g[u] = mc_ld_reduce_f32x4(p.grads_mc + gi);
            else g[u] = *reinterpret_cast<const float4*>(p.grads_local + gi);
          }
        }
        if (!MULTIMEM) {
          for (int r = 0; r < p.world; ++r) {
            if (r == p.rank) continue;
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
              const long long i4 = b4 + u * nthreads;
              if (i4 < n4) {
                const float4 o = *reinterpret_cast<const float4*>(p.grads_peer[r] + s.offset + lo + (i4 << 2));
                g[u].x += o.x; g[u].y += o.y; g[u].z += o.z; g[u].w += o.w;
              }
            }
          }
        }
      };
      float4 g[UNROLL];
      long long b4 = tid;
      if (b4 < n4) load_g(b4, g);
      while (b4 < n4) {
        const long long b4_next = b4 + nthreads * UNROLL;
        float4 gn[UNROLL];
        if (b4_next < n4) load_g(b4_next, gn);
        float4 w[UNROLL], v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const long long i4 = b4 + u * nthreads;
          if (i4 < n4) {
            const long long gi = s.offset + lo + (i4 << 2);
            w[u] = *reinterpret_cast<const float4*>(p.weights_local + gi);
            v[u] = *reinterpret_cast<const float4*>(p.vel_local + gi);
          }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const long long i4 = b4 + u * nthreads;
          if (i4 >= n4) continue;
          const long long i = lo + (i4 << 2);
          const long long gi = s.offset + i;
          float4 wv4 = w[u], vv = v[u];
          const float4 gg = g[u];
          wv4.x = gd_apply(wv4.x, gg.x, vv.x, s, p.grad_scale);
          wv4.y = gd_apply(wv4.y, gg.y, vv.y, s, p.grad_scale);
          wv4.z = gd_apply(wv4.z, gg.z, vv.z, s, p.grad_scale);
          wv4.w = gd_apply(wv4.w, gg.w, vv.w, s, p.grad_scale);
          *reinterpret_cast<float4*>(p.vel_local + gi) = vv;
          if (!bcast_master) {
            *reinterpret_cast<float4*>(p.weights_local + gi) = wv4;
          } else if (MULTIMEM) {
            mc_st_f32x4(p.weights_mc + gi, wv4);
          } else {
            for (int r = 0; r < p.world; ++r) *reinterpret_cast<float4*>(p.weights_peer[r] + gi) = wv4;
          }
          if (s.shadow_offset >= 0) {
            const uint32_t b0 = pack_bf16x2(wv4.x, wv4.y), b1 = pack_bf16x2(wv4.z, wv4.w);
            if (flat_shadow) {
              const long long so = s.shadow_offset + i;
              if (!p.bcast_shadow) {
                *reinterpret_cast<uint2*>(p.shadow_local + so) = make_uint2(b0, b1);
              } else if (MULTIMEM) {
                mc_st_f32x2(reinterpret_cast<float*>(p.shadow_mc + so), __uint_as_float(b0), __uint_as_float(b1));
              } else {
                for (int r = 0; r < p.world; ++r) *reinterpret_cast<uint2*>(p.shadow_peer[r] + so) = make_uint2(b0, b1);
              }
            } else {
              const float wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
#pragma unroll
              for (int q = 0; q < 4; ++q) {
                const long long e = i + q;
                const long long so = s.shadow_offset + (e / s.row_len) * s.row_pitch + (e % s.row_len);
                const __nv_bfloat16 bq = __float2bfloat16(wv[q]);
                for (int r = 0; r < p.world; ++r) p.shadow_peer[r][so] = bq;  // P2P scalar stores (tiny layers only)
              }
            }
          }
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) g[u] = gn[u];
        b4 = b4_next;
      }
      // scalar tail of my slice (numel not a multiple of 4)
      for (long long i = lo + (n4 << 2) + tid; i < hi; i += nthreads) {
        const long long gi = s.offset + i;
        float g = p.grads_local[gi];
        for (int r = 0; r < p.world; ++r)
          if (r != p.rank) g += p.grads_peer[r][gi];
        float v = p.vel_local[gi];
        const float w = gd_apply(p.weights_local[gi], g, v, s, p.grad_scale);
        p.vel_local[gi] = v;
        if (!bcast_master) p.weights_local[gi] = w;
        else
          for (int r = 0; r < p.world; ++r) p.weights_peer[r][gi] = w;
        if (s.shadow_offset >= 0) {
          const long long so = s.shadow_offset + (i / s.row_len) * s.row_pitch + (i % s.row_len);
          for (int r = 0; r < p.world; ++r) p.shadow_peer[r][so] = __float2bfloat16(w);
        }
      }
    }
  }

  // ---- completion: the last CTA tells the peers that all my stores are out and waits for theirs
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(p.cta_counter + p.channel, 1u);
    if (prev == gridDim.x - 1) {
      bool fine = ok;
      __threadfence_system();
      for (int r = 0; r < p.world; ++r)
        if (r != p.rank) st_release_sys(p.flags_peer[r] + done_slot + p.rank, epoch);
      if (fine) fine = wait_all(p, done_slot, epoch);
      if (!fine) atomicExch(p.status, 1 + p.channel);
      p.cta_counter[p.channel] = 0;
      p.epochs[p.channel] = epoch;
      __threadfence();
    }
  }
}

// ----------------------------------------------------------------------------------------------------------------
// PUSH form of the exchange (the default): no load ever crosses NVLink.
//   phase 1  every rank STORES, for each peer q, the part of its gradient that q owns into q's staging slab
//            (region `rank` of the segment there); stores are posted, so a few SMs keep the link busy, where the
//            pull form (multimem.ld_reduce / peer loads) stalls ~3 us per 64 KB of loads an SM can hold in registers
//            (measured at 2 ranks, 8 CTAs: 61 GB/s, 1.24 ms for AlexNet fc6 - longer than the backward pass it
//            was meant to hide behind);
//   handshake: the last CTA to finish phase 1 publishes "my pushes are out" to everybody (and to itself);
//   phase 2  once all contributions have arrived, the owner sums its own slice (local gradient + W - 1 staging
//            regions, all local HBM reads), applies the Znicz update to the fp32 master and momentum, and publishes
//            the bf16 operand copy (and flagged fp32 segments) with multimem.st / peer stores;
//   phase 3  clears the local gradient of segments whose producers accumulate, completion handshake.
// Gradients need no second parity buffer here: nobody reads another rank's gradient slab.
// Staging layout: a segment of numel n occupies the same [offset, offset + n) range in the staging slab; its W
// chunks of `chunk` = roundup4(ceil(n / W)) elements hold, on rank q, the contributions of ranks 0..W-1 to q's slice
// (W * chunk never exceeds n rounded up to the 64-element slab alignment).
// ----------------------------------------------------------------------------------------------------------------
template <bool MULTIMEM>
__device__ __forceinline__ void fused_push_gd_body(const FusedParams& p) {
  const int ready_slot = (p.channel * 2 + 0) * MAX_WORLD;
  const int done_slot = (p.channel * 2 + 1) * MAX_WORLD;
  const uint32_t epoch = p.epochs[p.channel] + 1;
  const long long tid = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  const long long nthreads = static_cast<long long>(gridDim.x) * blockDim.x;
  constexpr int UNROLL = 8;

  // ---- phase 1: push my contributions to the owners ----
  for (int si = 0; si < p.nseg; ++si) {
    const FusedSeg s = p.seg[si];
    long long chunk = (s.numel + p.world - 1) / p.world;
    chunk = (chunk + 3) & ~3LL;
    for (int q = 0; q < p.world; ++q) {
      if (q == p.rank) continue;
      const long long lo = min(s.numel, chunk * q), hi = min(s.numel, lo + chunk);
      const long long n4 = (hi - lo + 3) >> 2;      // whole 16-byte vectors (the slab is padded to 64 elements)
      const float4* src = reinterpret_cast<const float4*>(p.grads_local + s.offset + lo);
      float4* dst = reinterpret_cast<float4*>(p.stage_peer[q] + s.offset + chunk * p.rank);
      for (long long b4 = tid; b4 < n4; b4 += nthreads * UNROLL) {
        float4 g[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const long long i4 = b4 + u * nthreads;
          if (i4 < n4) g[u] = src[i4];
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
          const long long i4 = b4 + u * nthreads;
          if (i4 < n4) dst[i4] = g[u];
        }
      }
    }
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(p.cta_counter + p.channel, 1u);
    if (prev == gridDim.x - 1) {
      __threadfence_system();
      for (int r = 0; r < p.world; ++r) st_release_sys(p.flags_peer[r] + ready_slot + p.rank, epoch);   // me included
      p.cta_counter[p.channel] = 0;
    }
  }
  // ---- wait for every rank's pushes (and for all CTAs of this rank: its own flag) ----
  int my_ok = 1;
  if (threadIdx.x == 0) {
    const long long t0 = clock64();
    for (int r = 0; r < p.world && my_ok; ++r) {
      const uint32_t* f = p.flags_local + ready_slot + r;
      while (static_cast<int32_t>(ld_acquire_sys(f) - epoch) < 0) {
        if (clock64() - t0 > p.timeout_cycles) {
          my_ok = 0;
          break;
        }
        __nanosleep(64);
      }
    }
  }
  const bool ok = __syncthreads_and(my_ok) != 0;

  // ---- phase 2: reduce my slice from local memory, update, publish ----
  if (ok) {
    for (int si = 0; si < p.nseg; ++si) {
      const FusedSeg s = p.seg[si];
      long long chunk = (s.numel + p.world - 1) / p.world;
      chunk = (chunk + 3) & ~3LL;
      const long long lo = min(s.numel, chunk * p.rank);
      const long long hi = min(s.numel, lo + chunk);
      const long long n4 = (hi - lo) >> 2;
      const bool flat_shadow = s.shadow_offset >= 0 && s.row_len == s.row_pitch;
      const bool bcast_master = p.bcast_master || (s.flags & SEG_BCAST_MASTER);
      const float* stage = p.stage_local + s.offset;
      for (long long i4 = tid; i4 < n4; i4 += nthreads) {
        const long long i = lo + (i4 << 2);
        const long long gi = s.offset + i;
        float4 g = *reinterpret_cast<const float4*>(p.grads_local + gi);
        // (an accumulating segment restarts from zero: the thread that read an element clears it)
        if (s.flags & SEG_ZERO_OTHER) *reinterpret_cast<float4*>(p.grads_local + gi) = make_float4(0, 0, 0, 0);
        for (int r = 0; r < p.world; ++r) {
          if (r == p.rank) continue;
          const float4 o = *reinterpret_cast<const float4*>(stage + chunk * r + (i4 << 2));
          g.x += o.x; g.y += o.y; g.z += o.z; g.w += o.w;
        }
        float4 wv4 = *reinterpret_cast<const float4*>(p.weights_local + gi);
        float4 vv = *reinterpret_cast<const float4*>(p.vel_local + gi);
        wv4.x = gd_apply(wv4.x, g.x, vv.x, s, p.grad_scale);
        wv4.y = gd_apply(wv4.y, g.y, vv.y, s, p.grad_scale);
        wv4.z = gd_apply(wv4.z, g.z, vv.z, s, p.grad_scale);
        wv4.w = gd_apply(wv4.w, g.w, vv.w, s, p.grad_scale);
        *reinterpret_cast<float4*>(p.vel_local + gi) = vv;
        if (!bcast_master) {
          *reinterpret_cast<float4*>(p.weights_local + gi) = wv4;
        } else if (MULTIMEM) {
          mc_st_f32x4(p.weights_mc + gi, wv4);
        } else {
          for (int r = 0; r < p.world; ++r) *reinterpret_cast<float4*>(p.weights_peer[r] + gi) = wv4;
        }
        if (s.shadow_offset >= 0) {
          const uint32_t b0 = pack_bf16x2(wv4.x, wv4.y), b1 = pack_bf16x2(wv4.z, wv4.w);
          if (flat_shadow) {
            const long long so = s.shadow_offset + i;
            if (!p.bcast_shadow) {
              *reinterpret_cast<uint2*>(p.shadow_local + so) = make_uint2(b0, b1);
            } else if (MULTIMEM) {
              mc_st_f32x2(reinterpret_cast<float*>(p.shadow_mc + so), __uint_as_float(b0), __uint_as_float(b1));
            } else {
              for (int r = 0; r < p.world; ++r) *reinterpret_cast<uint2*>(p.shadow_peer[r] + so) = make_uint2(b0, b1);
            }
          } else {
            const float wv[4] = {wv4.x, wv4.y, wv4.z, wv4.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              const long long e = i + q;
              const long long so = s.shadow_offset + (e / s.row_len) * s.row_pitch + (e % s.row_len);
              const __nv_bfloat16 bq = __float2bfloat16(wv[q]);
              for (int r = 0; r < p.world; ++r) p.shadow_peer[r][so] = bq;
            }
          }
        }
      }
      // scalar tail of my slice (numel not a multiple of 4)
      for (long long i = lo + (n4 << 2) + tid; i < hi; i += nthreads) {
        const long long gi = s.offset + i;
        float g = p.grads_local[gi];
        if (s.flags & SEG_ZERO_OTHER) p.grads_local[gi] = 0.0f;
        for (int r = 0; r < p.world; ++r)
          if (r != p.rank) g += stage[chunk * r + (i - lo)];
        float v = p.vel_local[gi];
        const float w = gd_apply(p.weights_local[gi], g, v, s, p.grad_scale);
        p.vel_local[gi] = v;
        if (!bcast_master) p.weights_local[gi] = w;
        else
          for (int r = 0; r < p.world; ++r) p.weights_peer[r][gi] = w;
        if (s.shadow_offset >= 0) {
          const long long so = s.shadow_offset + (i / s.row_len) * s.row_pitch + (i % s.row_len);
          for (int r = 0; r < p.world; ++r) p.shadow_peer[r][so] = __float2bfloat16(w);
        }
      }
    }
    // ---- phase 3: the parts of accumulating segments that OTHER ranks own restart from zero too ----
    // (their last readers were this rank's phase-1 pushes, complete on every CTA: the rank's own "ready" flag was
    // awaited above; the own slice was cleared element by element in phase 2)
    for (int si = 0; si < p.nseg; ++si) {
      const FusedSeg s = p.seg[si];
      if (!(s.flags & SEG_ZERO_OTHER)) continue;
      long long chunk = (s.numel + p.world - 1) / p.world;
      chunk = (chunk + 3) & ~3LL;
      const long long lo = min(s.numel, chunk * p.rank);
      const long long hi = min(s.numel, lo + chunk);
      float* z = p.grads_local + s.offset;
      for (long long i = tid; i < s.numel; i += nthreads)
        if (i < lo || i >= hi) z[i] = 0.0f;
    }
  }

  // ---- completion: the last CTA tells the peers that all my stores are out and waits for theirs ----
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int prev = atomicAdd(p.cta_counter + 64 + p.channel, 1u);
    if (prev == gridDim.x - 1) {
      bool fine = ok;
      __threadfence_system();
      for (int r = 0; r < p.world; ++r)
        if (r != p.rank) st_release_sys(p.flags_peer[r] + done_slot + p.rank, epoch);
      if (fine) fine = wait_all(p, done_slot, epoch);
      if (!fine) atomicExch(p.status, 1 + p.channel);
      p.cta_counter[64 + p.channel] = 0;
      p.epochs[p.channel] = epoch;
      __threadfence();
    }
  }
}

template <bool MULTIMEM>
__global__ void __launch_bounds__(FUSED_THREADS, 1) fused_allreduce_gd_kernel(const FusedParams p) {
  if (p.push) fused_push_gd_body<MULTIMEM>(p);
  else fused_allreduce_gd_body<MULTIMEM>(p);
}

// All ranks of a group on ONE GPU, as one cooperative launch: blockIdx.y is the rank, every rank's CTAs run the very
// same body against that rank's parameter block (P2P path; the "peers" are other buffers of this GPU).  The ranks
// wait for each other's flags, so all CTAs must be resident at once - hence the cooperative launch.  This is how a
// single-GPU box exercises the flag protocol, the slice arithmetic and the per-segment options.
__global__ void __launch_bounds__(FUSED_THREADS, 1) fused_allreduce_gd_emulated_kernel(const FusedParams* ps) {
  if (ps[blockIdx.y].push) fused_push_gd_body<false>(ps[blockIdx.y]);
  else fused_allreduce_gd_body<false>(ps[blockIdx.y]);
}

}  // namespace vb

using namespace vb;

// Host ABI: all arrays are passed flat so that ctypes callers need no struct mirror.
//   segs: nseg x {offset, numel, shadow_offset, row_len, row_pitch, flags} as long long[6]
//         flags bit 0: zero the other parity's gradients of the segment; bit 1: always broadcast the fp32 master
//   hyper: nseg x {lr, wd, l1_vs_l2, moment} as float[4]
//   *_peer: world device pointers each
//   use_multimem: bit 0 = NVLS multimem path; bit 1 = keep the fp32 master slice local (no broadcast);
//                 bit 2 = keep the bf16 shadow slice local (flat-shadow segments only; peer-TMA weight delivery)
//                 bit 3 = wait for the peers' gradients in a separate 1-CTA launch ahead of the worker CTAs
//                 bit 4 = PUSH form: gradients are stored into the owners' staging slabs (stage_local / stage_peer:
//                         slabs laid out like the gradient slab); cta_counter then needs 128 words
extern "C" int vb_fused_allreduce_gd(int world, int rank, int channel, int nseg, const long long* segs,
                                     const float* hyper, float* grads_local, float* grads_mc, float** grads_peer,
                                     float* grads_other_local, float* weights_local, float* weights_mc,
                                     float** weights_peer, float* vel_local, void* shadow_local, void* shadow_mc,
                                     void** shadow_peer, unsigned int* flags_local, unsigned int** flags_peer,
                                     unsigned int* epochs, unsigned int* cta_counter, int* status, float grad_scale,
                                     int n_ctas, int use_multimem, double timeout_s, float* stage_local,
                                     float** stage_peer, cudaStream_t stream) {
  const int local_master = (use_multimem >> 1) & 1, local_shadow = (use_multimem >> 2) & 1;
  const int split_handshake = (use_multimem >> 3) & 1;
  const int push = (use_multimem >> 4) & 1;
  use_multimem &= 1;
  VB_REQUIRE(!push || (stage_local != nullptr && stage_peer != nullptr), "vb_fused_allreduce_gd: push mode needs the staging slabs");
  VB_REQUIRE(!push || !split_handshake, "vb_fused_allreduce_gd: push mode has no separate handshake launch");
  VB_REQUIRE(world >= 2 && world <= MAX_WORLD, "vb_fused_allreduce_gd: world size %d not in [2,%d]", world, MAX_WORLD);
  VB_REQUIRE(nseg >= 1 && nseg <= MAX_SEGS, "vb_fused_allreduce_gd: %d segments (max %d per launch)", nseg, MAX_SEGS);
  VB_REQUIRE(channel >= 0 && channel < 64, "vb_fused_allreduce_gd: channel %d out of range", channel);
  VB_REQUIRE(!use_multimem || (grads_mc && weights_mc), "vb_fused_allreduce_gd: multimem path needs multicast pointers");
  FusedParams p;
  p.world = world; p.rank = rank; p.nseg = nseg; p.channel = channel;
  for (int i = 0; i < nseg; ++i) {
    p.seg[i].offset = segs[i * 6 + 0];
    p.seg[i].numel = segs[i * 6 + 1];
    p.seg[i].shadow_offset = segs[i * 6 + 2];
    p.seg[i].row_len = static_cast<int>(segs[i * 6 + 3]);
    p.seg[i].row_pitch = static_cast<int>(segs[i * 6 + 4]);
    p.seg[i].flags = static_cast<int>(segs[i * 6 + 5]);
    p.seg[i].lr = hyper[i * 4 + 0];
    p.seg[i].wd = hyper[i * 4 + 1];
    p.seg[i].l12 = hyper[i * 4 + 2];
    p.seg[i].moment = hyper[i * 4 + 3];
    VB_REQUIRE((p.seg[i].offset & 3) == 0, "vb_fused_allreduce_gd: segment offset must be a multiple of 4 elements");
    VB_REQUIRE(p.seg[i].shadow_offset < 0 || p.seg[i].row_len > 0, "vb_fused_allreduce_gd: bad shadow geometry");
    if (use_multimem && p.seg[i].shadow_offset >= 0 && p.seg[i].row_len == p.seg[i].row_pitch)
      VB_REQUIRE(shadow_mc != nullptr, "vb_fused_allreduce_gd: shadow multicast pointer missing");
  }
  p.grads_local = grads_local; p.grads_mc = grads_mc; p.grads_other_local = grads_other_local;
  p.weights_local = weights_local; p.weights_mc = weights_mc; p.vel_local = vel_local;
  p.shadow_local = static_cast<__nv_bfloat16*>(shadow_local);
  p.shadow_mc = static_cast<__nv_bfloat16*>(shadow_mc);
  for (int r = 0; r < MAX_WORLD; ++r) {
    p.grads_peer[r] = r < world ? grads_peer[r] : nullptr;
    p.weights_peer[r] = r < world ? weights_peer[r] : nullptr;
    p.shadow_peer[r] = r < world && shadow_peer ? static_cast<__nv_bfloat16*>(shadow_peer[r]) : nullptr;
    p.flags_peer[r] = r < world ? flags_peer[r] : nullptr;
  }
  p.flags_local = flags_local; p.epochs = epochs; p.cta_counter = cta_counter; p.status = status;
  p.grad_scale = grad_scale;
  p.push = push;
  p.stage_local = stage_local;
  for (int r = 0; r < MAX_WORLD; ++r) p.stage_peer[r] = (push && r < world) ? stage_peer[r] : nullptr;
  p.bcast_master = !local_master;
  p.bcast_shadow = !local_shadow;
  p.timeout_cycles = static_cast<long long>(timeout_s * 1.9e9);
  if (n_ctas <= 0) n_ctas = 4;
  n_ctas = (n_ctas + 1) & ~1;   // whole 2-CTA clusters: a pair occupies one TPC, so GEMM CTA pairs lose no extra SMs
  if (split_handshake) {
    p.phase = 1;
    fused_allreduce_gd_kernel<false><<<1, 32, 0, stream>>>(p);
    VB_CHECK_CUDA(cudaGetLastError());
    p.phase = 2;
  } else {
    p.phase = 0;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(n_ctas);
  cfg.blockDim = dim3(FUSED_THREADS);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (use_multimem) VB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, fused_allreduce_gd_kernel<true>, p));
  else VB_CHECK_CUDA(cudaLaunchKernelEx(&cfg, fused_allreduce_gd_kernel<false>, p));
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

// Test / bring-up entry point: `world` ranks emulated on the current GPU (see fused_allreduce_gd_emulated_kernel).
// Per-rank arrays hold `world` device pointers each (rank r's gradient slab of this step's parity, the other
// parity's, weights, momentum, bf16 shadow, flag words [64][2][8], epoch words [64], CTA counters [64], status word).
// `params_dev` is caller-provided device scratch of at least world * vb_fused_params_bytes() bytes.  options bit 4 with
// `stage` (one staging slab per rank) selects the push form; cta_counter then needs 128 words per rank.
extern "C" long long vb_fused_params_bytes() { return static_cast<long long>(sizeof(FusedParams)); }

extern "C" int vb_fused_allreduce_gd_emulated(int world, int channel, int nseg, const long long* segs,
                                              const float* hyper, float** grads, float** grads_other, float** weights,
                                              float** vel, void** shadow, unsigned int** flags, unsigned int** epochs,
                                              unsigned int** cta_counter, int** status, float grad_scale, int n_ctas,
                                              int options, double timeout_s, void* params_dev, float** stage,
                                              cudaStream_t stream) {
  VB_REQUIRE(world >= 2 && world <= MAX_WORLD, "vb_fused_allreduce_gd_emulated: world size %d not in [2,%d]", world, MAX_WORLD);
  VB_REQUIRE(nseg >= 1 && nseg <= MAX_SEGS, "vb_fused_allreduce_gd_emulated: %d segments (max %d)", nseg, MAX_SEGS);
  VB_REQUIRE(channel >= 0 && channel < 64, "vb_fused_allreduce_gd_emulated: channel %d out of range", channel);
  VB_REQUIRE(params_dev != nullptr, "vb_fused_allreduce_gd_emulated: parameter scratch is required");
  if (n_ctas <= 0) n_ctas = 2;
  static FusedParams host[MAX_WORLD];
  for (int r = 0; r < world; ++r) {
    FusedParams& p = host[r];
    memset(&p, 0, sizeof(p));
    p.world = world; p.rank = r; p.nseg = nseg; p.channel = channel;
    for (int i = 0; i < nseg; ++i) {
      p.seg[i].offset = segs[i * 6 + 0];
      p.seg[i].numel = segs[i * 6 + 1];
      p.seg[i].shadow_offset = segs[i * 6 + 2];
      p.seg[i].row_len = static_cast<int>(segs[i * 6 + 3]);
      p.seg[i].row_pitch = static_cast<int>(segs[i * 6 + 4]);
      p.seg[i].flags = static_cast<int>(segs[i * 6 + 5]);
      p.seg[i].lr = hyper[i * 4 + 0];
      p.seg[i].wd = hyper[i * 4 + 1];
      p.seg[i].l12 = hyper[i * 4 + 2];
      p.seg[i].moment = hyper[i * 4 + 3];
      VB_REQUIRE((p.seg[i].offset & 3) == 0, "vb_fused_allreduce_gd_emulated: segment offset must be a multiple of 4");
    }
    p.grads_local = grads[r]; p.grads_other_local = grads_other ? grads_other[r] : nullptr;
    p.weights_local = weights[r]; p.vel_local = vel[r];
    p.shadow_local = static_cast<__nv_bfloat16*>(shadow ? shadow[r] : nullptr);
    for (int q = 0; q < world; ++q) {
      p.grads_peer[q] = grads[q];
      p.weights_peer[q] = weights[q];
      p.shadow_peer[q] = static_cast<__nv_bfloat16*>(shadow ? shadow[q] : nullptr);
      p.flags_peer[q] = flags[q];
    }
    p.flags_local = flags[r]; p.epochs = epochs[r]; p.cta_counter = cta_counter[r]; p.status = status[r];
    p.grad_scale = grad_scale;
    p.bcast_master = !((options >> 1) & 1);
    p.bcast_shadow = !((options >> 2) & 1);
    p.push = ((options >> 4) & 1) && stage != nullptr;
    if (p.push) {
      p.stage_local = stage[r];
      for (int q = 0; q < world; ++q) p.stage_peer[q] = stage[q];
    }
    p.timeout_cycles = static_cast<long long>(timeout_s * 1.9e9);
    p.phase = 0;
  }
  VB_CHECK_CUDA(cudaMemcpyAsync(params_dev, host, sizeof(FusedParams) * world, cudaMemcpyHostToDevice, stream));
  VB_CHECK_CUDA(cudaStreamSynchronize(stream));   // `host` is reused by the next call
  const FusedParams* ps = static_cast<const FusedParams*>(params_dev);
  void* args[1] = {&ps};
  VB_CHECK_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(fused_allreduce_gd_emulated_kernel),
                                            dim3(n_ctas, world), dim3(FUSED_THREADS), args, 0, stream));
  return 0;
}
