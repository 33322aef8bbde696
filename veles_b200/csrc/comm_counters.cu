// veles_b200 - device-side exchange of the small things a data-parallel group shares at class / epoch ends:
//   * the evaluator's counters (n_err, evaluated, loss sum, confusion matrix, max / min): one tiny all-reduce over
//     peer memory with the same epoch-flag handshake as the gradient exchange - no NCCL collective, no host barrier;
//   * the owners' slices of the sharded fp32 master weights / momentum: one gather kernel that reads every peer's
//     slice into the local slab.
// The reference's master collected these from its slaves' pickled "update" messages (veles/znicz/decision.py
// apply_data_from_slave; veles/server.py:401-430); here every rank ends up with the sums and a kernel with n = 0 is
// the group's device-side barrier.
//
// Protocol (one CTA per rank; the words live in symmetric memory, identical offsets on every rank):
//   e = epoch + 1; slot parity = e & 1
//   1. copy the local values into my slot[parity]; fence; store e into flag[my rank] on EVERY peer (st.release.sys)
//   2. wait until flag[r] >= e for every rank r (ld.acquire.sys)
//   3. out[i] = op over r = 0..world-1 (fixed order: every rank computes the same bits) of r's slot[parity][i]
// A rank rewrites slot[parity] only at e + 2, which it cannot enter before every peer has signalled e + 1, i.e. has
// finished reading e: two parities are enough.
#include "common.cuh"

#include <cstring>

namespace vb {

constexpr int CNT_MAX_WORLD = 8;
constexpr int CNT_THREADS = 256;
constexpr int CNT_FLAG_WORDS = 64;   // 256 bytes ahead of the slots

struct CounterParams {
  int world, rank, n, op;            // op 0: sum, 1: max
  long long slot_elems;              // doubles per parity
  const double* in;
  double* out;
  unsigned char* base_local;         // [CNT_FLAG_WORDS u32][2][slot_elems double]
  unsigned char* base_peer[CNT_MAX_WORLD];
  uint32_t* epoch;                   // local: last completed exchange
  int* status;
  long long timeout_cycles;
};

__device__ __forceinline__ void cnt_st_release(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t cnt_ld_acquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ double cnt_ld(const double* p) {
  double v;
  asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
  return v;
}

__device__ __forceinline__ void counter_allreduce_body(const CounterParams& p) {
  __shared__ int s_ok;
  const uint32_t e = *p.epoch + 1;
  double* my_slot = reinterpret_cast<double*>(p.base_local + CNT_FLAG_WORDS * 4) + (e & 1) * p.slot_elems;
  for (int i = threadIdx.x; i < p.n; i += CNT_THREADS) my_slot[i] = p.in[i];
  if (threadIdx.x == 0) s_ok = 1;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x < p.world)
    cnt_st_release(reinterpret_cast<uint32_t*>(p.base_peer[threadIdx.x]) + p.rank, e);
  if (threadIdx.x < p.world) {
    const uint32_t* f = reinterpret_cast<const uint32_t*>(p.base_local) + threadIdx.x;
    const long long t0 = clock64();
    while (static_cast<int32_t>(cnt_ld_acquire(f) - e) < 0) {
      if (clock64() - t0 > p.timeout_cycles) {
        s_ok = 0;
        break;
      }
      __nanosleep(100);
    }
  }
  __syncthreads();
  if (!s_ok) {
    if (threadIdx.x == 0) *p.status = 1000 + 1;   // channel numbers of the gradient exchange are < 64
    return;                                        // epoch not advanced: the caller raises
  }
  for (int i = threadIdx.x; i < p.n; i += CNT_THREADS) {
    double acc = 0.0;
    for (int r = 0; r < p.world; ++r) {
      const double* slot = reinterpret_cast<const double*>(p.base_peer[r] + CNT_FLAG_WORDS * 4) + (e & 1) * p.slot_elems;
      const double v = (r == p.rank) ? my_slot[i] : cnt_ld(slot + i);
      acc = (r == 0) ? v : (p.op == 1 ? fmax(acc, v) : acc + v);
    }
    p.out[i] = acc;
  }
  __syncthreads();
  if (threadIdx.x == 0) *p.epoch = e;
}

__global__ void __launch_bounds__(CNT_THREADS, 1) counter_allreduce_kernel(const CounterParams p) {
  counter_allreduce_body(p);
}

// `world` ranks on ONE GPU: CTA r plays rank r (cooperative launch = all co-resident), for the single-GPU test
__global__ void __launch_bounds__(CNT_THREADS, 1) counter_allreduce_emulated_kernel(const CounterParams* ps) {
  counter_allreduce_body(ps[blockIdx.x]);
}

constexpr int GATHER_MAX_SEGS = 16;
struct GatherParams {
  int world, rank, nseg;
  long long offset[GATHER_MAX_SEGS], numel[GATHER_MAX_SEGS];
  float* local;
  const float* peer[CNT_MAX_WORLD];
};

// local[slice of r] = peer r's [slice of r], for every segment and every r != rank (slice arithmetic of comm_fused.cu)
__global__ void __launch_bounds__(256) gather_slices_kernel(const GatherParams p) {
  for (int s = 0; s < p.nseg; ++s) {
    const long long numel = p.numel[s];
    long long chunk = (numel + p.world - 1) / p.world;
    chunk = (chunk + 3) & ~3LL;
    for (int r = 0; r < p.world; ++r) {
      if (r == p.rank) continue;
      const long long lo = min(numel, chunk * r), hi = min(numel, lo + chunk);
      const float* src = p.peer[r] + p.offset[s];
      float* dst = p.local + p.offset[s];
      const bool vec = (((p.offset[s] + lo) & 3) == 0);
      const long long n4 = vec ? (hi - lo) / 4 : 0;
      for (long long i = blockIdx.x * 256LL + threadIdx.x; i < n4; i += gridDim.x * 256LL) {
        float4 v;
        asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                     : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
                     : "l"(src + lo + 4 * i)
                     : "memory");
        *reinterpret_cast<float4*>(dst + lo + 4 * i) = v;
      }
      for (long long i = lo + 4 * n4 + blockIdx.x * 256LL + threadIdx.x; i < hi; i += gridDim.x * 256LL) {
        float v;
        asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(src + i) : "memory");
        dst[i] = v;
      }
    }
  }
}

static long long cycles_of(double timeout_s) {
  int dev = 0, khz = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
  if (khz <= 0) khz = 1900000;
  return static_cast<long long>(timeout_s * 1e3 * khz);
}

}  // namespace vb

using namespace vb;

// Bytes of symmetric memory the exchange needs for `slot_elems` doubles per parity.
extern "C" long long vb_counter_region_bytes(long long slot_elems) {
  return CNT_FLAG_WORDS * 4 + 2 * slot_elems * 8;
}

// All-reduce n doubles (op 0 sum / 1 max) over the group; n = 0 is a device-side barrier.  `base_peer[r]` is rank r's
// mapping of the region (base_peer[rank] == base_local), zero-initialised once before the first call.
extern "C" int vb_counter_allreduce(int world, int rank, int n, int op, long long slot_elems, const double* in,
                                    double* out, void* base_local, void** base_peer, uint32_t* epoch, int* status,
                                    double timeout_s, cudaStream_t stream) {
  VB_REQUIRE(world >= 1 && world <= CNT_MAX_WORLD && rank >= 0 && rank < world, "counter all-reduce: bad world / rank");
  VB_REQUIRE(n >= 0 && n <= slot_elems, "counter all-reduce: more values than the slot holds");
  CounterParams p;
  memset(&p, 0, sizeof(p));
  p.world = world; p.rank = rank; p.n = n; p.op = op; p.slot_elems = slot_elems;
  p.in = in; p.out = out;
  p.base_local = static_cast<unsigned char*>(base_local);
  for (int r = 0; r < world; ++r) p.base_peer[r] = static_cast<unsigned char*>(base_peer[r]);
  p.epoch = epoch; p.status = status;
  p.timeout_cycles = cycles_of(timeout_s);
  counter_allreduce_kernel<<<1, CNT_THREADS, 0, stream>>>(p);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" long long vb_counter_params_bytes() { return static_cast<long long>(sizeof(CounterParams)); }

// The same exchange with all ranks on this GPU (one cooperative launch); params_dev: world * vb_counter_params_bytes().
extern "C" int vb_counter_allreduce_emulated(int world, int n, int op, long long slot_elems, const double** ins,
                                             double** outs, void** bases, uint32_t** epochs, int** status,
                                             double timeout_s, void* params_dev, cudaStream_t stream) {
  VB_REQUIRE(world >= 1 && world <= CNT_MAX_WORLD, "counter all-reduce: bad world");
  VB_REQUIRE(n >= 0 && n <= slot_elems, "counter all-reduce: more values than the slot holds");
  CounterParams ps[CNT_MAX_WORLD];
  memset(ps, 0, sizeof(ps));
  for (int r = 0; r < world; ++r) {
    CounterParams& p = ps[r];
    p.world = world; p.rank = r; p.n = n; p.op = op; p.slot_elems = slot_elems;
    p.in = ins[r]; p.out = outs[r];
    p.base_local = static_cast<unsigned char*>(bases[r]);
    for (int q = 0; q < world; ++q) p.base_peer[q] = static_cast<unsigned char*>(bases[q]);
    p.epoch = epochs[r]; p.status = status[r];
    p.timeout_cycles = cycles_of(timeout_s);
  }
  VB_CHECK_CUDA(cudaMemcpyAsync(params_dev, ps, sizeof(CounterParams) * world, cudaMemcpyHostToDevice, stream));
  VB_CHECK_CUDA(cudaStreamSynchronize(stream));   // ps lives on this stack frame
  const CounterParams* dp = static_cast<const CounterParams*>(params_dev);
  void* args[] = {&dp};
  VB_CHECK_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(counter_allreduce_emulated_kernel), dim3(world),
                                            dim3(CNT_THREADS), args, 0, stream));
  return 0;
}

// Reassemble sharded state: for every segment copy each peer's slice out of its slab (segs: offset, numel pairs in
// fp32 elements; the caller orders this after a vb_counter_allreduce barrier).
extern "C" int vb_gather_slices(int world, int rank, int nseg, const long long* segs, float* local, void** peers,
                                int n_ctas, cudaStream_t stream) {
  VB_REQUIRE(world >= 1 && world <= CNT_MAX_WORLD && nseg >= 0 && nseg <= GATHER_MAX_SEGS, "gather: bad world / nseg");
  if (nseg == 0 || world == 1) return 0;
  GatherParams p;
  memset(&p, 0, sizeof(p));
  p.world = world; p.rank = rank; p.nseg = nseg; p.local = local;
  for (int s = 0; s < nseg; ++s) {
    p.offset[s] = segs[2 * s];
    p.numel[s] = segs[2 * s + 1];
  }
  for (int r = 0; r < world; ++r) p.peer[r] = static_cast<const float*>(peers[r]);
  gather_slices_kernel<<<n_ctas > 0 ? n_ctas : 32, 256, 0, stream>>>(p);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
