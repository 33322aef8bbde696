// veles_b200 — device PRNGs with the reference's exact bit streams.
//
//   xorshift1024*  : one generator per thread, 16 x u64 state kept in registers,
//                    output interleaved with stride n_states so that
//                    out[round*16*S + i*S + state] is draw (round*16+i) of `state`
//                    (behaviour of cuda/random.cu:45-73; oracle veles/prng/uniform.py:129-163)
//   xorshift128+   : one draw per state (behaviour of cuda/random.cu:107-128)
//
// Both are also available as device functions to consumers (dropout, stochastic
// pooling, weight init) so random bytes need not round-trip through HBM.
#include "common.cuh"

namespace vb {

__global__ void xorshift1024star_kernel(uint64_t* __restrict__ states, int n_states, int rounds,
                                        uint64_t* __restrict__ out) {
  vb::pdl_sync();
  const int id = blockIdx.x * blockDim.x + threadIdx.x;
  if (id >= n_states) return;
  uint64_t s[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) s[i] = states[id * 16 + i];
  long long offs = id;
  for (int r = 0; r < rounds; ++r) {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      // position p == i: fully unrolled so the state stays in registers
      uint64_t s0 = s[i];
      uint64_t s1 = s[(i + 1) & 15];
      s1 ^= s1 << 31;
      s1 ^= s1 >> 11;
      s0 ^= s0 >> 30;
      const uint64_t v = s0 ^ s1;
      s[(i + 1) & 15] = v;
      out[offs] = v * 1181783497276652981ULL;
      offs += n_states;
    }
  }
#pragma unroll
  for (int i = 0; i < 16; ++i) states[id * 16 + i] = s[i];
}

__global__ void xorshift128plus_kernel(ulonglong2* __restrict__ states, long long n, uint64_t* __restrict__ out) {
  vb::pdl_sync();
  const long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= n) return;
  ulonglong2 st = states[i];
  uint64_t s1 = st.x;
  const uint64_t s0 = st.y;
  st.x = s0;
  s1 ^= s1 << 23;
  st.y = s1 ^ s0 ^ (s1 >> 17) ^ (s0 >> 26);
  states[i] = st;
  out[i] = st.y + s0;
}

// Uniform / normal fill for weight initialisation straight from xorshift128+ draws:
// kind 0: uniform in [-amp, amp); kind 1: gaussian(0, amp) via Box-Muller, clipped to 5 sigma
// (the reference clips host-side normal fills the same way: veles/prng/random_generator.py:185-198).
__global__ void fill_random_kernel(ulonglong2* __restrict__ states, int n_states, float* __restrict__ out,
                                   long long n, int kind, float amp) {
  vb::pdl_sync();
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_states) return;
  ulonglong2 st = states[t];
  uint64_t a = st.x, b = st.y;
  for (long long base = 2LL * t; base < n; base += 2LL * n_states) {
    uint64_t x = a;
    const uint64_t y = b;
    a = y;
    x ^= x << 23;
    b = x ^ y ^ (x >> 17) ^ (y >> 26);
    const uint64_t r = b + y;
    const float u0 = (static_cast<float>(static_cast<uint32_t>(r) >> 8) + 0.5f) * (1.0f / 16777216.0f);
    const float u1 = (static_cast<float>(static_cast<uint32_t>(r >> 32) >> 8) + 0.5f) * (1.0f / 16777216.0f);
    float v0, v1;
    if (kind == 0) {
      v0 = (2.0f * u0 - 1.0f) * amp;
      v1 = (2.0f * u1 - 1.0f) * amp;
    } else {
      const float rad = sqrtf(-2.0f * __logf(u0));
      float sn, cs;
      __sincosf(6.28318530718f * u1, &sn, &cs);
      v0 = fminf(fmaxf(rad * cs, -5.0f), 5.0f) * amp;
      v1 = fminf(fmaxf(rad * sn, -5.0f), 5.0f) * amp;
    }
    out[base] = v0;
    if (base + 1 < n) out[base + 1] = v1;
  }
  states[t] = make_ulonglong2(a, b);
}

}  // namespace vb

using namespace vb;

extern "C" int vb_xorshift1024star(void* states, int n_states, int rounds, void* out, cudaStream_t s) {
  if (n_states == 0 || rounds == 0) return 0;
  // block <= 32 threads like the reference launch (64 live u64 registers per thread)
  vb::launch_k(xorshift1024star_kernel, dim3(ceil_div(n_states, 32)), dim3(32), 0, s, static_cast<uint64_t*>(states), n_states, rounds,
                                                               static_cast<uint64_t*>(out));
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_xorshift128plus(void* states, long long n, void* out, cudaStream_t s) {
  if (n == 0) return 0;
  vb::launch_k(xorshift128plus_kernel, dim3(static_cast<int>(ceil_div_ll(n, 256))), dim3(256), 0, s, static_cast<ulonglong2*>(states), n,
                                                                              static_cast<uint64_t*>(out));
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_fill_random(void* states, int n_states, float* out, long long n, int kind, float amp,
                              cudaStream_t s) {
  if (n == 0) return 0;
  VB_REQUIRE(n_states > 0, "vb_fill_random: need at least one generator state");
  vb::launch_k(fill_random_kernel, dim3(ceil_div(n_states, 128)), dim3(128), 0, s, static_cast<ulonglong2*>(states), n_states, out, n, kind, amp);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
