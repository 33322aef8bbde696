// veles_b200 — recurrent cell kernels (Elman RNN and LSTM), forward and backward through time.
//
// The reference lists RNN and LSTM units among Znicz's algorithms (docs/source/manualrst_veles_algorithms.rst:
// 108-135, "all necessary units developed", no source in the checkout).  Here one time step is
//     Z_t = [x_t | h_{t-1}] . W^T + b        one tcgen05 GEMM over the concatenated operand (gemm_tcgen05.cu)
//     cell kernel (this file)                 gates, cell state, h_t written straight into the h-slot of step t+1's
//                                             operand and into the unit's output
// and one backward step is the mirrored pair (cell backward -> dZ_t, then dZ_t . W gives dx_t and dh_{t-1} at once);
// the weight gradient is ONE GEMM over all T*B rows after the loop.  The cell kernels are pure element-wise work:
// one thread per (sample, hidden unit), all four gate pre-activations of that unit read by the same thread.
#include "common.cuh"

namespace vb {

__device__ __forceinline__ float rec_load(const void* p, long long i, int dt) {
  return dt == 1 ? __bfloat162float(static_cast<const __nv_bfloat16*>(p)[i]) : static_cast<const float*>(p)[i];
}
__device__ __forceinline__ void rec_store(void* p, long long i, int dt, float v) {
  if (dt == 1) static_cast<__nv_bfloat16*>(p)[i] = __float2bfloat16(v);
  else static_cast<float*>(p)[i] = v;
}
__device__ __forceinline__ float sigmoidf(float x) { return 1.f / (1.f + __expf(-x)); }

__global__ void __launch_bounds__(256)
lstm_cell_forward_kernel(const float* __restrict__ z, long long ldz, const float* __restrict__ c_prev,
                         float* __restrict__ c, float* __restrict__ gates, long long ldg,
                         __nv_bfloat16* __restrict__ h_next, long long ld_next, void* __restrict__ out, int out_dt,
                         long long ld_out, int B, int H) {
  vb::pdl_sync();
  const long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (idx >= static_cast<long long>(B) * H) return;
  const int b = static_cast<int>(idx / H), j = static_cast<int>(idx - static_cast<long long>(b) * H);
  const float* zr = z + b * ldz;
  const float gi = sigmoidf(zr[j]);
  const float gf = sigmoidf(zr[H + j]);
  const float gg = tanhf(zr[2 * H + j]);
  const float go = sigmoidf(zr[3 * H + j]);
  const float cp = c_prev ? c_prev[idx] : 0.f;
  const float cn = gf * cp + gi * gg;
  const float h = go * tanhf(cn);
  c[idx] = cn;
  float* gr = gates + b * ldg;
  gr[j] = gi;
  gr[H + j] = gf;
  gr[2 * H + j] = gg;
  gr[3 * H + j] = go;
  if (h_next) h_next[b * ld_next + j] = __float2bfloat16(h);
  if (out) rec_store(out, b * ld_out + j, out_dt, h);
}

// dc: running dL/dc_t (already multiplied by f_{t+1} on the way in); `first` = last time step (nothing flows in yet)
__global__ void __launch_bounds__(256)
lstm_cell_backward_kernel(const void* __restrict__ dh_ext, int ext_dt, long long ld_ext,
                          const float* __restrict__ dh_rec, long long ld_rec, float* __restrict__ dc, int first,
                          const float* __restrict__ gates, long long ldg, const float* __restrict__ c_prev,
                          const float* __restrict__ c, __nv_bfloat16* __restrict__ dz, long long ld_dz, int B, int H) {
  vb::pdl_sync();
  const long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (idx >= static_cast<long long>(B) * H) return;
  const int b = static_cast<int>(idx / H), j = static_cast<int>(idx - static_cast<long long>(b) * H);
  float dh = 0.f;
  if (dh_ext) dh += rec_load(dh_ext, b * ld_ext + j, ext_dt);
  if (dh_rec) dh += dh_rec[b * ld_rec + j];
  const float* gr = gates + b * ldg;
  const float gi = gr[j], gf = gr[H + j], gg = gr[2 * H + j], go = gr[3 * H + j];
  const float tc = tanhf(c[idx]);
  const float dct = (first ? 0.f : dc[idx]) + dh * go * (1.f - tc * tc);
  const float cp = c_prev ? c_prev[idx] : 0.f;
  dc[idx] = dct * gf;
  __nv_bfloat16* dr = dz + b * ld_dz;
  dr[j] = __float2bfloat16(dct * gg * gi * (1.f - gi));
  dr[H + j] = __float2bfloat16(dct * cp * gf * (1.f - gf));
  dr[2 * H + j] = __float2bfloat16(dct * gi * (1.f - gg * gg));
  dr[3 * H + j] = __float2bfloat16(dh * tc * go * (1.f - go));
}

__global__ void __launch_bounds__(256)
rnn_cell_forward_kernel(const float* __restrict__ z, long long ldz, float* __restrict__ h_save,
                        __nv_bfloat16* __restrict__ h_next, long long ld_next, void* __restrict__ out, int out_dt,
                        long long ld_out, int B, int H) {
  vb::pdl_sync();
  const long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (idx >= static_cast<long long>(B) * H) return;
  const int b = static_cast<int>(idx / H), j = static_cast<int>(idx - static_cast<long long>(b) * H);
  const float h = tanhf(z[b * ldz + j]);
  h_save[idx] = h;
  if (h_next) h_next[b * ld_next + j] = __float2bfloat16(h);
  if (out) rec_store(out, b * ld_out + j, out_dt, h);
}

__global__ void __launch_bounds__(256)
rnn_cell_backward_kernel(const void* __restrict__ dh_ext, int ext_dt, long long ld_ext,
                         const float* __restrict__ dh_rec, long long ld_rec, const float* __restrict__ h_save,
                         __nv_bfloat16* __restrict__ dz, long long ld_dz, int B, int H) {
  vb::pdl_sync();
  const long long idx = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (idx >= static_cast<long long>(B) * H) return;
  const int b = static_cast<int>(idx / H), j = static_cast<int>(idx - static_cast<long long>(b) * H);
  float dh = 0.f;
  if (dh_ext) dh += rec_load(dh_ext, b * ld_ext + j, ext_dt);
  if (dh_rec) dh += dh_rec[b * ld_rec + j];
  const float h = h_save[idx];
  dz[b * ld_dz + j] = __float2bfloat16(dh * (1.f - h * h));
}

}  // namespace vb

using namespace vb;

static inline int rec_blocks(int B, int H) { return static_cast<int>((static_cast<long long>(B) * H + 255) / 256); }

extern "C" int vb_lstm_cell_forward(const float* z, long long ldz, const float* c_prev, float* c, float* gates,
                                    long long ldg, void* h_next, long long ld_next, void* out, int out_dtype,
                                    long long ld_out, int B, int H, cudaStream_t s) {
  VB_REQUIRE(B > 0 && H > 0, "vb_lstm_cell_forward: empty cell");
  VB_REQUIRE(out_dtype == 0 || out_dtype == 1, "vb_lstm_cell_forward: output must be fp32 or bf16");
  vb::launch_k(lstm_cell_forward_kernel, dim3(rec_blocks(B, H)), dim3(256), 0, s, z, ldz, c_prev, c, gates, ldg,
                                                           static_cast<__nv_bfloat16*>(h_next), ld_next, out, out_dtype,
                                                           ld_out, B, H);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_lstm_cell_backward(const void* dh_ext, int ext_dtype, long long ld_ext, const float* dh_rec,
                                     long long ld_rec, float* dc, int first, const float* gates, long long ldg,
                                     const float* c_prev, const float* c, void* dz, long long ld_dz, int B, int H,
                                     cudaStream_t s) {
  VB_REQUIRE(B > 0 && H > 0, "vb_lstm_cell_backward: empty cell");
  VB_REQUIRE(ext_dtype == 0 || ext_dtype == 1, "vb_lstm_cell_backward: err_output must be fp32 or bf16");
  vb::launch_k(lstm_cell_backward_kernel, dim3(rec_blocks(B, H)), dim3(256), 0, s, dh_ext, ext_dtype, ld_ext, dh_rec, ld_rec, dc, first, gates,
                                                            ldg, c_prev, c, static_cast<__nv_bfloat16*>(dz), ld_dz, B,
                                                            H);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_rnn_cell_forward(const float* z, long long ldz, float* h_save, void* h_next, long long ld_next,
                                   void* out, int out_dtype, long long ld_out, int B, int H, cudaStream_t s) {
  VB_REQUIRE(B > 0 && H > 0, "vb_rnn_cell_forward: empty cell");
  VB_REQUIRE(out_dtype == 0 || out_dtype == 1, "vb_rnn_cell_forward: output must be fp32 or bf16");
  vb::launch_k(rnn_cell_forward_kernel, dim3(rec_blocks(B, H)), dim3(256), 0, s, z, ldz, h_save, static_cast<__nv_bfloat16*>(h_next), ld_next,
                                                          out, out_dtype, ld_out, B, H);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}

extern "C" int vb_rnn_cell_backward(const void* dh_ext, int ext_dtype, long long ld_ext, const float* dh_rec,
                                    long long ld_rec, const float* h_save, void* dz, long long ld_dz, int B, int H,
                                    cudaStream_t s) {
  VB_REQUIRE(B > 0 && H > 0, "vb_rnn_cell_backward: empty cell");
  VB_REQUIRE(ext_dtype == 0 || ext_dtype == 1, "vb_rnn_cell_backward: err_output must be fp32 or bf16");
  vb::launch_k(rnn_cell_backward_kernel, dim3(rec_blocks(B, H)), dim3(256), 0, s, dh_ext, ext_dtype, ld_ext, dh_rec, ld_rec, h_save,
                                                           static_cast<__nv_bfloat16*>(dz), ld_dz, B, H);
  VB_CHECK_CUDA(cudaGetLastError());
  return 0;
}
