/* veles_b200 — C ABI of the sm_100a kernel library (lib/libveles_b200.so).
 *
 * Every entry point enqueues work on `stream` and returns 0 on success or a negative code
 * (message via vb_last_error()).  Pointers are device pointers unless stated otherwise.
 * dtype codes: 0 = float32, 1 = bfloat16, 2 = uint8, 3 = int32.
 * Activation codes: 0 linear, 1 tanh (1.7159*tanh(0.6666 s)), 2 sigmoid, 3 "relu" = log(1+e^s),
 *                   4 strict relu, 5 log, 6 tanhlog, 7 sincos, 8 mul.
 * Reference citations name the file of devbib/veles each entry point replaces.
 */
#ifndef VELES_B200_H_
#define VELES_B200_H_

#include <cuda_runtime.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime -------------------------------------------------------------------------------- */
const char* vb_last_error(void);
int vb_version(void);
int vb_sm_count(void);
int vb_device_cc(void);

/* ---- tcgen05/TMEM/TMA GEMM: D[M,N] = epilogue(sum_k A(m,k) B(n,k))
 * replaces cuBLAS sgemm/dgemm call sites (veles/accelerated_units.py:761-766, veles/backends.py:829-836)
 * and the OpenCL tiled GEMM + STORE_OUTPUT epilogue (ocl/matrix_multiplication_precise.cl:43-185).
 * a_mn_major/b_mn_major: 0 = operand stored [M|N, K] row-major, 1 = stored [K, M|N] row-major.
 * Z/deriv: optional saved activation output, D *= f'(Z).  split_k: 0 = auto, >1 needs accumulate. */
int vb_gemm_bf16(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                 void* D, long long ldd, int M, int N, int K, const float* bias, int act, int out_fp32,
                 int accumulate, const void* Z, long long ldz, int deriv, float alpha, int split_k, int block_n,
                 cudaStream_t stream);

/* fp32 operands multiplied as TF32 (tcgen05.mma.kind::tf32), fp32 accumulate: the "float" precision mode
 * (precision_type of veles/config.py:236-241).  A, B and Z are float; leading dimensions multiples of 4. */
int vb_gemm_tf32(const float* A, long long lda, int a_mn_major, const float* B, long long ldb, int b_mn_major, void* D,
                 long long ldd, int M, int N, int K, const float* bias, int act, int out_fp32, int accumulate,
                 const float* Z, long long ldz, int deriv, float alpha, int split_k, int block_n, cudaStream_t stream);

/* The same product with the weight operand B row-sharded over `world` replicas (one device pointer per rank to the
 * same [N,K] / [K,N] matrix in symmetric memory): rows [r*shard_rows, (r+1)*shard_rows) are TMA-loaded from
 * b_peers[r].  Fuses the all-gather of updated weight shards into the GEMM (no counterpart in the reference, whose
 * master pushed pickled weights over ZeroMQ: veles/server.py:401-430). */
int vb_gemm_bf16_sharded_b(const void* A, long long lda, int a_mn_major, const void* const* b_peers, int world,
                           int shard_rows, int own_rank, long long ldb, int b_mn_major, void* D, long long ldd, int M,
                           int N, int K, const float* bias, int act, int out_fp32, int accumulate, const void* Z,
                           long long ldz, int deriv, float alpha, int split_k, int block_n, cudaStream_t stream);

/* Implicit-GEMM convolution (Znicz conv / gd_conv, absent from the checkout: SURVEY 2.4): one operand is fetched by
 * TMA in im2col mode from the NHWC bf16 tensor T (tB x tH x tW pixels, t_pitch elements between pixels, t_ch
 * channels visible, group offset applied by the caller).  mode 1 forward: D[pixels, N] = act(im2col(T) W^T + bias),
 * other = W [other_rows, ky*kx*cg]; mode 2 err_input (unit stride): T = dY, D[tB*oH*oW, cg] = convT * f'(Z);
 * mode 3 weight gradient: other = dY [pixels, M], D[M, ky*kx*cg] fp32 (+)=.  (low_h, low_w) = -padding
 * (mode 2: -(k-1-padding)); oH x oW = traversal positions per image. */
int vb_conv_gemm_bf16(int mode, const void* T, int tB, int tH, int tW, long long t_pitch, int t_ch, int ky, int kx,
                      int sy, int sx, int low_h, int low_w, int oH, int oW, const void* other, long long ld_other,
                      int other_rows, int cg, void* D, long long ldd, int M, int N, const float* bias, int act,
                      int out_fp32, int accumulate, const void* Z, long long ldz, int deriv, float alpha, int split_k,
                      int block_n, cudaStream_t stream);
/* All groups of a grouped convolution in ONE launch (modes 1, 2): T / other / D are the whole tensors, t_ch and N the
 * total widths; per-group channel and kernel counts must be multiples of 64. */
int vb_conv_gemm_grouped_bf16(int mode, int groups, const void* T, int tB, int tH, int tW, long long t_pitch, int t_ch,
                              int ky, int kx, int sy, int sx, int low_h, int low_w, int oH, int oW, const void* other,
                              long long ld_other, int other_rows, int cg, void* D, long long ldd, int M, int N,
                              const float* bias, int act, int out_fp32, const void* Z, long long ldz, int deriv,
                              int block_n, cudaStream_t stream);
/* Halo-tile variant for unit-stride layers (OW + kx - 1 <= 64): every 64-channel block of the walked tensor is loaded
 * once per output tile (4-D TMA box with halo) and all ky*kx taps are issued from that shared-memory tile.
 * mode 1 forward / 2 err_input as above; D is [tB, oH, oW, N] with pixel pitch d_pitch.  Returns 1 when unsupported. */
int vb_conv_halo_bf16(int mode, const void* T, int tB, int tH, int tW, long long t_pitch, int t_ch, int ky, int kx,
                      int low_h, int low_w, int oH, int oW, const void* W, long long ldw, int w_rows, int cg, void* D,
                      long long d_pitch, int N, const float* bias, int act, const void* Z, long long ldz, int deriv,
                      cudaStream_t stream);
/* x [B,H,W,C] (dtype 0 f32 / 1 bf16) -> bf16 [B, ceil(H/sy), ceil(W/sx), sy*sx*C]: strided few-channel convolutions
 * become unit-stride ones the implicit GEMM can walk. */
int vb_space_to_depth(const void* x, int dtype, void* xs, int B, int H, int W, int C, int sy, int sx,
                      cudaStream_t stream);
/* Minibatch gather (rows by shuffled index, cast, (x - mean) * rdisp, zero rows past `count`) fused with that
 * transform: xs [max_rows, ceil(H/sy), ceil(W/sx), sy*sx*C] bf16 is written directly and the NHWC minibatch never
 * exists.  src_dtype: 0 f32, 1 bf16, 2 u8.  Returns 1 (nothing launched) when the geometry is not covered. */
int vb_gather_s2d(const void* src, int src_dtype, long long src_ld, const int* indices, int start, int count,
                  int max_rows, int H, int W, int C, int sy, int sx, void* xs, const float* mean, const float* rdisp,
                  const int* labels, int* mb_labels, int* mb_indices, cudaStream_t stream);
/* dst[r, j] (+)= src[r, map[j]] (0 where map[j] < 0); dtypes 0 f32 / 1 bf16. */
int vb_gather_cols(const void* src, int src_dtype, long long ld_src, void* dst, int dst_dtype, long long ld_dst,
                   const int* map, int rows, int cols, int accumulate, cudaStream_t stream);
/* Test hooks: force the per-thread store epilogue; 0 cost model / 1 single-CTA tiles / 2 CTA-pair tiles. */
/* Second phase of a split-K layer GEMM: out = act(acc + bias) [* f'(z)]; acc (fp32, what the K-splits reduce-added
 * into) is re-zeroed when zero_acc. */
int vb_bias_act_2d(float* acc, long long ld_acc, const float* bias, int act, const void* z, long long ldz, int deriv,
                   void* out, int out_dtype, long long ld_out, int M, int N, int zero_acc, cudaStream_t stream);
int vb_gemm_plan(int M, int N, int K, int b_mn_major, int can_split, int block_n, int* bn_out, int* split_out,
                 int* pair_out);
int vb_gemm_set_dynamic_schedule(int on);
int vb_gemm_set_direct_epilogue(int on);
int vb_gemm_set_pair_mode(int mode);

/* ---- loader path: cuda/fullbatch_loader.cu:9-61, cuda/mean_disp_normalizer.cu:11-19, cuda/join.jcu:10-35 */
int vb_gather_minibatch(const void* src, int src_dtype, long long src_ld, const int* indices, int start, int count,
                        int max_rows, int row_size, void* dst, int dst_dtype, long long dst_ld, const float* mean,
                        const float* rdisp, const int* labels, int* mb_labels, int* mb_indices, cudaStream_t stream);
int vb_mean_disp_normalize(const void* in, int in_dtype, const float* mean, const float* rdisp, void* out,
                           int out_dtype, int batch, int sample_size, cudaStream_t stream);
int vb_copy_2d(const void* src, int src_dtype, long long src_ld, void* dst, int dst_dtype, long long dst_ld,
               int rows, int cols, cudaStream_t stream);
int vb_cast(const void* in, int in_dtype, void* out, int out_dtype, long long n, cudaStream_t stream);

/* ---- Znicz element-wise units (sources absent from the reference checkout) -------- */
int vb_activation_forward(int act, const void* x, void* y, long long n, int dtype, float factor, cudaStream_t stream);
int vb_activation_backward(int act, const void* err_out, const void* x, const void* y, void* err_in, long long n,
                           int dtype, float factor, cudaStream_t stream);
int vb_dropout(const void* in, void* out, long long n, int dtype, const void* states_in, void* states_out,
               int n_states, float ratio, cudaStream_t stream);
int vb_gd_update(float* w, float* grad, float* vel, float* acc, void* w_bf16, long long n, float lr, float wd,
                 float l1_vs_l2, float moment, float grad_scale, int solver, float eps, float rho, int zero_grad,
                 cudaStream_t stream);

/* ---- Znicz spatial units -------------------------------------------------------------------------------- */
int vb_im2col(const void* x, void* cols, int B, int H, int W, int C, int OH, int OW, int ky, int kx, int sy, int sx,
              int pad_top, int pad_left, int groups, long long ld, cudaStream_t stream);
int vb_col2im(const void* dcols, void* dx, int B, int H, int W, int C, int OH, int OW, int ky, int kx, int sy,
              int sx, int pad_top, int pad_left, int groups, long long ld, const void* z, int deriv,
              cudaStream_t stream);
int vb_pool_forward(int kind, const void* x, void* y, int* offs, int dtype, int B, int H, int W, int C, int OH,
                    int OW, int ky, int kx, int sy, int sx, const void* rnd_u16, cudaStream_t stream);
int vb_pool_backward(int kind, const void* dy, const int* offs, void* dx, int dtype, int B, int H, int W, int C,
                     int OH, int OW, int ky, int kx, int sy, int sx, cudaStream_t stream);
/* backward variants that also multiply by f'(z) of the activation which produced the layer input */
int vb_pool_backward_fused(int kind, const void* dy, const int* offs, void* dx, int dtype, int B, int H, int W, int C,
                           int OH, int OW, int ky, int kx, int sy, int sx, const void* z, int deriv,
                           cudaStream_t stream);
/* Compact argmax records (bf16, C % 8 == 0): one byte per output = window-local index (row*kx + col) instead of the
 * 4-byte global offset; return 1 if the shape is unsupported (use the int32 entry points), 0 on success.
 * vb_pool_expand_offsets rebuilds the Znicz-visible int32 input_offset on demand. */
int vb_pool_forward_local(int kind, const void* x, void* y, int* offs, unsigned char* offs8, int B, int H, int W,
                          int C, int OH, int OW, int ky, int kx, int sy, int sx, cudaStream_t stream);
int vb_pool_backward_local(int kind, const void* dy, const unsigned char* offs8, void* dx, int B, int H, int W, int C,
                           int OH, int OW, int ky, int kx, int sy, int sx, const void* z, int deriv,
                           cudaStream_t stream);
int vb_pool_expand_offsets(const unsigned char* offs8, int* offs, int B, int H, int W, int C, int OH, int OW, int ky,
                           int kx, int sy, int sx, cudaStream_t stream);
int vb_lrn_backward_fused(const void* x, const void* dy, void* dx, int dtype, long long pixels, int C, int n,
                          float alpha, float beta, float k, int deriv, cudaStream_t stream);
int vb_lrn_forward(const void* x, void* y, int dtype, long long pixels, int C, int n, float alpha, float beta,
                   float k, cudaStream_t stream);
int vb_lrn_backward(const void* x, const void* dy, void* dx, int dtype, long long pixels, int C, int n, float alpha,
                    float beta, float k, cudaStream_t stream);
int vb_softmax_eval(const void* sm_in, int in_dtype, long long lds, void* y, long long ldy, void* err, long long lde,
                    int out_dtype, const int* labels, int* max_idx, int batch, int N, float err_scale, int* n_err,
                    int* confusion, float* max_err_sum, float* loss_sum, cudaStream_t stream);
int vb_mse_eval(const void* y, long long ldy, const void* target, long long ldt, void* err, long long lde, int dtype,
                int batch, int batch_valid, int N, float err_scale, float* metrics, float* mse_per_sample,
                cudaStream_t stream);

/* ---- reductions: cuda/matrix_reduce.cu:35-81 ---------------------------------------------------------------- */
int vb_row_sum(const void* a, int dtype, long long lda, int rows, int cols, float* out, cudaStream_t stream);
long long vb_col_sum_scratch_elems(int rows, int cols);
/* out[j] (+)= alpha * sum_i a[i,j]; `scratch` holds vb_col_sum_scratch_elems(rows, cols) floats. */
int vb_col_sum(const void* a, int dtype, long long lda, int rows, int cols, float* out, float alpha, int accumulate,
               float* scratch, cudaStream_t stream);

/* vb_col_sum whose result is used on the spot as the gradient of the vector w (a layer's bias): momentum GD step
 * v = moment * v - lr * (g + wd * ((1 - l1_vs_l2) * w + 0.5 * l1_vs_l2 * sign w)), w += v, applied by the reduction's
 * second phase; `grad` (needed only with accumulate) is left zeroed. */
int vb_col_sum_gd(const void* a, int dtype, long long lda, int rows, int cols, float* grad, float alpha, int accumulate,
                  float* scratch, float* w, float* vel, float lr, float wd, float l1_vs_l2, float moment,
                  cudaStream_t stream);

/* Column sum for callers that cannot allocate device memory (fixed memory budgets, code inside a stream
 * capture, latency-critical loops where cudaMalloc is forbidden).  Same result as vb_col_sum, but the
 * inter-block scratch is a fixed 4 MiB array inside the library, so only inputs and outputs are passed.
 * This entry point has no counterpart in the reference.
 * Limits: ceil(rows/64) * cols <= vb_col_sum_static_max_elems() (1 Mi floats) and ceil(rows/64) <= 65535
 * (rows <= 4 194 240: the slabs are the grid's y dimension); larger inputs are rejected with an error before
 * anything is launched.  The first call on each device looks up the array's device
 * address (no allocation) — make it before any stream capture.  All calls on a device share the array:
 * calls whose kernels can overlap on the GPU race with each other; nothing needs zeroing between calls. */
long long vb_col_sum_static_max_elems(void);
int vb_col_sum_static(const void* a, int dtype, long long lda, int rows, int cols, float* out, float alpha,
                      int accumulate, cudaStream_t stream);

/* ---- PRNG: cuda/random.cu:45-73 (xorshift1024*), :107-128 (xorshift128+) ------------------------------------- */
int vb_xorshift1024star(void* states, int n_states, int rounds, void* out, cudaStream_t stream);
int vb_xorshift128plus(void* states, long long n, void* out, cudaStream_t stream);
int vb_fill_random(void* states, int n_states, float* out, long long n, int kind, float amp, cudaStream_t stream);

/* ---- batched signal kernels (csrc/signal.cu): device side of libVeles/simd ------------------------------------- */
/* Rows are fp32, `batch` of them, pitches in elements.  out has x_length + h_length - 1 elements per row;
 * h_pitch 0 shares one filter between all rows; reverse = 1 gives cross-correlation (correlate.h:58-135). */
int vb_sig_convolve(const float* x, long long x_pitch, int x_length, const float* h, long long h_pitch, int h_length,
                    int reverse, float* out, long long out_pitch, int batch, cudaStream_t stream);
/* One wavelet analysis step (wavelet.h:96-137).  lowpass / highpass are HOST arrays of `order` (<= 76) taps.
 * stationary = 0: hi / lo get length / 2 outputs; 1: `length` outputs with the taps 2^(level-1) apart.
 * ext: 0 periodic, 1 mirror, 2 constant, 3 zero. */
int vb_sig_wavelet(const float* x, long long x_pitch, int length, const float* lowpass, const float* highpass,
                   int order, int level, int stationary, int ext, float* hi, float* lo, long long out_pitch,
                   int batch, cudaStream_t stream);
/* out[2 r] = min, out[2 r + 1] = max of row r (normalize.h:84-89). */
int vb_sig_minmax_f32(const float* x, long long pitch, int length, int batch, float* out, cudaStream_t stream);
/* uint8 planes -> (v - min) / ((max - min) / 2) - 1 (normalize.h:48-82).  minmax: int32 [batch][2], computed here
 * unless have_minmax; dst may be null (range pass only). */
int vb_sig_normalize_u8(const unsigned char* src, long long plane_pitch, int stride, int width, int height, int batch,
                        int* minmax, int have_minmax, float* dst, long long dst_plane_pitch, int dst_stride,
                        cudaStream_t stream);
/* Strict interior extrema in index order (detect_peaks.h).  type 1 max, 2 min, 3 both.  counts: scratch of
 * batch * ceil(length / 1024) ints; positions / values: [batch][capacity]; totals[batch] = peaks found. */
int vb_sig_detect_peaks(const float* x, long long pitch, int length, int batch, int type, int* counts, int* positions,
                        float* values, int capacity, int* totals, cudaStream_t stream);
/* out = a * b (or a * conj(b)) * scale on interleaved complex rows; pitches in complex elements, b_pitch 0 broadcasts. */
int vb_sig_complex_mul(const float* a, long long a_pitch, const float* b, long long b_pitch, float* out,
                       long long out_pitch, int n_complex, int batch, int conj_b, float scale, cudaStream_t stream);
/* kind 0 int16->float, 1 float->int16 (truncate), 2 int32->float, 3 float->int32, 4 half->float (arithmetic.h). */
int vb_sig_convert(const void* src, void* dst, long long n, int kind, cudaStream_t stream);

/* ---- recurrent cells (csrc/recurrent.cu): RNN / LSTM units of docs/source/manualrst_veles_algorithms.rst:108-135 - */
/* z: [B, 4H] fp32 pre-activations (gate blocks i f g o, bias included).  Writes the gate activations (saved for the
 * backward pass), c, and h both into h_next (bf16 h-slot of the next step's [x | h] operand; may be null) and into
 * `out` rows (fp32 = 0 / bf16 = 1; may be null).  c_prev null = zero state. */
int vb_lstm_cell_forward(const float* z, long long ldz, const float* c_prev, float* c, float* gates, long long ldg,
                         void* h_next, long long ld_next, void* out, int out_dtype, long long ld_out, int B, int H,
                         cudaStream_t stream);
/* dh = dh_ext + dh_rec (either may be null); dc is the running dL/dc (ignored on input when first != 0);
 * dz: bf16 [B, 4H] gradient w.r.t. the pre-activations. */
int vb_lstm_cell_backward(const void* dh_ext, int ext_dtype, long long ld_ext, const float* dh_rec, long long ld_rec,
                          float* dc, int first, const float* gates, long long ldg, const float* c_prev, const float* c,
                          void* dz, long long ld_dz, int B, int H, cudaStream_t stream);
int vb_rnn_cell_forward(const float* z, long long ldz, float* h_save, void* h_next, long long ld_next, void* out,
                        int out_dtype, long long ld_out, int B, int H, cudaStream_t stream);
int vb_rnn_cell_backward(const void* dh_ext, int ext_dtype, long long ld_ext, const float* dh_rec, long long ld_rec,
                         const float* h_save, void* dz, long long ld_dz, int B, int H, cudaStream_t stream);

/* ---- data-parallel exchange (csrc/comm_fused.cu): gradient all-reduce FUSED with the Znicz GD update and the
 * updated-weight broadcast, one launch per bucket of layers; replaces the master-slave exchange of
 * veles/server.py:141-169, 401-430 + veles/client.py:138-160, 333-342 + veles/workflow.py:513-529.
 * All slabs are symmetric memory (same offsets on every rank); *_mc are NVLS multicast aliases (or null),
 * *_peer arrays hold `world` device pointers.  segs: nseg x {offset, numel, shadow_offset, row_len, row_pitch,
 * flags} (flags bit 0: zero the other parity's gradients; bit 1: always publish the fp32 master); hyper: nseg x
 * {lr, wd, l1_vs_l2, moment}.  use_multimem bits: 0 multimem path, 1 sharded fp32 master, 2 shadow stays on its
 * owner (peer-TMA delivery), 3 handshake in its own 1-CTA launch, 4 PUSH form (the default of parallel/dp.py): every
 * rank stores the gradient parts its peers own into their staging slabs (stage_local / stage_peer, laid out like the
 * gradient slab), the owner reduces locally - no load crosses NVLink; cta_counter then has 128 words.
 * n_ctas is rounded up to whole 2-CTA clusters. */
int vb_fused_allreduce_gd(int world, int rank, int channel, int nseg, const long long* segs, const float* hyper,
                          float* grads_local, float* grads_mc, float** grads_peer, float* grads_other_local,
                          float* weights_local, float* weights_mc, float** weights_peer, float* vel_local,
                          void* shadow_local, void* shadow_mc, void** shadow_peer, unsigned int* flags_local,
                          unsigned int** flags_peer, unsigned int* epochs, unsigned int* cta_counter, int* status,
                          float grad_scale, int n_ctas, int use_multimem, double timeout_s, float* stage_local,
                          float** stage_peer, cudaStream_t stream);
/* The same kernel body with every rank emulated on the current GPU by one cooperative launch (blockIdx.y = rank,
 * P2P path): how a single-GPU box tests the flag protocol and the slice arithmetic.  params_dev: device scratch of
 * world * vb_fused_params_bytes() bytes.  No counterpart in the reference. */
long long vb_fused_params_bytes(void);
int vb_fused_allreduce_gd_emulated(int world, int channel, int nseg, const long long* segs, const float* hyper,
                                   float** grads, float** grads_other, float** weights, float** vel, void** shadow,
                                   unsigned int** flags, unsigned int** epochs, unsigned int** cta_counter,
                                   int** status, float grad_scale, int n_ctas, int options, double timeout_s,
                                   void* params_dev, float** stage, cudaStream_t stream);
/* Class / epoch ends of a data-parallel group, device-side (csrc/comm_counters.cu): all-reduce n doubles (op 0 sum,
 * 1 max; n = 0: barrier) through a region of symmetric memory of vb_counter_region_bytes(slot_elems) bytes per rank
 * (zeroed once; base_peer[r] = rank r's mapping), flags by st.release.sys / ld.acquire.sys; a missing peer sets
 * *status = 1001 after timeout_s and leaves *epoch alone.  The reference's master summed these out of its slaves'
 * pickled updates (veles/server.py:401-430). */
long long vb_counter_region_bytes(long long slot_elems);
int vb_counter_allreduce(int world, int rank, int n, int op, long long slot_elems, const double* in, double* out,
                         void* base_local, void** base_peer, unsigned int* epoch, int* status, double timeout_s,
                         cudaStream_t stream);
/* All ranks of that exchange on the current GPU in one cooperative launch (single-GPU tests). */
long long vb_counter_params_bytes(void);
int vb_counter_allreduce_emulated(int world, int n, int op, long long slot_elems, const double** ins, double** outs,
                                  void** bases, unsigned int** epochs, int** status, double timeout_s,
                                  void* params_dev, cudaStream_t stream);
/* Reassemble sharded fp32 state: for each (offset, numel) pair of segs copy every peer's slice (the update kernel's
 * slice arithmetic) out of peers[r] into local; order it after a vb_counter_allreduce barrier. */
int vb_gather_slices(int world, int rank, int nseg, const long long* segs, float* local, void** peers, int n_ctas,
                     cudaStream_t stream);

/* SMs that the persistent tcgen05 kernels leave free from now on (0 = none): the exchange kernel runs there beside
 * the backward pass.  Read at launch / graph-capture time; returns the previous value. */
int vb_set_sm_reserve(int n);

/* LRN backward (bf16) that also leaves per-block partial column sums of dx in partials[*blocks_out][C] (fp32, at most
 * max_blocks rows): the bias gradient of the convolution in front of the LRN, so that layer does not read dx again.
 * Returns 1 (nothing launched) when the shape has no summing kernel (C % 8, C <= 256, window <= 9). */
int vb_lrn_backward_colsum(const void* x, const void* dy, void* dx, long long pixels, int C, int n, float alpha,
                           float beta, float k, int deriv, float* partials, int max_blocks, int* blocks_out,
                           cudaStream_t stream);
/* Second phase of the column-sum kernels alone: grad[c] (+)= sum of `slabs` partial rows; with w / vel the momentum
 * GD step of that bias is applied instead and grad is left zeroed. */
int vb_col_sum_finish(const float* scratch, int slabs, int cols, float* grad, int accumulate, float* w, float* vel,
                      float lr, float wd, float l1_vs_l2, float moment, cudaStream_t stream);
/* Max-pooling backward (3x3 / stride 2, byte argmax) fused into LRN backward (bf16 NHWC): the dense gradient of the
 * LRN output is neither written nor read.  Returns 1 when the geometry is not covered (run the two kernels). */
int vb_lrn_pool_backward(const void* x, const void* dpool, const unsigned char* offs8, void* dx, int B, int H, int W,
                         int C, int OH, int OW, int ky, int kx, int sy, int sx, int n, float alpha, float beta, float k,
                         int deriv, cudaStream_t stream);
/* factor_ortho regulariser of the Znicz GD units: grad[i,j] += f * (sums[j] - w[i,j])  (by_row: sums[i]). */
int vb_ortho_grad(const float* w, long long ldw, float* grad, long long ldg, const float* sums, long long rows,
                  long long cols, float f, int by_row, cudaStream_t stream);
/* Halo-tile convolution tuning / test hook: sub-tiles per CTA tile (0 = choose, 1, 2); resident weights (-1 = choose,
 * 0 = never, 1 = whenever they fit). */
int vb_conv_halo_set_options(int mt, int resident);
/* Profiling hook of the halo-tile kernel: a device buffer of 24 x 8 long longs that CTA 0 fills with clock64 stamps at
 * the hand-over points of its first 24 tiles (scripts/trace_halo.py); null switches it off. */
int vb_conv_halo_set_trace(long long* buf);
/* Weight-gradient GEMM dW = A^T B whose epilogue applies the Znicz momentum step to w / vel (and refreshes the bf16
 * shadow) instead of storing the gradient; opt-in (measured slower than GEMM + vb_gd_update on this machine). */
int vb_gemm_bf16_fused_gd(const void* A, long long lda, int a_mn_major, const void* B, long long ldb, int b_mn_major,
                          int M, int N, int K, float* w, long long ldw, float* vel, void* shadow, long long ld_shadow,
                          float lr, float wd, float l1_vs_l2, float moment, float alpha, int block_n,
                          cudaStream_t stream);
/* 1: every kernel of the library is launched with programmatic stream serialisation from now on (each kernel starts
 * with griddepcontrol.wait, so results are unchanged); 0: ordinary launches.  Returns the previous value. */
int vb_set_pdl(int on);

#ifdef __cplusplus
}
#endif
#endif /* VELES_B200_H_ */
