// Internal prototypes of the bandwidth-tuned kernel variants (fast_paths.cu).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace vb {

struct FastConvGeom {
  int B, H, W, C, OH, OW, ky, kx, sy, sx, pt, pl, groups;
  long long ld;
};
struct FastPoolGeom {
  int B, H, W, C, OH, OW, ky, kx, sy, sx;
};

// each returns 1 = launched, 0 = shape not supported (use the generic kernel), <0 = error
int fast_im2col(const void* x, void* cols, const FastConvGeom& g, cudaStream_t s);
int fast_col2im(const void* dcols, void* dx, const FastConvGeom& g, const void* z, int deriv, cudaStream_t s);
// offs8 (optional): one byte per output element, the window-local argmax (row * kx + col)
int fast_pool_forward(int kind, const void* x, void* y, int* offs, uint8_t* offs8, const FastPoolGeom& g,
                      cudaStream_t s);
int fast_pool_backward(int kind, const void* dy, const int* offs, const uint8_t* offs8, void* dx,
                       const FastPoolGeom& g, const void* z, int deriv, cudaStream_t s);
int fast_pool_expand_offsets(const uint8_t* offs8, int* offs, const FastPoolGeom& g, cudaStream_t s);
int fast_lrn_forward(const void* x, void* y, long long pixels, int C, int n, float alpha, float beta, float k,
                     cudaStream_t s);
// colsum (optional): [blocks][C] fp32 partial rows of the per-channel sums of dx, one per launched block (at most
// max_blocks; the count comes back in *blocks_out); returns 0 (nothing launched) when the shape's kernel cannot
// produce them
int fast_lrn_backward(const void* x, const void* dy, void* dx, long long pixels, int C, int n, float alpha, float beta,
                      float k, int deriv, cudaStream_t s, float* colsum = nullptr, int max_blocks = 0,
                      int* blocks_out = nullptr);

// max pooling backward (3x3 / stride 2, byte argmax) fused into LRN backward: dpool / offs8 are the pooling layer's
// err_output and argmax bytes, x the LRN input, dx the LRN err_input
int fast_lrn_pool_backward(const void* x, const void* dpool, const uint8_t* offs8, void* dx, const FastPoolGeom& g, int n,
                           float alpha, float beta, float k, int deriv, cudaStream_t s);

}  // namespace vb
