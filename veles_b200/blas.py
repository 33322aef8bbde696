"""cuBLAS-signature GEMM facade over the tcgen05 kernels (successor of ``veles/ocl_blas.py:53-236`` and of the
per-thread cuBLAS handles in ``veles/backends.py:811-836``): units written against
``device.blas.sgemm(transA, transB, m, n, k, alpha, A, B, beta, C)`` keep working.

Column-major BLAS semantics are honoured by the usual identity ``C^T = op(B)^T op(A)^T``.  ``sgemm`` keeps fp32
operands and multiplies them as TF32 (``tcgen05.mma.kind::tf32``, fp32 accumulate); ``gemm`` takes bf16 operands
(others are converted, fp32 accumulate); ``dgemm`` refuses: Blackwell's tensor cores have no usable fp64 path and a
silent down-conversion would not be a dgemm.  ``beta`` must be 0 or 1.
"""
from __future__ import annotations

OP_N, OP_T = 0, 1


class BLAS(object):
    def __init__(self, device=None):
        self.device = device

    @staticmethod
    def _bf16(t):
        import torch
        return t if t.dtype == torch.bfloat16 else t.to(torch.bfloat16)

    def gemm(self, transA, transB, m, n, k, alpha, A, B, beta, C, lda=None, ldb=None, ldc=None):
        """Column-major: C[m,n] = alpha * op(A)[m,k] * op(B)[k,n] + beta * C.  ``A``, ``B``, ``C`` are 2-D torch
        tensors holding the column-major matrices as their row-major transposes (what Znicz passed)."""
        from .ops import kernels
        if beta not in (0, 0.0, 1, 1.0):
            raise ValueError("beta must be 0 or 1")
        # row-major view: C_rm[n,m] = op(B)_rm[n,k] . op(A)_rm[m,k]^T
        a = self._bf16(A)     # row-major [k,m] if transA == N else [m,k]
        b = self._bf16(B)     # row-major [n,k] if transB == N else [k,n]
        if beta and str(C.dtype) != "torch.float32":
            raise ValueError("beta=1 needs an fp32 C")
        kernels.gemm(b, a, C, a_mn=(transB != OP_N), b_mn=(transA == OP_N), alpha=float(alpha),
                     accumulate=bool(beta))
        return C

    def sgemm(self, transA, transB, m, n, k, alpha, A, B, beta, C, lda=None, ldb=None, ldc=None):
        """fp32 operands multiplied as TF32 (10-bit mantissa products, fp32 accumulate)."""
        import torch
        from .ops import kernels
        if beta not in (0, 0.0, 1, 1.0):
            raise ValueError("beta must be 0 or 1")
        if A.dtype != torch.float32 or B.dtype != torch.float32:
            return self.gemm(transA, transB, m, n, k, alpha, A, B, beta, C)
        if beta and C.dtype != torch.float32:
            raise ValueError("beta=1 needs an fp32 C")
        kernels.gemm_tf32(B, A, C, a_mn=(transB != OP_N), b_mn=(transA == OP_N), alpha=float(alpha),
                          accumulate=bool(beta))
        return C

    def dgemm(self, *args, **kwargs):
        raise NotImplementedError("dgemm: no fp64 tensor-core path on sm_100a in this library; use sgemm (TF32) or "
                                  "gemm (bf16), or torch.matmul on float64 tensors for an oracle")
