"""Run ONE GEMM shape a few times (for ncu captures of the tcgen05 kernel)."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from veles_b200.ops import kernels as k

SHAPES = {
    "conv3_fwd": (43264, 384, 2304, False, False),
    "square8k": (8192, 8192, 8192, False, False),
    "square4k": (4096, 4096, 4096, False, False),
    "conv3_dw": (384, 2304, 43264, True, True),
    "fc6_fwd": (256, 4096, 9216, False, False),
}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", default="conv3_fwd")
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--pair", type=int, default=0)
    ap.add_argument("--block-n", type=int, default=0)
    ap.add_argument("--bias-act", action="store_true")
    a = ap.parse_args()
    M, N, K, a_mn, b_mn = SHAPES[a.shape]
    dev = torch.device("cuda:0")
    A = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
    B = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
    acc = a_mn and b_mn
    out = torch.zeros(M, N, device=dev, dtype=torch.float32 if acc else torch.bfloat16)
    bias = torch.randn(N, device=dev) if a.bias_act else None
    k.set_pair_mode(a.pair)
    for _ in range(a.iters):
        k.gemm(A, B, out, a_mn=a_mn, b_mn=b_mn, accumulate=acc, split_k=0 if acc else 1, block_n=a.block_n,
               bias=bias, act=4 if a.bias_act else 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        k.gemm(A, B, out, a_mn=a_mn, b_mn=b_mn, accumulate=acc, split_k=0 if acc else 1, block_n=a.block_n,
               bias=bias, act=4 if a.bias_act else 0)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("%s: %.4f ms  %.1f TFLOP/s" % (a.shape, ms, 2.0 * M * N * K / ms / 1e9))
