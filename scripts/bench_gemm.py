"""Micro-benchmark of the tcgen05 GEMM family vs torch.matmul (cuBLAS) on the layer shapes
of the BASELINE configs.  CUDA-event timed, L2 flushed between iterations."""
import json
import sys
import os

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from veles_b200.ops import kernels as k

dev = torch.device("cuda:0")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0 = torch.cuda.Event(enable_timing=True)
        e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def case(name, M, N, K, a_mn, b_mn, split_k=1, block_n=0, out_dtype=torch.bfloat16):
    a = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
    b = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
    out = torch.zeros(M, N, device=dev, dtype=out_dtype)
    acc = split_k > 1
    ours = timeit(lambda: k.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, split_k=split_k, accumulate=acc, block_n=block_n))
    A = a.t() if a_mn else a
    B = b if b_mn else b.t()
    ref = timeit(lambda: torch.matmul(A, B))
    fl = 2.0 * M * N * K
    r = dict(name=name, M=M, N=N, K=K, a_mn=a_mn, b_mn=b_mn, split_k=split_k, block_n=block_n,
             ours_ms=round(ours, 4), cublas_ms=round(ref, 4), ours_tflops=round(fl / ours / 1e9, 1),
             cublas_tflops=round(fl / ref / 1e9, 1))
    print(json.dumps(r), flush=True)
    return r


if __name__ == "__main__":
    B = 256
    res = []
    res.append(case("square8k", 8192, 8192, 8192, False, False))
    res.append(case("square8k_bn128", 8192, 8192, 8192, False, False, block_n=128))
    res.append(case("square4k", 4096, 4096, 4096, False, False))
    res.append(case("fc6_fwd", B, 4096, 9216, False, False))
    res.append(case("fc6_dx", B, 9216, 4096, False, True))
    res.append(case("fc6_dw", 4096, 9216, B, True, True, out_dtype=torch.float32))
    res.append(case("fc7_fwd", B, 4096, 4096, False, False))
    res.append(case("fc7_dw", 4096, 4096, B, True, True, out_dtype=torch.float32))
    res.append(case("mlp4096_b1024_fwd", 1024, 4096, 4096, False, False))
    res.append(case("conv1_fwd", B * 3025, 96, 368, False, False))
    res.append(case("conv2g_fwd", B * 729, 128, 1200, False, False))
    res.append(case("conv3_fwd", B * 169, 384, 2304, False, False))
    res.append(case("conv3_dcols", B * 169, 2304, 384, False, True))
    for sk in (1, 4, 16):
        res.append(case("conv3_dw_sk%d" % sk, 384, 2304, B * 169, True, True, split_k=sk, out_dtype=torch.float32))
    for sk in (1, 8, 32):
        res.append(case("conv1_dw_sk%d" % sk, 96, 368, B * 3025, True, True, split_k=sk, out_dtype=torch.float32))
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/bench_gemm.json", "w"), indent=1)
