"""Epilogue sensitivity of the tcgen05 GEMM on the small-K conv shapes (bias / activation / dtype / tile)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from veles_b200.ops import kernels as k
from bench_gemm import timeit, dev


def case(name, M, N, K, ld=None, bias=False, act=0, out_dtype=torch.bfloat16, block_n=0, direct=False):
    ld = ld or K
    a = torch.randn(M, ld, device=dev).bfloat16()[:, :K]
    b = torch.randn(N, ld, device=dev).bfloat16()[:, :K]
    out = torch.zeros(M, N, device=dev, dtype=out_dtype)
    bv = torch.randn(N, device=dev) if bias else None
    k.set_direct_epilogue(direct)
    t = timeit(lambda: k.gemm(a, b, out, bias=bv, act=act, block_n=block_n))
    k.set_direct_epilogue(False)
    gb = (M * K * 2 + M * N * out.element_size()) / 1e9
    r = dict(name=name, M=M, N=N, K=K, ld=ld, bias=bias, act=act, out=str(out_dtype), block_n=block_n, direct=direct,
             ms=round(t, 4), gbps=round(gb / t * 1e3, 1))
    print(json.dumps(r), flush=True)
    return r


if __name__ == "__main__":
    B = 256
    res = []
    for kw in (dict(), dict(bias=True), dict(bias=True, act=4), dict(bias=True, act=4, direct=True),
               dict(bias=True, act=4, block_n=64), dict(out_dtype=torch.float32)):
        res.append(case("conv1", B * 3025, 96, 368, **kw))
    res.append(case("conv1_k363", B * 3025, 96, 363, ld=368, bias=True, act=4))
    for kw in (dict(), dict(bias=True, act=4)):
        res.append(case("conv2g", B * 729, 128, 1200, **kw))
        res.append(case("conv3", B * 169, 384, 2304, **kw))
    os.makedirs("../gpurun_out", exist_ok=True)
    json.dump(res, open("../gpurun_out/bench_gemm_epi.json", "w"), indent=1)
