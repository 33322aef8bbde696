"""Sweep tile width x CTA pairing x split-K of the tcgen05 GEMM over the small-batch fully connected shapes of
AlexNet / the wide MLP, against cuBLAS (torch) with the same output type.  CUDA-event timed, L2 flushed between
iterations.  Writes gpurun_out/sweep_fc_gemm.json; the scheduler's own choice is the row with block_n = split_k = 0."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from veles_b200.ops import kernels as k

dev = torch.device("cuda:0")
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)


def timeit(fn, iters=9, warm=2):
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


def sweep(name, M, N, K, a_mn, b_mn, out_dtype):
    a = torch.randn((K, M) if a_mn else (M, K), device=dev).bfloat16()
    b = torch.randn((K, N) if b_mn else (N, K), device=dev).bfloat16()
    out = torch.zeros(M, N, device=dev, dtype=out_dtype)
    acc32 = torch.zeros(M, N, device=dev, dtype=torch.float32)
    bias = torch.zeros(N, device=dev)
    A = a.t() if a_mn else a
    Bm = b if b_mn else b.t()
    try:
        ref = timeit(lambda: torch.mm(A, Bm, out_dtype=out_dtype))
    except Exception:  # noqa: BLE001
        ref = timeit(lambda: torch.matmul(A, Bm))
    rows = []
    k.set_pair_mode(0)
    auto = timeit(lambda: k.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn))
    plan = k.gemm_plan(M, N, K, b_mn=b_mn, can_split=False)
    rows.append(dict(name=name, block_n=0, split_k=0, pair="auto", plan=list(plan), us=round(auto * 1e3, 1)))
    for pair_mode in (1, 2):
        k.set_pair_mode(pair_mode)
        for bn in (64, 128, 192, 256):
            for sk in (1, 2, 3, 4, 6, 8):
                try:
                    got = k.gemm_plan(M, N, K, b_mn=b_mn, can_split=True, block_n=bn)
                    if pair_mode == 2 and not got[2]:
                        continue
                    if sk == 1:
                        t = timeit(lambda: k.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, block_n=bn))
                        t2 = None
                    else:
                        # split-K accumulates into fp32; the two-phase form adds the bias / activation pass
                        t = timeit(lambda: k.gemm(a, b, acc32, a_mn=a_mn, b_mn=b_mn, block_n=bn, split_k=sk,
                                                  accumulate=True))
                        if out_dtype == torch.float32:
                            t2 = None
                        else:
                            def both():
                                k.gemm(a, b, acc32, a_mn=a_mn, b_mn=b_mn, block_n=bn, split_k=sk, accumulate=True)
                                k.bias_act(acc32, out, bias=bias, act=k.ACT_LINEAR)
                            t2 = timeit(both)
                    rows.append(dict(name=name, block_n=bn, split_k=sk, pair=pair_mode == 2, us=round(t * 1e3, 1),
                                     two_phase_us=None if t2 is None else round(t2 * 1e3, 1)))
                except Exception as e:  # noqa: BLE001
                    rows.append(dict(name=name, block_n=bn, split_k=sk, pair=pair_mode == 2, error=str(e)[:80]))
    k.set_pair_mode(0)
    ok = [r for r in rows if "us" in r and r["split_k"] != 0]
    best = min(ok, key=lambda r: r["two_phase_us"] if r.get("two_phase_us") else r["us"])
    print(json.dumps(dict(name=name, M=M, N=N, K=K, cublas_us=round(ref * 1e3, 1), auto_us=rows[0]["us"],
                          auto_plan=rows[0]["plan"], best=best)), flush=True)
    return dict(name=name, M=M, N=N, K=K, a_mn=a_mn, b_mn=b_mn, out=str(out_dtype), cublas_us=round(ref * 1e3, 1),
                rows=rows)


if __name__ == "__main__":
    B = 256
    res = [
        sweep("fc6_fwd", B, 4096, 9216, False, False, torch.bfloat16),
        sweep("fc6_dx", B, 9216, 4096, False, True, torch.bfloat16),
        sweep("fc6_dw", 4096, 9216, B, True, True, torch.float32),
        sweep("fc7_fwd", B, 4096, 4096, False, False, torch.bfloat16),
        sweep("fc7_dx", B, 4096, 4096, False, True, torch.bfloat16),
        sweep("fc7_dw", 4096, 4096, B, True, True, torch.float32),
        sweep("fc8_fwd", B, 1000, 4096, False, False, torch.bfloat16),
        sweep("fc8_dx", B, 4096, 1000, False, True, torch.bfloat16),
        sweep("fc8_dw", 1000, 4096, B, True, True, torch.float32),
        sweep("mlp_fwd", 1024, 4096, 4096, False, False, torch.bfloat16),
        sweep("mlp_dw", 4096, 4096, 1024, True, True, torch.float32),
    ]
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open("gpurun_out/sweep_fc_gemm.json", "w"), indent=1)
