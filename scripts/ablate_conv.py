"""Which resource bounds each implicit-GEMM convolution call?  Times every AlexNet convolution shape with parts of the
kernel switched off (vb_gemm_set_ablation: 1 no A loads, 2 no B loads, 4 no MMAs, 8 no epilogue body) - results of an
ablated run are garbage, only its duration means something.  CUDA events, L2 flushed between iterations."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from veles_b200.ops import kernels as k
from veles_b200.ops._lib import lib
from bench_gemm import timeit, dev
from bench_conv import LAYERS

MASKS = (0, 8, 1, 2, 3)

if __name__ == "__main__":
    res = []
    os.makedirs("gpurun_out", exist_ok=True)
    which = sys.argv[1:] or list(LAYERS)
    for name in which:
        B, H, W, C, n, ky, kx, p = LAYERS[name]
        OHW = H + 2 * p - ky + 1
        g = dict(OH=OHW, OW=OHW, ky=ky, kx=kx, sy=1, sx=1, pt=p, pl=p, H=H, W=W)
        x = torch.randn(B, H, W, C, device=dev).bfloat16()
        w = (torch.randn(n, ky * kx * C, device=dev) * 0.05).bfloat16()
        y = torch.empty(B * OHW * OHW, n, device=dev, dtype=torch.bfloat16)
        dy = torch.randn(B, OHW, OHW, n, device=dev).bfloat16()
        dx = torch.empty(B * H * W, C, device=dev, dtype=torch.bfloat16)
        gw = torch.zeros(n, ky * kx * C, device=dev)
        fl = 2.0 * B * OHW * OHW * n * ky * kx * C
        for op, fn in (("fwd", lambda: k.conv_forward(x, w, y, g, act=4)),
                       ("dx", lambda: k.conv_err_input(dy, w, dx, g)),
                       ("dw", lambda: k.conv_weight_grad(x, dy.view(-1, n), gw, g))):
            row = dict(layer=name, op=op, gflop=round(fl / 1e9, 2))
            for m in MASKS:
                lib().vb_gemm_set_ablation(m)
                try:
                    row["m%d_us" % m] = round(timeit(fn, iters=5, warm=2) * 1e3, 1)
                finally:
                    lib().vb_gemm_set_ablation(0)
                print("  %s %s mask %d: %.1f us" % (name, op, m, row["m%d_us" % m]), flush=True)
            print(json.dumps(row), flush=True)
            res.append(row)
            json.dump(res, open("gpurun_out/ablate_conv.json", "w"), indent=1)
