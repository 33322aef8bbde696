"""Micro-benchmark of the implicit-GEMM convolution kernels on the AlexNet layer shapes (per group), sweeping the tile
width.  CUDA-event timed, L2 flushed between iterations."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from veles_b200.ops import kernels as k
from bench_gemm import timeit, dev

LAYERS = {   # name: (B, H, W, Cg, n_g, ky, kx, pad)
    "conv1s2d": (256, 57, 57, 48, 96, 3, 3, 0),
    "conv2": (256, 27, 27, 48, 128, 5, 5, 2),
    "conv3": (256, 13, 13, 256, 384, 3, 3, 1),
    "conv4": (256, 13, 13, 192, 192, 3, 3, 1),
    "conv5": (256, 13, 13, 192, 128, 3, 3, 1),
}

if __name__ == "__main__":
    res = []
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
        for bn in (0,):
            for pair in (0,):
                k.set_pair_mode(pair)
                row = dict(layer=name, block_n=bn, pair=pair == 2)
                for op, fn in (("fwd", lambda: k.conv_forward(x, w, y, g, act=4, block_n=bn)),
                               ("dx", lambda: k.conv_err_input(dy, w, dx, g, block_n=bn)),
                               ("dw", lambda: k.conv_weight_grad(x, dy.view(-1, n), gw, g, block_n=bn))):
                    try:
                        t = timeit(fn, iters=5, warm=2)
                        row[op + "_us"] = round(t * 1e3, 1)
                        row[op + "_tflops"] = round(fl / t / 1e9, 1)
                    except Exception as e:   # tile width not available for this operand layout
                        row[op + "_us"] = None
                print(json.dumps(row), flush=True)
                res.append(row)
        k.set_pair_mode(0)
        for mt, resident in ((0, -1), (1, 0), (2, 0)):
            k.conv_halo_set_options(mt, resident)
            row = dict(layer=name, kernel="halo", mt=mt, resident=resident)
            for op, fn in (("fwd", lambda: k.conv_forward_halo(x, w, y, g, act=4)),
                           ("dx", lambda: k.conv_err_input_halo(dy, w, dx, g))):
                if fn():
                    t = timeit(fn, iters=5, warm=2)
                    row[op + "_us"] = round(t * 1e3, 1)
                    row[op + "_tflops"] = round(fl / t / 1e9, 1)
            print(json.dumps(row), flush=True)
            res.append(row)
        k.conv_halo_set_options(0, -1)
    k.set_pair_mode(0)
    out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    os.makedirs(out_dir, exist_ok=True)
    json.dump(res, open(os.path.join(out_dir, "bench_conv.json"), "w"), indent=1)
