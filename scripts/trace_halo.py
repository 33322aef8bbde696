"""Where does a halo-tile convolution CTA spend its time?  CTA 0 stamps clock64 at the hand-over points of its first
24 tiles (vb_conv_halo_set_trace); this prints the per-tile intervals in cycles."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from ctypes import c_void_p
from veles_b200.ops import kernels as k
from veles_b200.ops._lib import lib
from bench_conv import LAYERS

dev = torch.device("cuda:0")
for name in (sys.argv[1:] or ["conv1s2d", "conv2"]):
    B, H, W, C, n, ky, kx, p = LAYERS[name]
    OHW = H + 2 * p - ky + 1
    g = dict(OH=OHW, OW=OHW, ky=ky, kx=kx, sy=1, sx=1, pt=p, pl=p, H=H, W=W)
    x = torch.randn(B, H, W, C, device=dev).bfloat16()
    w = (torch.randn(n, ky * kx * C, device=dev) * 0.05).bfloat16()
    y = torch.empty(B * OHW * OHW, n, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(B, OHW, OHW, n, device=dev).bfloat16()
    dx = torch.empty(B * H * W, C, device=dev, dtype=torch.bfloat16)
    for op, fn in (("fwd", lambda: k.conv_forward_halo(x, w, y, g, act=4)), ("dx", lambda: k.conv_err_input_halo(dy, w, dx, g))):
        fn()
        torch.cuda.synchronize()
        buf = torch.zeros(24 * 8, dtype=torch.int64, device=dev)
        lib().vb_conv_halo_set_trace(c_void_p(buf.data_ptr()))
        fn()
        torch.cuda.synchronize()
        lib().vb_conv_halo_set_trace(c_void_p(0))
        t = buf.cpu().view(24, 8)
        t0 = int(t[0, 0])
        print("== %s %s: tile | wait acc | wait A | issue | (mma done - issue end) | epi wait | epi work | tile period | A load issued rel. tile start" % (name, op))
        for i in range(2, 14):
            r = [int(v) for v in t[i]]
            if r[3] == 0:
                break
            prev = int(t[i - 1, 0])
            print("%3d | %6d | %6d | %6d | %6d | %6d | %6d | %6d | %6d" % (
                i, r[1] - r[0], r[2] - r[1], r[3] - r[2], r[5] - r[3], r[5] - r[4], r[6] - r[5], r[0] - prev, r[7] - r[0]))
