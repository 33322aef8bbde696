"""One launch of each convolution kernel variant on one AlexNet layer shape, for `ncu --set full`:
    ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16|conv_halo" -o gpurun_out/conv_prof \
        python scripts/profile_conv.py conv2
Launch order per convolution layer named on the command line: im2col forward, err_input, weight gradient, halo
forward, halo err_input; "fc" = AlexNet fc6 forward, err_input, weight gradient (plain GEMMs)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
from veles_b200.ops import kernels as k
from bench_conv import LAYERS

dev = torch.device("cuda:0")


def conv_layer(name):
    B, H, W, C, n, ky, kx, p = LAYERS[name]
    OHW = H + 2 * p - ky + 1
    g = dict(OH=OHW, OW=OHW, ky=ky, kx=kx, sy=1, sx=1, pt=p, pl=p, H=H, W=W)
    x = torch.randn(B, H, W, C, device=dev).bfloat16()
    w = (torch.randn(n, ky * kx * C, device=dev) * 0.05).bfloat16()
    y = torch.empty(B * OHW * OHW, n, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(B, OHW, OHW, n, device=dev).bfloat16()
    dx = torch.empty(B * H * W, C, device=dev, dtype=torch.bfloat16)
    gw = torch.zeros(n, ky * kx * C, device=dev)
    k.conv_forward(x, w, y, g, act=4)
    k.conv_err_input(dy, w, dx, g)
    k.conv_weight_grad(x, dy.view(-1, n), gw, g)
    k.conv_forward_halo(x, w, y, g, act=4)
    k.conv_err_input_halo(dy, w, dx, g)


def fc_layer(B=256, n_in=9216, n_out=4096):
    x = torch.randn(B, n_in, device=dev).bfloat16()
    w = (torch.randn(n_out, n_in, device=dev) * 0.01).bfloat16()
    y = torch.empty(B, n_out, device=dev, dtype=torch.bfloat16)
    dy = torch.randn(B, n_out, device=dev).bfloat16()
    dx = torch.empty(B, n_in, device=dev, dtype=torch.bfloat16)
    gw = torch.zeros(n_out, n_in, device=dev)
    k.gemm(x, w, y, act=4)                                  # forward
    k.gemm(dy, w, dx, b_mn=True)                            # err_input
    k.gemm(dy, x, gw, a_mn=True, b_mn=True, split_k=1)      # weight gradient


for name in (sys.argv[1:] or ["conv2"]):
    if name == "fc":
        fc_layer()
    else:
        conv_layer(name)
torch.cuda.synchronize()
print("done")
