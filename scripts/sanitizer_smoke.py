"""Every kernel family once on small, awkward shapes - the program compute-sanitizer is pointed at:

    compute-sanitizer --tool memcheck  python scripts/sanitizer_smoke.py     (out-of-bounds / misaligned accesses)
    compute-sanitizer --tool racecheck python scripts/sanitizer_smoke.py     (shared-memory hazards)
    compute-sanitizer --tool synccheck python scripts/sanitizer_smoke.py     (barrier misuse)

One tool per gpurun call; copy the tail of the log into profiles/sanitizer_<tool>.txt.  Successor of the
reference's only runtime guards (NaN-tail overrun detector tests/doubling_reset.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from veles_b200.ops import kernels as k

dev = torch.device("cuda:0")


def bf(*shape, scale=0.3):
    return (torch.randn(*shape, device=dev) * scale).bfloat16()


def main():
    torch.manual_seed(0)
    # GEMM family: all operand layouts, tails in M/N/K, both tile kinds, split-K, fused epilogues
    for pair in (1, 2):
        k.set_pair_mode(pair)
        for a_mn, b_mn in ((False, False), (False, True), (True, True), (True, False)):
            M, N, K = 264, 136, 200
            a = bf(K, M) if a_mn else bf(M, K)
            b = bf(K, N) if b_mn else bf(N, K)
            out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
            k.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, bias=torch.randn(N, device=dev), act=1)
            acc = torch.zeros(M, N, device=dev)
            k.gemm(a, b, acc, a_mn=a_mn, b_mn=b_mn, accumulate=True, split_k=3)
        z = bf(264, 136)
        k.gemm(bf(264, 64), bf(64, 136), torch.empty(264, 136, device=dev, dtype=torch.bfloat16), b_mn=True, z=z, deriv=2)
        # implicit-GEMM convolution: forward / err_input / weight gradient, grouped + padded K blocks
        B, H, W, C, n, ky, kx, p = 2, 9, 7, 48, 40, 3, 3, 1
        g = dict(OH=H, OW=W, ky=ky, kx=kx, sy=1, sx=1, pt=p, pl=p, H=H, W=W)
        x, w = bf(B, H, W, C), bf(n, ky * kx * C)
        y = torch.empty(B * H * W, n, device=dev, dtype=torch.bfloat16)
        k.conv_forward(x, w, y, g, bias=torch.randn(n, device=dev), act=4)
        dx = torch.empty(B * H * W, C, device=dev, dtype=torch.bfloat16)
        k.conv_err_input(y.view(B, H, W, n), w, dx, g, z=x.view(-1, C), deriv=4)
        k.conv_weight_grad(x, y, torch.zeros(n, ky * kx * C, device=dev), g)
    k.set_pair_mode(0)
    k.set_direct_epilogue(True)
    k.gemm(bf(100, 64), bf(30, 64), torch.empty(100, 31, device=dev, dtype=torch.bfloat16)[:, 1:])
    k.set_direct_epilogue(False)
    # memory-bound kernels
    x = bf(3, 11, 9, 3)
    xs = torch.empty(3, 3, 3, 48, device=dev, dtype=torch.bfloat16)
    k.space_to_depth(x, xs, 4, 4)
    geom = dict(OH=5, OW=3, ky=3, kx=3, sy=2, sx=3, pt=0, pl=0, groups=1)
    cols = torch.zeros(3 * 5 * 3, 32, device=dev, dtype=torch.bfloat16)
    k.im2col(x, cols[:, :27], geom)
    k.col2im(cols[:, :27], torch.empty_like(x), geom)
    x = bf(2, 7, 9, 16)
    y = torch.empty(2, 3, 4, 16, device=dev, dtype=torch.bfloat16)
    offs = torch.zeros(2, 3, 4, 16, device=dev, dtype=torch.int32)
    offs8 = torch.zeros(2, 3, 4, 16, device=dev, dtype=torch.uint8)
    k.pool_forward(0, x, y, offs, 3, 3, 2, 2)
    k.pool_forward_local(0, x, y, None, offs8, 3, 3, 2, 2)
    k.pool_expand_offsets(offs8, offs, x.shape, 3, 3, 2, 2)
    k.pool_backward(0, y, offs, torch.empty_like(x), 3, 3, 2, 2, z=x, deriv=4)
    k.pool_backward_local(0, y, offs8, torch.empty_like(x), 3, 3, 2, 2, z=x, deriv=4)
    x = bf(2, 5, 5, 24)
    k.lrn_forward(x, torch.empty_like(x), 5, 1e-2, 0.75, 2.0)
    k.lrn_backward(x, bf(2, 5, 5, 24), torch.empty_like(x), 5, 1e-2, 0.75, 2.0, deriv=4)
    a = bf(1000, 96)
    scratch = torch.empty(k.col_sum_scratch_elems(1000, 96), device=dev)
    k.col_sum(a, torch.zeros(96, device=dev), scratch)
    k.row_sum(a, torch.zeros(1000, device=dev))
    logits = torch.randn(33, 10, device=dev)
    labels = torch.randint(0, 10, (33,), device=dev, dtype=torch.int32)
    k.softmax_eval(logits, torch.empty(33, 10, device=dev, dtype=torch.bfloat16),
                   torch.empty(33, 10, device=dev, dtype=torch.bfloat16), labels,
                   torch.zeros(33, device=dev, dtype=torch.int32), n_err=torch.zeros(2, device=dev, dtype=torch.int32))
    w = torch.randn(1003, device=dev)
    k.gd_update(w[:1000], torch.randn(1000, device=dev), torch.zeros(1000, device=dev), None,
                torch.empty(1000, device=dev, dtype=torch.bfloat16), lr=0.1, wd=1e-3, l1_vs_l2=0.5, moment=0.9)
    st = torch.randint(1, 1 << 40, (256 * 2,), device=dev, dtype=torch.int64)
    xin = bf(64, 100)
    k.dropout(xin, torch.empty_like(xin), st, st.clone(), 0.5)
    torch.cuda.synchronize()
    print("SANITIZER_SMOKE_DONE")


if __name__ == "__main__":
    main()
