"""Device-vs-fp32-reference numerics for every sm_100a kernel (pattern of the reference's
GPU-vs-numpy oracle tests: veles/tests/test_mean_disp_normalizer.py:79-98,
veles/tests/test_ocl_blas.py:58-109, veles/tests/test_random.py:120-148)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def K():
    from veles_b200.ops import kernels
    return kernels


def dev():
    return torch.device("cuda:0")


def rel_err(a, b):
    a = a.float()
    b = b.float()
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


GEMM_SHAPES = [
    (128, 128, 64), (128, 128, 256), (100, 100, 784), (256, 10, 100), (17, 231, 1999), (1, 8, 8),
    (300, 4096, 512), (1024, 1000, 4096), (7777, 24, 216), (513, 257, 129 * 8),
]


@pytest.mark.parametrize("M,N,K_", GEMM_SHAPES)
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_layouts(M, N, K_, a_mn, b_mn):
    k = K()
    torch.manual_seed(M * 7 + N * 3 + K_)
    pad = lambda v: (v + 7) // 8 * 8
    A = torch.randn(M, K_, device=dev())
    B = torch.randn(N, K_, device=dev())
    a_src = A.t().contiguous() if a_mn else A
    b_src = B.t().contiguous() if b_mn else B
    a = k.pitched_empty(a_src.shape[0], a_src.shape[1], torch.bfloat16, dev())
    b = k.pitched_empty(b_src.shape[0], b_src.shape[1], torch.bfloat16, dev())
    a.copy_(a_src)
    b.copy_(b_src)
    ref = (a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t())
    out = k.pitched_empty(M, N, torch.float32, dev(), align=4)
    out.fill_(float("nan"))
    k.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    assert torch.isfinite(out).all()
    assert rel_err(out, ref) < 2e-3, (M, N, K_, a_mn, b_mn)


@pytest.mark.parametrize("block_n", [64, 128, 192, 256])
def test_gemm_block_n_and_bf16_out(block_n):
    k = K()
    torch.manual_seed(block_n)
    M, N, K_ = 384, 520, 328
    a = torch.randn(M, K_, device=dev()).bfloat16()
    b = torch.randn(N, K_, device=dev()).bfloat16()
    out = torch.empty(M, N, device=dev(), dtype=torch.bfloat16)
    k.gemm(a, b, out, block_n=block_n)
    ref = a.float() @ b.float().t()
    assert rel_err(out, ref) < 1e-2


@pytest.mark.parametrize("act", [0, 1, 2, 3, 4])
def test_gemm_bias_activation_epilogue(act):
    k = K()
    torch.manual_seed(act)
    M, N, K_ = 200, 136, 192
    a = (torch.randn(M, K_, device=dev()) * 0.2).bfloat16()
    b = (torch.randn(N, K_, device=dev()) * 0.2).bfloat16()
    bias = torch.randn(N, device=dev())
    out = torch.empty(M, N, device=dev(), dtype=torch.float32)
    k.gemm(a, b, out, bias=bias, act=act)
    s = a.float() @ b.float().t() + bias
    ref = [s, 1.7159 * torch.tanh(0.6666 * s), torch.sigmoid(s), torch.nn.functional.softplus(s), torch.relu(s)][act]
    assert rel_err(out, ref) < 2e-3


@pytest.mark.parametrize("deriv", [1, 2, 3, 4])
def test_gemm_fused_activation_backward(deriv):
    k = K()
    torch.manual_seed(10 + deriv)
    M, N, K_ = 130, 264, 96
    a = (torch.randn(M, K_, device=dev()) * 0.3).bfloat16()
    b = (torch.randn(K_, N, device=dev()) * 0.3).bfloat16()  # MN-major B: dX = dY * W
    z = torch.rand(M, N, device=dev()).bfloat16()
    out = torch.empty(M, N, device=dev(), dtype=torch.bfloat16)
    k.gemm(a, b, out, b_mn=True, z=z, deriv=deriv)
    zf = z.float()
    d = {1: 0.6666 * (1.7159 - zf * zf / 1.7159), 2: zf * (1 - zf), 3: 1 - torch.exp(-zf), 4: (zf > 0).float()}[deriv]
    ref = (a.float() @ b.float()) * d
    assert rel_err(out, ref) < 1e-2


def test_gemm_split_k_accumulate():
    k = K()
    torch.manual_seed(5)
    M, N, K_ = 96, 200, 20000  # dW-like: tiny output, huge reduction
    a = (torch.randn(K_, M, device=dev()) * 0.1).bfloat16()
    b = (torch.randn(K_, N, device=dev()) * 0.1).bfloat16()
    out = torch.zeros(M, N, device=dev(), dtype=torch.float32)
    k.gemm(a, b, out, a_mn=True, b_mn=True, accumulate=True, split_k=32)
    ref = a.float().t() @ b.float()
    assert rel_err(out, ref) < 2e-3
    k.gemm(a, b, out, a_mn=True, b_mn=True, accumulate=True, split_k=7, alpha=-1.0)
    torch.cuda.synchronize()
    assert out.abs().max().item() < 2e-2 * ref.abs().max().item()


def test_gemm_column_slice_output_grouped_conv_style():
    k = K()
    torch.manual_seed(6)
    M, Kg, Ng = 500, 200, 64
    cols = torch.randn(M, 2 * Kg, device=dev()).bfloat16()
    w = torch.randn(2 * Ng, Kg, device=dev()).bfloat16()
    y = torch.zeros(M, 2 * Ng, device=dev(), dtype=torch.bfloat16)
    bias = torch.randn(2 * Ng, device=dev())
    for g in range(2):
        k.gemm(cols[:, g * Kg:(g + 1) * Kg], w[g * Ng:(g + 1) * Ng], y[:, g * Ng:(g + 1) * Ng],
               bias=bias[g * Ng:(g + 1) * Ng], act=4)
    ref = torch.cat([torch.relu(cols[:, g * Kg:(g + 1) * Kg].float() @ w[g * Ng:(g + 1) * Ng].float().t()
                                + bias[g * Ng:(g + 1) * Ng]) for g in range(2)], dim=1)
    assert rel_err(y, ref) < 1e-2


@pytest.mark.parametrize("direct", [False, True])
@pytest.mark.parametrize("out_dtype", [torch.float32, torch.bfloat16])
def test_gemm_epilogue_paths_respect_tile_tails(direct, out_dtype):
    """TMA store / reduce-add epilogue vs the per-thread fallback: same numbers, nothing written outside D."""
    k = K()
    torch.manual_seed(21)
    M, N, K_ = 333, 200, 264
    a = (torch.randn(M, K_, device=dev()) * 0.2).bfloat16()
    b = (torch.randn(N, K_, device=dev()) * 0.2).bfloat16()
    bias = torch.randn(N, device=dev())
    big = torch.full((M + 64, N + 56), float("nan"), device=dev(), dtype=out_dtype)
    out = big[:M, :N]
    k.set_direct_epilogue(direct)
    try:
        k.gemm(a, b, out, bias=bias, act=1)
        ref = 1.7159 * torch.tanh(0.6666 * (a.float() @ b.float().t() + bias))
        assert rel_err(out, ref) < (2e-3 if out_dtype == torch.float32 else 1e-2)
        assert torch.isnan(big[M:].float()).all() and torch.isnan(big[:, N:].float()).all()
        if out_dtype == torch.float32:
            out.zero_()
            k.gemm(a, b, out, accumulate=True)
            k.gemm(a, b, out, accumulate=True, split_k=3)
            assert rel_err(out, 2 * (a.float() @ b.float().t())) < 2e-3
            assert torch.isnan(big[M:]).all() and torch.isnan(big[:, N:]).all()
    finally:
        k.set_direct_epilogue(False)


def test_gemm_unaligned_output_uses_fallback():
    k = K()
    torch.manual_seed(22)
    M, N, K_ = 100, 30, 64
    a = torch.randn(M, K_, device=dev()).bfloat16()
    b = torch.randn(N, K_, device=dev()).bfloat16()
    out = torch.empty(M, N + 1, device=dev(), dtype=torch.bfloat16)[:, 1:]   # 2-byte aligned base, odd pitch
    k.gemm(a, b, out)
    assert rel_err(out, a.float() @ b.float().t()) < 1e-2


CONV_CASES = [
    # B, H, W, C, n, ky, kx, sy, sx, pt, pl, pb, pr, groups
    (4, 13, 13, 256, 384, 3, 3, 1, 1, 1, 1, 1, 1, 1),     # AlexNet conv3
    (3, 13, 13, 384, 256, 3, 3, 1, 1, 1, 1, 1, 1, 2),     # AlexNet conv5 (grouped, 192 channels per group)
    (2, 27, 27, 96, 256, 5, 5, 1, 1, 2, 2, 2, 2, 2),      # AlexNet conv2 (48 channels per group: padded K blocks)
    (5, 9, 11, 16, 40, 3, 2, 1, 1, 0, 1, 2, 0, 1),        # asymmetric padding, odd sizes, tiny channels
    (3, 17, 15, 64, 72, 3, 3, 2, 2, 1, 1, 1, 1, 1),       # strided (forward / weights only)
    (2, 23, 23, 8, 24, 5, 5, 3, 2, 0, 0, 0, 0, 1),        # mixed strides
]


def _conv_ref(x, w, n, ky, kx, sy, sx, pt, pl, pb, pr, G):
    """x [B,H,W,C] fp32, w [n, ky*kx*Cg] with (tap row, tap col, channel) columns -> y [B,OH,OW,n]."""
    B, H, W, C = x.shape
    Cg = C // G
    wt = w.view(n, ky, kx, Cg).permute(0, 3, 1, 2).contiguous()
    xp = torch.nn.functional.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    return torch.nn.functional.conv2d(xp, wt, stride=(sy, sx), groups=G).permute(0, 2, 3, 1).contiguous()


@pytest.fixture(params=[1, 2], ids=["single_cta", "cta_pair"])
def pair_mode(request):
    """Runs a GEMM test with single-CTA tiles and again with cta_group::2 CTA-pair tiles forced."""
    k = K()
    k.set_pair_mode(request.param)
    yield request.param
    k.set_pair_mode(0)


@pytest.mark.parametrize("M,N,K_", [(256, 512, 384), (300, 136, 100), (1028, 64, 36), (640, 1000, 4096)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_tf32_layouts(M, N, K_, a_mn, b_mn, pair_mode):
    """fp32 operands through tcgen05 kind::tf32 vs an fp64 reference: ~1e-3 (10-bit mantissa products), far tighter than bf16."""
    k = K()
    torch.manual_seed(M + N)
    a = torch.randn((K_, M) if a_mn else (M, K_), device=dev()) * 0.2
    b = torch.randn((K_, N) if b_mn else (N, K_), device=dev()) * 0.2
    bias = torch.randn(N, device=dev())
    A = a.double().t() if a_mn else a.double()
    B = b.double() if b_mn else b.double().t()
    ref = torch.sigmoid(A @ B + bias.double())
    big = torch.full((M + 8, N + 8), float("nan"), device=dev())
    k.gemm_tf32(a, b, big[:M, :N], a_mn=a_mn, b_mn=b_mn, bias=bias, act=2)
    assert rel_err(big[:M, :N], ref) < 3e-3
    assert torch.isnan(big[M:]).all() and torch.isnan(big[:, N:]).all()
    out = torch.zeros(M, N, device=dev())
    k.gemm_tf32(a, b, out, a_mn=a_mn, b_mn=b_mn, accumulate=True, split_k=0)
    z = torch.rand(M, N, device=dev())
    out2 = torch.empty(M, N, device=dev())
    k.gemm_tf32(a, b, out2, a_mn=a_mn, b_mn=b_mn, z=z, deriv=2)
    lin = A @ B
    assert rel_err(out, lin) < 3e-3
    assert rel_err(out2, lin * (z * (1 - z)).double()) < 3e-3
    # and it IS tighter than the bf16 path on the same data
    if not a_mn and not b_mn and K_ % 8 == 0:
        o16 = torch.empty(M, N, device=dev())
        k.gemm(a.bfloat16(), b.bfloat16(), o16)
        assert rel_err(out, lin) < rel_err(o16, lin)


@pytest.fixture(params=[False, True], ids=["static_sched", "clc_sched"])
def dyn_sched(request):
    """Static persistent round-robin vs cluster-launch-control work stealing."""
    k = K()
    k.set_dynamic_schedule(request.param)
    yield request.param
    k.set_dynamic_schedule(False)


def test_gemm_dynamic_schedule_many_tiles(pair_mode):
    """Far more tiles than SMs, so almost every tile is obtained through clusterlaunchcontrol.try_cancel."""
    k = K()
    k.set_dynamic_schedule(True)
    try:
        torch.manual_seed(3)
        M, N, K_ = 128 * 70, 128 * 9, 192
        a = (torch.randn(M, K_, device=dev()) * 0.2).bfloat16()
        b = (torch.randn(N, K_, device=dev()) * 0.2).bfloat16()
        out = torch.full((M, N), float("nan"), device=dev(), dtype=torch.bfloat16)
        for _ in range(3):
            k.gemm(a, b, out, block_n=128)
        torch.cuda.synchronize()
        assert rel_err(out, a.float() @ b.float().t()) < 1e-2
    finally:
        k.set_dynamic_schedule(False)


@pytest.mark.parametrize("M,N,K_", [(1024, 512, 320), (304, 264, 200), (776, 136, 1032), (4104, 96, 128)])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("block_n", [0, 128, 256])
def test_gemm_cta_pair_layouts(M, N, K_, a_mn, b_mn, block_n, pair_mode, dyn_sched):
    k = K()
    torch.manual_seed(M + N + K_)
    a = torch.randn((K_, M) if a_mn else (M, K_), device=dev()).bfloat16()
    b = torch.randn((K_, N) if b_mn else (N, K_), device=dev()).bfloat16()
    bias = torch.randn(N, device=dev())
    ref = torch.relu((a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t()) + bias)
    big = torch.full((M + 40, N + 24), float("nan"), device=dev(), dtype=torch.bfloat16)
    k.gemm(a, b, big[:M, :N], a_mn=a_mn, b_mn=b_mn, bias=bias, act=4, block_n=block_n)
    assert rel_err(big[:M, :N], ref) < 1e-2
    assert torch.isnan(big[M:].float()).all() and torch.isnan(big[:, N:].float()).all()
    out = torch.zeros(M, N, device=dev())
    k.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, accumulate=True, split_k=0, block_n=block_n)
    k.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, accumulate=True, split_k=2, block_n=block_n)
    assert rel_err(out, 2 * ((a.float().t() if a_mn else a.float()) @ (b.float() if b_mn else b.float().t()))) < 3e-3


HALO_CASES = [
    # B, H, W, C, n, ky, kx, pt, pl, pb, pr, groups
    (3, 27, 27, 96, 256, 5, 5, 2, 2, 2, 2, 2),     # AlexNet conv2: P = 32, 48 channels / 128 kernels per group
    (4, 13, 13, 256, 384, 3, 3, 1, 1, 1, 1, 1),    # conv3: P = 16, two image rows per epilogue warp, 3 N tiles
    (2, 13, 13, 384, 256, 3, 3, 1, 1, 1, 1, 2),    # conv5 (grouped)
    (2, 55, 55, 48, 96, 3, 3, 0, 0, 2, 2, 1),      # conv1 after space-to-depth: P = 64
    (3, 9, 11, 16, 40, 3, 2, 0, 1, 2, 0, 1),       # asymmetric padding, odd sizes, N tail
    (2, 30, 7, 8, 24, 7, 3, 3, 1, 3, 1, 1),        # tall filter: many halo rows
]


@pytest.fixture(params=[(0, -1), (1, 0), (2, 0), (1, 1)], ids=["auto", "mt1_stream", "mt2_stream", "mt1_resident"])
def halo_options(request):
    k = K()
    k.conv_halo_set_options(*request.param)
    yield request.param
    k.conv_halo_set_options(0, -1)


@pytest.mark.parametrize("case", HALO_CASES)
def test_halo_tile_conv_forward_and_err_input(case, halo_options):
    """Halo-tile kernel (all taps from one smem tile, shifted UMMA descriptors, 4-D TMA boxes; one or two sub-tiles
    per weight tile, streamed or resident weights) vs torch conv2d."""
    k = K()
    B, H, W, C, n, ky, kx, pt, pl, pb, pr, G = case
    torch.manual_seed(sum(case))
    Cg, ng = C // G, n // G
    OH, OW = H + pt + pb - ky + 1, W + pl + pr - kx + 1
    g = dict(OH=OH, OW=OW, ky=ky, kx=kx, sy=1, sx=1, pt=pt, pl=pl, H=H, W=W)
    x = (torch.randn(B, H, W, C, device=dev()) * 0.5).bfloat16()
    w = (torch.randn(n, ky * kx * Cg, device=dev()) * 0.1).bfloat16()
    bias = torch.randn(n, device=dev())
    xr = x.float().requires_grad_(True)
    yr = torch.relu(_conv_ref(xr, w.float(), n, ky, kx, 1, 1, pt, pl, pb, pr, G) + bias)
    y = torch.full((B * OH * OW + 5, n), float("nan"), device=dev(), dtype=torch.bfloat16)
    for gi in range(G):
        assert k.conv_forward_halo(x[..., gi * Cg:(gi + 1) * Cg], w[gi * ng:(gi + 1) * ng],
                                   y[:B * OH * OW, gi * ng:(gi + 1) * ng], g, bias=bias[gi * ng:(gi + 1) * ng], act=4)
    assert rel_err(y[:B * OH * OW], yr.detach().reshape(-1, n)) < 1e-2
    assert torch.isnan(y[B * OH * OW:].float()).all()
    e = (torch.randn(B, OH, OW, n, device=dev()) * 0.3).bfloat16()
    yl = _conv_ref(xr, w.float(), n, ky, kx, 1, 1, pt, pl, pb, pr, G)
    yl.backward(e.float())
    dx = torch.full((B * H * W + 3, C), float("nan"), device=dev(), dtype=torch.bfloat16)
    for gi in range(G):
        assert k.conv_err_input_halo(e[..., gi * ng:(gi + 1) * ng], w[gi * ng:(gi + 1) * ng],
                                     dx[:B * H * W, gi * Cg:(gi + 1) * Cg], g,
                                     z=x.reshape(-1, C)[:, gi * Cg:(gi + 1) * Cg], deriv=1)
    ref = xr.grad * (0.6666 * (1.7159 - x.float() ** 2 / 1.7159))
    assert rel_err(dx[:B * H * W], ref.reshape(-1, C)) < 1e-2
    assert torch.isnan(dx[B * H * W:].float()).all()


@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("world", [2, 8])
def test_gemm_row_sharded_weight_operand(b_mn, world, pair_mode):
    """Peer-TMA weight delivery on one GPU: replica r holds valid data only in the rows rank r owns (NaN elsewhere);
    the GEMM must pick every weight tile from its owner.  Forward (K-major B) and err_input (MN-major B) forms."""
    k = K()
    torch.manual_seed(world)
    rows, inner = 1024, 320          # weight matrix [rows, inner]; shard = rows / world
    shard = rows // world
    w = (torch.randn(rows, inner, device=dev()) * 0.2).bfloat16()
    reps = []
    for r in range(world):
        t = torch.full_like(w, float("nan"))
        t[r * shard:(r + 1) * shard] = w[r * shard:(r + 1) * shard]
        reps.append(t)
    peers = ([t.data_ptr() for t in reps], shard, 1)
    if not b_mn:      # Y = X W^T
        x = (torch.randn(300, inner, device=dev()) * 0.2).bfloat16()
        out = torch.empty(300, rows, device=dev(), dtype=torch.bfloat16)
        k.gemm(x, reps[1], out, b_peers=peers)
        ref = x.float() @ w.float().t()
    else:             # dX = dY W
        dy = (torch.randn(300, rows, device=dev()) * 0.2).bfloat16()
        out = torch.empty(300, inner, device=dev(), dtype=torch.bfloat16)
        k.gemm(dy, reps[1], out, b_mn=True, b_peers=peers)
        ref = dy.float() @ w.float()
    assert torch.isfinite(out.float()).all()
    assert rel_err(out, ref) < 1e-2


@pytest.mark.parametrize("case", CONV_CASES)
def test_implicit_gemm_conv_forward_backward(case, pair_mode, dyn_sched):
    """TMA-im2col implicit GEMM (forward, err_input, weight gradient) vs torch conv2d autograd in fp32."""
    k = K()
    B, H, W, C, n, ky, kx, sy, sx, pt, pl, pb, pr, G = case
    torch.manual_seed(sum(case))
    Cg, ng = C // G, n // G
    OH = (H + pt + pb - ky) // sy + 1
    OW = (W + pl + pr - kx) // sx + 1
    g = dict(OH=OH, OW=OW, ky=ky, kx=kx, sy=sy, sx=sx, pt=pt, pl=pl, H=H, W=W)
    x = (torch.randn(B, H, W, C, device=dev()) * 0.5).bfloat16()
    w = (torch.randn(n, ky * kx * Cg, device=dev()) * 0.1).bfloat16()
    bias = torch.randn(n, device=dev())
    xr = x.float().requires_grad_(True)
    wr = w.float().requires_grad_(True)
    yr = torch.tanh(0.6666 * (_conv_ref(xr, wr, n, ky, kx, sy, sx, pt, pl, pb, pr, G) + bias)) * 1.7159
    dy = (torch.randn_like(yr) * 0.3).bfloat16()
    yr.backward(dy.float())
    # forward
    y = torch.full((B * OH * OW + 3, n), float("nan"), device=dev(), dtype=torch.bfloat16)
    for gi in range(G):
        k.conv_forward(x[..., gi * Cg:(gi + 1) * Cg], w[gi * ng:(gi + 1) * ng], y[:B * OH * OW, gi * ng:(gi + 1) * ng], g,
                       bias=bias[gi * ng:(gi + 1) * ng], act=1)
    assert rel_err(y[:B * OH * OW], yr.detach().reshape(-1, n)) < 1e-2
    assert torch.isnan(y[B * OH * OW:].float()).all()
    # weight gradient of the pre-activation error e = dy * f'(y)
    e = (dy.float() * (0.6666 * (1.7159 - yr.detach() ** 2 / 1.7159))).bfloat16()
    yr2 = _conv_ref(xr, wr, n, ky, kx, sy, sx, pt, pl, pb, pr, G)
    xr.grad = None
    wr.grad = None
    yr2.backward(e.float())
    e2 = e.reshape(-1, n)
    gw = torch.zeros(n, ky * kx * Cg, device=dev())
    for gi in range(G):
        k.conv_weight_grad(x[..., gi * Cg:(gi + 1) * Cg], e2[:, gi * ng:(gi + 1) * ng], gw[gi * ng:(gi + 1) * ng], g)
    assert rel_err(gw, wr.grad) < 5e-3
    for gi in range(G):   # accumulate: a second pass with alpha=-1 cancels
        k.conv_weight_grad(x[..., gi * Cg:(gi + 1) * Cg], e2[:, gi * ng:(gi + 1) * ng], gw[gi * ng:(gi + 1) * ng], g,
                           alpha=-1.0, split_k=3)
    assert gw.abs().max().item() < 2e-2 * wr.grad.abs().max().item()
    # err_input (unit stride), fused with the previous layer's tanh derivative through z = x
    if sx == 1 and sy == 1:
        dx = torch.full((B * H * W, C), float("nan"), device=dev(), dtype=torch.bfloat16)
        for gi in range(G):
            k.conv_err_input(e[..., gi * ng:(gi + 1) * ng], w[gi * ng:(gi + 1) * ng], dx[:, gi * Cg:(gi + 1) * Cg], g,
                             z=x.reshape(-1, C)[:, gi * Cg:(gi + 1) * Cg], deriv=1)
        ref = xr.grad * (0.6666 * (1.7159 - x.float() ** 2 / 1.7159))
        assert rel_err(dx, ref.reshape(-1, C)) < 1e-2


@pytest.mark.parametrize("src_dtype", [torch.uint8, torch.float32])
@pytest.mark.parametrize("dst_dtype", [torch.float32, torch.bfloat16])
def test_gather_minibatch(src_dtype, dst_dtype):
    k = K()
    g = torch.Generator(device="cpu").manual_seed(3)
    n, size, mb, count = 1000, 3 * 32 * 32 + 3, 128, 100
    if src_dtype == torch.uint8:
        src = torch.randint(0, 256, (n, size), generator=g, dtype=torch.uint8).to(dev())
    else:
        src = torch.randn(n, size, generator=g).to(dev())
    labels = torch.randint(0, 10, (n,), generator=g, dtype=torch.int32).to(dev())
    idx = torch.randperm(n, generator=g).int().to(dev())
    mean = torch.randn(size, generator=g).to(dev())
    rdisp = (torch.rand(size, generator=g) + 0.5).to(dev())
    dst = torch.full((mb + mb, size), float("nan"), device=dev(), dtype=dst_dtype)  # guard band (doubling_reset pattern)
    mbl = torch.full((mb,), 77, device=dev(), dtype=torch.int32)
    mbi = torch.full((mb,), 77, device=dev(), dtype=torch.int32)
    start = 250
    k.gather_minibatch(src, idx, start, count, dst[:mb], mean, rdisp, labels, mbl, mbi)
    sel = idx[start:start + count].long()
    ref = (src[sel].float() - mean) * rdisp
    tol = 1e-5 if dst_dtype == torch.float32 else 1e-2
    assert rel_err(dst[:count], ref) < tol
    assert (dst[count:mb] == 0).all()
    assert torch.isnan(dst[mb:].float()).all(), "kernel wrote past the minibatch"
    assert (mbl[:count] == labels[sel]).all() and (mbl[count:] == -1).all()
    assert (mbi[:count] == idx[start:start + count]).all() and (mbi[count:] == -1).all()


@pytest.mark.parametrize("M,N,KK", [(256, 512, 128), (1000, 4096, 256), (100, 784, 64), (130, 70, 100), (64, 33, 40)])
@pytest.mark.parametrize("moment,wd,l12", [(0.0, 0.0, 0.0), (0.9, 0.0005, 0.0), (0.5, 0.01, 0.3)])
def test_gemm_with_fused_gd_epilogue(M, N, KK, moment, wd, l12):
    """dW = A^T B never leaves the GEMM: its epilogue applies the momentum step to w / velocity and refreshes the bf16
    shadow.  Reference: the same GEMM in fp32 + the update rule of vb_gd_update."""
    k = K()
    g = torch.Generator(device="cpu").manual_seed(M + N + KK)
    a = k.pitched_empty(KK, M, torch.bfloat16, dev())          # err_output: [batch, n_out], read MN-major
    b = k.pitched_empty(KK, N, torch.bfloat16, dev())          # input:      [batch, n_in]
    a.copy_(torch.randn(KK, M, generator=g).to(dev()) * 0.1)
    b.copy_(torch.randn(KK, N, generator=g).to(dev()))
    w = (torch.randn(M, N, generator=g) * 0.1).to(dev())
    w[0, :min(N, 5)] = 0.0                                   # sign(0) = 0 in the L1 term
    vel = (torch.randn(M, N, generator=g) * 0.01).to(dev())
    shadow = k.pitched_empty(M, N, torch.bfloat16, dev())
    lr = 0.05
    grad = a.float().t() @ b.float()
    gref = grad + wd * ((1 - l12) * w + 0.5 * l12 * torch.sign(w))
    vref = moment * vel - lr * gref
    wref = w + vref
    before = k.LAUNCHES[0]
    k.gemm_fused_gd(a, b, w, vel, shadow, lr=lr, wd=wd, l1_vs_l2=l12, moment=moment)
    assert k.LAUNCHES[0] == before + 1
    scale = vref.abs().max().item() + 1e-9
    assert (vel - vref).abs().max().item() < 2e-3 * scale + 1e-6
    assert (w - wref).abs().max().item() < 2e-3 * scale + 1e-6
    assert torch.equal(shadow, w.bfloat16())
    pair_w, pair_v = w.clone(), vel.clone()
    # second step from the updated state, CTA-pair tiles forced when the shape allows them
    k.set_pair_mode(2)
    try:
        k.gemm_fused_gd(a, b, w, vel, shadow, lr=lr, wd=wd, l1_vs_l2=l12, moment=moment)
    finally:
        k.set_pair_mode(0)
    g2 = grad + wd * ((1 - l12) * pair_w + 0.5 * l12 * torch.sign(pair_w))
    v2 = moment * pair_v - lr * g2
    assert (vel - v2).abs().max().item() < 2e-3 * (v2.abs().max().item() + 1e-9) + 1e-6
    assert (w - (pair_w + v2)).abs().max().item() < 2e-3 * (v2.abs().max().item() + 1e-9) + 1e-6


@pytest.mark.parametrize("rows,cols,dtype", [(256, 1000, torch.bfloat16), (43264, 384, torch.bfloat16),
                                             (100, 10, torch.float32), (777, 96, torch.bfloat16)])
def test_col_sum_with_fused_bias_update(rows, cols, dtype):
    k = K()
    g = torch.Generator(device="cpu").manual_seed(rows + cols)
    a = k.pitched_empty(rows, cols, dtype, dev(), align=8 if dtype == torch.bfloat16 else 4)
    a.copy_(torch.randn(rows, cols, generator=g).to(dev()) * 0.1)
    w = torch.randn(cols, generator=g).to(dev())
    vel = (torch.randn(cols, generator=g) * 0.01).to(dev())
    grad = (torch.randn(cols, generator=g) * 0.01).to(dev())       # something already accumulated
    scratch = torch.empty(k.col_sum_scratch_elems(rows, cols), dtype=torch.float32, device=dev())
    lr, wd, l12, mom = 0.1, 0.01, 0.25, 0.9
    gref = a.float().sum(0) + grad
    gref = gref + wd * ((1 - l12) * w + 0.5 * l12 * torch.sign(w))
    vref = mom * vel - lr * gref
    wref = w + vref
    k.col_sum_gd(a, grad, scratch, w, vel, lr=lr, wd=wd, l1_vs_l2=l12, moment=mom)
    tol = 1e-4 * (a.float().abs().sum(0).max().item() + 1.0)
    assert (vel - vref).abs().max().item() < tol and (w - wref).abs().max().item() < tol
    assert (grad == 0).all()
    # and the plain column sum still agrees with the reference through the re-shaped second phase
    out = torch.zeros(cols, device=dev())
    k.col_sum(a, out, scratch)
    assert (out - a.float().sum(0)).abs().max().item() < tol


def test_gather_u8_to_bf16_is_exact_for_odd_row_sizes():
    """The flat 8-element gather path: odd row sizes (unaligned rows), partial minibatch, no normalisation."""
    k = K()
    g = torch.Generator(device="cpu").manual_seed(5)
    for size, mb, count in ((227 * 3 * 5 + 1, 16, 16), (4099, 7, 5), (17, 33, 32), (16, 8, 8)):
        n = 50
        src = torch.randint(0, 256, (n, size), generator=g, dtype=torch.uint8).to(dev())
        idx = torch.randperm(n, generator=g).int().to(dev())
        dst = torch.full((mb + 3, size), float("nan"), device=dev(), dtype=torch.bfloat16)
        k.gather_minibatch(src, idx, 3, count, dst[:mb], None, None, None, None, None)
        sel = idx[3:3 + count].long()
        assert torch.equal(dst[:count].float(), src[sel].float()), size
        assert (dst[count:mb] == 0).all()
        assert torch.isnan(dst[mb:].float()).all(), "kernel wrote past the minibatch"


@pytest.mark.parametrize("src_dtype", [torch.uint8, torch.float32])
@pytest.mark.parametrize("H,W,C,sy,sx", [(227, 227, 3, 4, 4), (32, 30, 2, 2, 2), (17, 9, 1, 4, 2), (12, 12, 4, 1, 2),
                                         (10, 7, 3, 2, 4), (9, 11, 6, 4, 2), (33, 250, 3, 4, 4)])
def test_gather_fused_with_space_to_depth(src_dtype, H, W, C, sy, sx):
    """``vb_gather_s2d`` against a plain fp32 restatement: gather by index, normalise, then x[b, h2*sy+dy, w2*sx+dx, c]
    -> xs[b, h2, w2, (dy*sx+dx)*C + c] with zeros past the image edge and for rows past ``count``."""
    k = K()
    g = torch.Generator(device="cpu").manual_seed(H + W + C)
    n, mb, count, start = 40, 19, 13, 5
    size = H * W * C
    if src_dtype == torch.uint8:
        src = torch.randint(0, 256, (n, size), generator=g, dtype=torch.uint8).to(dev())
    else:
        src = torch.randn(n, size, generator=g).to(dev())
    labels = torch.randint(0, 10, (n,), generator=g, dtype=torch.int32).to(dev())
    idx = torch.randperm(n, generator=g).int().to(dev())
    mean = torch.randn(size, generator=g).to(dev())
    rdisp = (torch.rand(size, generator=g) + 0.5).to(dev())
    H2, W2, C2 = -(-H // sy), -(-W // sx), sy * sx * C
    xs = torch.full((mb + 2, H2, W2, C2), float("nan"), device=dev(), dtype=torch.bfloat16)
    mbl = torch.full((mb,), 77, device=dev(), dtype=torch.int32)
    mbi = torch.full((mb,), 77, device=dev(), dtype=torch.int32)
    for m_, r_ in ((mean, rdisp), (None, None)):
        assert k.gather_s2d(src, idx, start, count, xs[:mb], H, W, C, sy, sx, m_, r_, labels, mbl, mbi)
        sel = idx[start:start + count].long()
        x = src[sel].float()
        if m_ is not None:
            x = (x - mean) * rdisp
        x = x.reshape(count, H, W, C)
        pad = torch.zeros(mb, H2 * sy, W2 * sx, C, device=dev())
        pad[:count, :H, :W] = x
        ref = pad.reshape(mb, H2, sy, W2, sx, C).permute(0, 1, 3, 2, 4, 5).reshape(mb, H2, W2, C2)
        got = xs[:mb].float()
        assert (got - ref).abs().max().item() <= 1e-2 * max(1.0, ref.abs().max().item())
        assert (got[count:] == 0).all() and torch.equal(got == 0, got == 0)
        assert (got[ref == 0] == 0).all()                       # padding positions are exact zeros
        assert torch.isnan(xs[mb:].float()).all(), "kernel wrote past the minibatch"
        assert (mbl[:count] == labels[sel]).all() and (mbl[count:] == -1).all()
        assert (mbi[:count] == idx[start:start + count]).all() and (mbi[count:] == -1).all()
    # the two-kernel path it replaces gives the same bits for bf16-exact inputs
    if src_dtype == torch.uint8:
        nhwc = torch.zeros(mb, size, device=dev(), dtype=torch.bfloat16)
        k.gather_minibatch(src, idx, start, count, nhwc, None, None, None, None, None)
        xs2 = torch.zeros(mb, H2, W2, C2, device=dev(), dtype=torch.bfloat16)
        k.space_to_depth(nhwc.view(mb, H, W, C), xs2, sy, sx)
        assert torch.equal(xs2, xs[:mb])


def test_gather_s2d_refuses_uncovered_geometry():
    k = K()
    src = torch.zeros(4, 40 * 400 * 3, dtype=torch.uint8, device=dev())
    idx = torch.arange(4, dtype=torch.int32, device=dev())
    xs = torch.zeros(4, 10, 100, 48, device=dev(), dtype=torch.bfloat16)
    assert k.gather_s2d(src, idx, 0, 4, xs, 40, 400, 3, 4, 4) is False      # strip of 4 x 400 x 3 > 3072 elements


def test_mean_disp_normalizer():
    k = K()
    g = torch.Generator(device="cpu").manual_seed(4)
    inp = torch.randint(0, 256, (37, 3, 21, 20), generator=g, dtype=torch.uint8).to(dev())
    mean = torch.rand(3 * 21 * 20, generator=g).to(dev()) * 255
    rdisp = (torch.rand(3 * 21 * 20, generator=g) + 0.1).to(dev())
    out = torch.empty(37, 3 * 21 * 20, device=dev())
    k.mean_disp_normalize(inp, mean, rdisp, out)
    ref = (inp.reshape(37, -1).float() - mean) * rdisp
    assert (out - ref).abs().max().item() < 1e-4


def test_copy_2d_joiner():
    k = K()
    a = torch.randn(50, 30, device=dev())
    b = torch.randn(50, 17, device=dev())
    out = torch.zeros(50, 47, device=dev())
    k.copy_2d(a, out[:, :30])
    k.copy_2d(b, out[:, 30:])
    assert torch.equal(out, torch.cat([a, b], dim=1))


@pytest.mark.parametrize("act", list(range(1, 9)))
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_activations(act, dtype):
    k = K()
    torch.manual_seed(act)
    x = (torch.randn(1000, 33, device=dev()) * 2).to(dtype)
    y = torch.empty_like(x)
    factor = 0.37
    k.activation_forward(act, x, y, factor)
    xd = x.double().cpu().requires_grad_(True)
    idx = torch.arange(xd.numel()).reshape(xd.shape) & 1

    def f(v):
        if act == 1: return 1.7159 * torch.tanh(0.6666 * v)
        if act == 2: return torch.sigmoid(v)
        if act == 3: return torch.nn.functional.softplus(v)
        if act == 4: return torch.relu(v)
        if act == 5: return torch.log(v + torch.sqrt(v * v + 1))
        if act == 6:
            d, a, b = 3.0, 0.242528761112, 305.459953195
            inner = 1.7159 * torch.tanh(0.6666 * v)
            return torch.where(v > d, torch.log(v.clamp(min=1e-9) * b) * a,
                               torch.where(v < -d, -torch.log((-v).clamp(min=1e-9) * b) * a, inner))
        if act == 7: return torch.where(idx == 1, torch.sin(v), torch.cos(v))
        if act == 8: return v * factor
    ref = f(xd)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert rel_err(y.cpu(), ref.detach()) < tol
    eo = torch.randn_like(x)
    ei = torch.empty_like(x)
    k.activation_backward(act, eo, x, y, ei, factor)
    (ref * eo.double().cpu()).sum().backward()
    tolb = 2e-4 if dtype == torch.float32 else 6e-2
    assert rel_err(ei.cpu(), xd.grad) < tolb


def test_dropout_reproducible_mask_and_scale():
    k = K()
    torch.manual_seed(1)
    n_states = 4096
    st = torch.randint(1, 2 ** 62, (n_states, 2), dtype=torch.int64, device=dev())
    st2 = torch.empty_like(st)
    x = torch.randn(1001, 333, device=dev())
    y = torch.empty_like(x)
    k.dropout(x, y, st, st2, 0.5)
    kept = (y != 0)
    frac = kept.float().mean().item()
    assert 0.48 < frac < 0.52
    assert torch.allclose(y[kept], x[kept] * 2.0)
    # backward with the saved states regenerates the identical mask
    e = torch.ones_like(x)
    ei = torch.empty_like(x)
    k.dropout(e, ei, st, None, 0.5)
    assert torch.equal(ei != 0, kept)
    assert not torch.equal(st, st2)


@pytest.mark.parametrize("solver", [0, 1, 2])
def test_gd_update(solver):
    k = K()
    torch.manual_seed(solver)
    n = 100003
    w = torch.randn(n, device=dev())
    g = torch.randn(n, device=dev())
    v = torch.randn(n, device=dev()) * 0.1 if solver != 1 else torch.zeros(n, device=dev())
    if solver == 2:
        v = v.abs()
    acc = torch.rand(n, device=dev())
    wb = torch.empty(n, device=dev(), dtype=torch.bfloat16)
    w0, g0, v0, a0 = w.clone(), g.clone(), v.clone(), acc.clone()
    lr, wd, l12, mom, gs, eps, rho = 0.03, 0.0005, 0.3, 0.9, 0.5, 1e-6, 0.95
    k.gd_update(w, g, v, acc, wb, lr=lr, wd=wd, l1_vs_l2=l12, moment=mom, grad_scale=gs, solver=solver, eps=eps,
                rho=rho, zero_grad=True)
    gg = g0 * gs + wd * ((1 - l12) * w0 + 0.5 * l12 * torch.sign(w0))
    if solver == 0:
        vr = mom * v0 - lr * gg
        wr = w0 + vr
        assert torch.allclose(v, vr, atol=1e-6)
    elif solver == 1:
        ar = a0 + gg * gg
        wr = w0 - lr * gg / torch.sqrt(ar + eps)
        assert torch.allclose(acc, ar, atol=1e-5)
    else:
        ar = rho * a0 + (1 - rho) * gg * gg
        d = -torch.sqrt(v0 + eps) / torch.sqrt(ar + eps) * gg
        wr = w0 + d
    assert torch.allclose(w, wr, atol=1e-4, rtol=1e-4)
    assert (g == 0).all()
    assert torch.allclose(wb.float(), w, atol=1e-2, rtol=1e-2)


def _conv_geom(H, W, ky, kx, sy, sx, pt, pl, pb, pr, groups=1):
    OH = (H + pt + pb - ky) // sy + 1
    OW = (W + pl + pr - kx) // sx + 1
    return dict(OH=OH, OW=OW, ky=ky, kx=kx, sy=sy, sx=sx, pt=pt, pl=pl, groups=groups)


@pytest.mark.parametrize("C,groups,k_,s_,p_", [(3, 1, 5, 1, 2), (16, 1, 3, 1, 1), (32, 2, 5, 2, 2), (3, 1, 11, 4, 0),
                                               (1, 1, 3, 2, 1), (4, 1, 4, 3, 0), (6, 2, 3, 1, 1)])
def test_im2col_col2im_match_unfold(C, groups, k_, s_, p_):
    k = K()
    torch.manual_seed(C)
    B, H, W = 3, 19, 23
    x = torch.randn(B, H, W, C, device=dev()).bfloat16()
    geom = _conv_geom(H, W, k_, k_, s_, s_, p_, p_, p_, p_, groups)
    rows = B * geom["OH"] * geom["OW"]
    kc = k_ * k_ * C
    pitch = (kc + 7) // 8 * 8
    full = torch.full((rows, pitch), float("nan"), device=dev(), dtype=torch.bfloat16)
    cols = full[:, :kc]
    k.im2col(x, cols, geom)
    # oracle: unfold on NCHW, reorder to (g, tap, cg)
    un = torch.nn.functional.unfold(x.float().permute(0, 3, 1, 2), k_, padding=p_, stride=s_)  # [B, C*k*k, L]
    Cg = C // groups
    un = un.reshape(B, groups, Cg, k_ * k_, -1).permute(0, 4, 1, 3, 2).reshape(rows, kc)
    assert torch.equal(cols.float(), un)
    assert (full[:, kc:] == 0).all(), "K tail must be zero-filled"
    # col2im == fold (adjoint of unfold)
    dcols = torch.randn(rows, kc, device=dev()).bfloat16()
    dfull = torch.zeros_like(full)
    dfull[:, :kc] = dcols
    dx = torch.empty_like(x)
    k.col2im(dfull[:, :kc], dx, geom)
    dun = dcols.float().reshape(B, -1, groups, k_ * k_, Cg).permute(0, 2, 4, 3, 1).reshape(B, C * k_ * k_, -1)
    ref = torch.nn.functional.fold(dun, (H, W), k_, padding=p_, stride=s_).permute(0, 2, 3, 1)
    assert rel_err(dx, ref) < 1e-2


@pytest.mark.parametrize("kind", ["max", "maxabs", "avg"])
@pytest.mark.parametrize("kx,sx", [(2, 2), (3, 2)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_pooling(kind, kx, sx, dtype):
    k = K()
    torch.manual_seed(kx)
    B, H, W, C = 2, 13, 14, 24
    x = torch.randn(B, H, W, C, device=dev()).to(dtype)
    OH = (H - kx + sx - 1) // sx + 1 if (H - kx) % sx else (H - kx) // sx + 1
    OW = (W - kx + sx - 1) // sx + 1 if (W - kx) % sx else (W - kx) // sx + 1
    y = torch.empty(B, OH, OW, C, device=dev(), dtype=dtype)
    offs = torch.empty(B, OH, OW, C, device=dev(), dtype=torch.int32)
    kid = k.POOL_KINDS[kind]
    k.pool_forward(kid, x, y, offs, kx, kx, sx, sx)
    xf = x.float()
    ref = torch.empty(B, OH, OW, C, device=dev())
    dy = torch.randn(B, OH, OW, C, device=dev()).to(dtype)
    dref = torch.zeros(B, H, W, C, device=dev())
    for oh in range(OH):
        for ow in range(OW):
            win = xf[:, oh * sx:min(oh * sx + kx, H), ow * sx:min(ow * sx + kx, W), :]
            wh, ww = win.shape[1], win.shape[2]
            flat = win.reshape(B, wh * ww, C)
            if kind == "avg":
                ref[:, oh, ow] = flat.mean(1)
                dref[:, oh * sx:oh * sx + wh, ow * sx:ow * sx + ww, :] += (dy[:, oh, ow].float() / (wh * ww))[:, None, None, :]
            else:
                key = flat.abs() if kind == "maxabs" else flat
                am = key.argmax(1)
                ref[:, oh, ow] = flat.gather(1, am[:, None, :])[:, 0]
                hh = oh * sx + am // ww
                wwi = ow * sx + am % ww
                for b in range(B):
                    dref[b, hh[b], wwi[b], torch.arange(C)] += dy[b, oh, ow].float()
    assert rel_err(y, ref) < (1e-6 if dtype == torch.float32 else 8e-3)
    dx = torch.empty_like(x)
    k.pool_backward(kid, dy, offs, dx, kx, kx, sx, sx)
    assert rel_err(dx, dref) < (1e-5 if dtype == torch.float32 else 2e-2)


def test_stochastic_pooling_picks_window_member():
    k = K()
    torch.manual_seed(0)
    B, H, W, C = 2, 8, 8, 16
    x = torch.rand(B, H, W, C, device=dev())
    y = torch.empty(B, 4, 4, C, device=dev())
    offs = torch.empty(B, 4, 4, C, device=dev(), dtype=torch.int32)
    rnd = torch.randint(0, 65536, (B, 4, 4, C), device=dev(), dtype=torch.int32).to(torch.int16)
    k.pool_forward(3, x, y, offs, 2, 2, 2, 2, rnd)
    assert torch.equal(x.flatten()[offs.long().flatten()], y.flatten())
    mx = torch.nn.functional.max_pool2d(x.permute(0, 3, 1, 2), 2).permute(0, 2, 3, 1)
    assert (y <= mx + 1e-6).all() and (y < mx - 1e-6).any()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_lrn_forward_backward(dtype):
    k = K()
    torch.manual_seed(2)
    B, H, W, C = 2, 5, 6, 32
    n, alpha, beta, kk = 5, 1e-2, 0.75, 2.0
    x = torch.randn(B, H, W, C, device=dev()).to(dtype)
    y = torch.empty_like(x)
    k.lrn_forward(x, y, n, alpha, beta, kk)
    xr = x.double().requires_grad_(True)
    sq = torch.nn.functional.pad(xr * xr, (n // 2, n // 2))
    s = kk + alpha * sum(sq[..., i:i + C] for i in range(n))
    ref = xr * s ** (-beta)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert rel_err(y, ref.detach()) < tol
    dy = torch.randn_like(x)
    dx = torch.empty_like(x)
    k.lrn_backward(x, dy, dx, n, alpha, beta, kk)
    (ref * dy.double()).sum().backward()
    assert rel_err(dx, xr.grad) < (1e-4 if dtype == torch.float32 else 2e-2)


@pytest.mark.parametrize("C,n", [(96, 5), (256, 5), (8, 3), (40, 9), (24, 7), (20, 5), (64, 11)])
def test_lrn_vector_paths(C, n):
    """bf16 LRN: the 8-channel register kernels (n <= 9, C % 8 == 0) and the row fallback, large alpha so the
    window term matters; fused previous-layer strict-ReLU derivative in the backward."""
    k = K()
    torch.manual_seed(C + n)
    B, H, W = 3, 7, 5
    alpha, beta, kk = 0.3, (0.75 if C != 40 else 0.6), 1.5   # 0.75 takes the rsqrt path, anything else exp2/log2
    x = torch.randn(B, H, W, C, device=dev()).bfloat16()
    y = torch.empty_like(x)
    k.lrn_forward(x, y, n, alpha, beta, kk)
    xr = x.double().requires_grad_(True)
    sq = torch.nn.functional.pad(xr * xr, (n // 2, n // 2))
    s = kk + alpha * sum(sq[..., i:i + C] for i in range(n))
    ref = xr * s ** (-beta)
    assert rel_err(y, ref.detach()) < 1e-2
    dy = torch.randn_like(x)
    dx = torch.empty_like(x)
    k.lrn_backward(x, dy, dx, n, alpha, beta, kk, deriv=4)
    (ref * dy.double()).sum().backward()
    assert rel_err(dx, xr.grad * (x > 0).double()) < 2e-2


@pytest.mark.parametrize("B,H,W,C,n,deriv", [(3, 13, 13, 96, 5, 4), (8, 27, 27, 256, 5, 4), (2, 8, 9, 40, 3, 0),
                                              (1, 7, 7, 8, 9, 4), (64, 20, 20, 96, 5, 4)])
def test_lrn_backward_accumulates_bias_column_sums(B, H, W, C, n, deriv):
    """LRN backward with ``partials``: same dx as without, plus per-block partial column sums of dx that
    ``col_sum_finish`` folds into the bias gradient of the convolution in front (plain accumulate, and with the fused
    momentum step), against an fp32 sum of the stored dx."""
    k = K()
    torch.manual_seed(B + H + C)
    alpha, beta, kk = 0.3, 0.75, 1.5
    x = torch.randn(B, H, W, C, device=dev()).bfloat16()
    dy = torch.randn_like(x)
    dx_ref = torch.empty_like(x)
    k.lrn_backward(x, dy, dx_ref, n, alpha, beta, kk, deriv=deriv)
    dx = torch.full((B * H * W + 2, C), float("nan"), device=dev(), dtype=torch.bfloat16)
    for rows in (1184, 3):                         # few rows: the grid shrinks to what the scratch holds
        partials = torch.full((rows + 1, C), float("nan"), device=dev())
        slabs = k.lrn_backward(x, dy, dx[:B * H * W].view_as(x), n, alpha, beta, kk, deriv=deriv,
                               partials=partials[:rows])
        assert slabs is not None and 1 <= slabs <= rows
        assert torch.equal(dx[:B * H * W].view_as(x), dx_ref)
        assert torch.isnan(dx[B * H * W:].float()).all() and torch.isnan(partials[slabs:]).all()
        colsum = dx_ref.float().reshape(-1, C).sum(0)
        # the kernel sums the fp32 values before they are rounded to bf16: allow the rounding of B*H*W terms
        tol = 2e-3 * dx_ref.float().abs().reshape(-1, C).sum(0).max().item() + 1e-3
        grad = torch.ones(C, device=dev())
        k.col_sum_finish(partials, slabs, grad)
        assert (grad - (1.0 + colsum)).abs().max().item() < tol
        w = torch.randn(C, device=dev())
        vel = torch.randn(C, device=dev()) * 0.1
        w0, v0 = w.clone(), vel.clone()
        grad = torch.full((C,), 0.5, device=dev())
        k.col_sum_finish(partials, slabs, grad, w, vel, lr=0.01, wd=0.1, l1_vs_l2=0.0, moment=0.9)
        g = 0.5 + colsum + 0.1 * w0
        v1 = 0.9 * v0 - 0.01 * g
        assert (vel - v1).abs().max().item() < 0.01 * tol + 1e-6 and (w - (w0 + v1)).abs().max().item() < 0.01 * tol + 1e-6
        assert (grad == 0).all()
    # shapes without a summing kernel report it and launch nothing
    xw = torch.randn(2, 3, 3, 264, device=dev()).bfloat16()
    assert k.lrn_backward(xw, torch.randn_like(xw), torch.empty_like(xw), 5, alpha, beta, kk,
                          partials=torch.zeros(8, 264, device=dev())) is None


@pytest.mark.parametrize("B,H,W,C,n", [(3, 13, 13, 96, 5), (2, 27, 27, 256, 5), (2, 8, 9, 40, 3), (1, 7, 7, 8, 9)])
def test_pool_backward_fused_into_lrn_backward(B, H, W, C, n):
    """One kernel = max-pooling backward (3x3 / 2, byte argmax) followed by LRN backward with the strict-ReLU
    derivative, vs the two separate kernels (each of which is checked against autograd on its own)."""
    k = K()
    torch.manual_seed(B * H + C)
    alpha, beta, kk = 0.3, 0.75, 1.5
    x = torch.randn(B, H, W, C, device=dev()).bfloat16()
    y = torch.empty_like(x)
    k.lrn_forward(x, y, n, alpha, beta, kk)
    OH, OW = (H - 3 + 1) // 2 + 1, (W - 3 + 1) // 2 + 1          # ceil mode, as the pooling units compute it
    p = torch.empty(B, OH, OW, C, device=dev(), dtype=torch.bfloat16)
    offs8 = torch.zeros(B, OH, OW, C, device=dev(), dtype=torch.uint8)
    assert k.pool_forward_local(k.POOL_KINDS["max"], y, p, None, offs8, 3, 3, 2, 2)
    dp = torch.randn_like(p)
    # reference: the two kernels in sequence
    dy = torch.zeros_like(x)
    assert k.pool_backward_local(k.POOL_KINDS["max"], dp, offs8, dy, 3, 3, 2, 2)
    dx_ref = torch.empty_like(x)
    k.lrn_backward(x, dy, dx_ref, n, alpha, beta, kk, deriv=4)
    dx = torch.full((B * H * W + 2, C), float("nan"), device=dev(), dtype=torch.bfloat16)
    assert k.lrn_pool_backward(x, dp, offs8, dx[:B * H * W].view(B, H, W, C), 3, 3, 2, 2, n, alpha, beta, kk, deriv=4)
    assert rel_err(dx[:B * H * W].view_as(x), dx_ref) < 2e-2
    assert torch.isnan(dx[B * H * W:].float()).all()


@pytest.mark.parametrize("N", [10, 1000])
@pytest.mark.parametrize("in_dtype", [torch.float32, torch.bfloat16])
def test_softmax_evaluator(N, in_dtype):
    k = K()
    torch.manual_seed(N)
    B, valid = 100, 91
    s = (torch.randn(B, N, device=dev()) * 3).to(in_dtype)
    labels = torch.randint(0, N, (B,), device=dev(), dtype=torch.int32)
    labels[valid:] = -1
    y = torch.empty(B, N, device=dev())
    err = torch.empty(B, N, device=dev())
    max_idx = torch.empty(B, device=dev(), dtype=torch.int32)
    n_err = torch.zeros(2, device=dev(), dtype=torch.int32)
    conf = torch.zeros(N, N, device=dev(), dtype=torch.int32)
    mes = torch.zeros(1, device=dev())
    loss = torch.zeros(1, device=dev())
    k.softmax_eval(s, y, err, labels, max_idx, err_scale=1.0 / valid, n_err=n_err, confusion=conf, max_err_sum=mes,
                   loss_sum=loss)
    p = torch.softmax(s.float(), 1)
    assert rel_err(y, p) < 1e-5
    am = p.argmax(1)
    assert torch.equal(max_idx.long(), am)
    oh = torch.zeros(B, N, device=dev())
    oh[torch.arange(valid), labels[:valid].long()] = 1
    eref = (p - oh) / valid
    eref[valid:] = 0
    assert (err - eref).abs().max().item() < 1e-6
    assert n_err[0].item() == (am[:valid] != labels[:valid].long()).sum().item()
    assert n_err[1].item() == valid
    cref = torch.zeros(N, N, device=dev(), dtype=torch.int32)
    for i in range(valid):
        cref[am[i], labels[i].long()] += 1
    assert torch.equal(conf, cref)
    assert abs(mes.item() - eref.abs().sum(1).max().item()) < 1e-6
    lref = -torch.log(p[torch.arange(valid), labels[:valid].long()]).sum().item()
    assert abs(loss.item() - lref) < 1e-3 * max(1.0, abs(lref))


def test_mse_evaluator():
    k = K()
    torch.manual_seed(9)
    B, valid, N = 64, 50, 77
    y = torch.randn(B, N, device=dev())
    t = torch.randn(B, N, device=dev())
    err = torch.empty(B, N, device=dev())
    metrics = torch.tensor([0.0, 0.0, 1e30], device=dev())
    per = torch.empty(B, device=dev())
    k.mse_eval(y, t, err, valid, err_scale=0.5, metrics=metrics, mse_per_sample=per)
    e = (y - t)
    e[valid:] = 0
    rm = torch.sqrt((e * e).mean(1))
    assert torch.allclose(err, e * 0.5, atol=1e-6)
    assert abs(metrics[0].item() - rm[:valid].sum().item()) < 1e-3
    assert abs(metrics[1].item() - rm[:valid].max().item()) < 1e-6
    assert abs(metrics[2].item() - rm[:valid].min().item()) < 1e-6


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_matrix_reduce(dtype):
    k = K()
    torch.manual_seed(3)
    a = torch.randn(777, 203, device=dev()).to(dtype)
    rs = torch.empty(777, device=dev())
    k.row_sum(a, rs)
    assert rel_err(rs, a.float().sum(1)) < 1e-5
    cs = torch.ones(203, device=dev())
    scratch = torch.empty(k.col_sum_scratch_elems(777, 203), device=dev())
    k.col_sum(a, cs, scratch, alpha=2.0, accumulate=True)
    assert rel_err(cs, 1 + 2 * a.float().sum(0)) < 1e-5
    k.col_sum(a, cs, scratch)
    assert rel_err(cs, a.float().sum(0)) < 1e-5


def _xs1024_oracle(states, rounds):
    """Pure-numpy xorshift1024* with the reference's output interleaving."""
    S = states.shape[0]
    st = states.copy()
    out = np.zeros(rounds * 16 * S, dtype=np.uint64)
    mult = np.uint64(1181783497276652981)
    with np.errstate(over="ignore"):
        for r in range(rounds):
            for i in range(16):
                s0 = st[:, i].copy()
                s1 = st[:, (i + 1) & 15].copy()
                s1 ^= s1 << np.uint64(31)
                s1 ^= s1 >> np.uint64(11)
                s0 ^= s0 >> np.uint64(30)
                v = s0 ^ s1
                st[:, (i + 1) & 15] = v
                out[(r * 16 + i) * S:(r * 16 + i + 1) * S] = v * mult
    return out, st


def test_xorshift1024star_bit_exact():
    k = K()
    rng = np.random.RandomState(7)
    S, rounds = 256, 3
    states = rng.randint(0, 2 ** 63, size=(S, 16), dtype=np.int64).astype(np.uint64)
    ref, st_ref = _xs1024_oracle(states, rounds)
    d_states = torch.from_numpy(states.view(np.int64).copy()).to(dev())
    out = torch.empty(rounds * 16 * S, dtype=torch.int64, device=dev())
    k.xorshift1024star(d_states, rounds, out)
    assert np.array_equal(out.cpu().numpy().view(np.uint64), ref)
    assert np.array_equal(d_states.cpu().numpy().view(np.uint64), st_ref)


def test_xorshift128plus_bit_exact():
    k = K()
    rng = np.random.RandomState(8)
    n = 5000
    st = rng.randint(1, 2 ** 63, size=(n, 2), dtype=np.int64).astype(np.uint64)
    d = torch.from_numpy(st.view(np.int64).copy()).to(dev())
    out = torch.empty(n, dtype=torch.int64, device=dev())
    k.xorshift128plus(d, out)
    with np.errstate(over="ignore"):
        s1 = st[:, 0].copy()
        s0 = st[:, 1].copy()
        s1 ^= s1 << np.uint64(23)
        ny = s1 ^ s0 ^ (s1 >> np.uint64(17)) ^ (s0 >> np.uint64(26))
        ref = ny + s0
    assert np.array_equal(out.cpu().numpy().view(np.uint64), ref)
    back = d.cpu().numpy().view(np.uint64)
    assert np.array_equal(back[:, 0], s0) and np.array_equal(back[:, 1], ny)


def test_fill_random_moments():
    k = K()
    st = torch.randint(1, 2 ** 62, (2048, 2), dtype=torch.int64, device=dev())
    out = torch.empty(1_000_001, device=dev())
    k.fill_random(st, out, 0, 0.05)
    assert out.abs().max().item() <= 0.05 and abs(out.mean().item()) < 1e-3
    assert abs(out.std().item() - 0.05 / 3 ** 0.5) < 1e-3
    k.fill_random(st, out, 1, 0.01)
    assert abs(out.mean().item()) < 1e-4 and abs(out.std().item() - 0.01) < 2e-4


def test_two_phase_split_k_bias_activation_pass():
    """Split-K GEMM into an fp32 scratch followed by vb_bias_act_2d equals the fused epilogue; the scratch is re-zeroed."""
    k = K()
    torch.manual_seed(9)
    M, N, K_ = 256, 520, 4096
    a = (torch.randn(M, K_, device=dev()) * 0.05).bfloat16()
    b = (torch.randn(N, K_, device=dev()) * 0.05).bfloat16()
    bias = torch.randn(N, device=dev())
    z = torch.rand(M, N, device=dev()).bfloat16()
    acc = torch.zeros(M, N, device=dev())
    out = torch.empty(M, N, device=dev(), dtype=torch.bfloat16)
    assert k.gemm_plan(M, N, K_)[1] >= 2
    k.gemm(a, b, acc, accumulate=True, split_k=0)
    k.bias_act(acc, out, bias=bias, act=1, z=z, deriv=2)
    ref = 1.7159 * torch.tanh(0.6666 * (a.float() @ b.float().t() + bias)) * (z.float() * (1 - z.float()))
    assert rel_err(out, ref) < 1e-2
    assert (acc == 0).all()


@pytest.mark.parametrize("C,n,G", [(384, 384, 2), (384, 256, 2), (256, 512, 4)])
def test_grouped_conv_single_launch(C, n, G, pair_mode):
    """All groups of a grouped convolution in one implicit-GEMM launch (tile -> group) vs torch conv2d."""
    k = K()
    torch.manual_seed(C + n + G)
    B, H, W, ky, kx, p = 3, 13, 13, 3, 3, 1
    Cg = C // G
    assert k.grouped_conv_ok(C, n, G)
    g = dict(OH=H, OW=W, ky=ky, kx=kx, sy=1, sx=1, pt=p, pl=p, H=H, W=W)
    x = (torch.randn(B, H, W, C, device=dev()) * 0.5).bfloat16()
    w = (torch.randn(n, ky * kx * Cg, device=dev()) * 0.1).bfloat16()
    bias = torch.randn(n, device=dev())
    xr = x.float().requires_grad_(True)
    yr = torch.relu(_conv_ref(xr, w.float(), n, ky, kx, 1, 1, p, p, p, p, G) + bias)
    y = torch.full((B * H * W + 2, n), float("nan"), device=dev(), dtype=torch.bfloat16)
    k.conv_forward_grouped(x, w, y[:B * H * W], g, G, bias=bias, act=4)
    assert rel_err(y[:B * H * W], yr.detach().reshape(-1, n)) < 1e-2
    assert torch.isnan(y[B * H * W:].float()).all()
    e = (torch.randn(B, H, W, n, device=dev()) * 0.3).bfloat16()
    _conv_ref(xr, w.float(), n, ky, kx, 1, 1, p, p, p, p, G).backward(e.float())
    dx = torch.full((B * H * W + 2, C), float("nan"), device=dev(), dtype=torch.bfloat16)
    k.conv_err_input_grouped(e, w, dx[:B * H * W], g, G, z=x.reshape(-1, C), deriv=1)
    ref = xr.grad * (0.6666 * (1.7159 - x.float() ** 2 / 1.7159))
    assert rel_err(dx[:B * H * W], ref.reshape(-1, C)) < 1e-2
    assert torch.isnan(dx[B * H * W:].float()).all()
