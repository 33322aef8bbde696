"""Python entry points for the sm_100a kernel library.

Every function takes ``torch`` CUDA tensors, launches on the current torch stream and
returns immediately (no synchronisation).  There is no eager fallback: if the library
cannot be loaded the call raises :class:`veles_b200.ops._lib.KernelError`.

Counts launches in :data:`LAUNCHES` so ``bench.py`` can report ``gpu_launches``.
"""
from __future__ import annotations

import ctypes
import threading
from ctypes import c_float, c_int, c_longlong, c_void_p

import torch

from . import _lib

# activation codes shared with csrc/common.cuh
ACT_LINEAR, ACT_TANH, ACT_SIGMOID, ACT_RELU, ACT_STRICT_RELU, ACT_LOG, ACT_TANHLOG, ACT_SINCOS, ACT_MUL = range(9)
ACT_CODES = {
    "linear": ACT_LINEAR, "tanh": ACT_TANH, "sigmoid": ACT_SIGMOID, "relu": ACT_RELU,
    "strict_relu": ACT_STRICT_RELU, "str": ACT_STRICT_RELU, "log": ACT_LOG, "tanhlog": ACT_TANHLOG,
    "sincos": ACT_SINCOS, "mul": ACT_MUL,
}
POOL_KINDS = {"max": 0, "maxabs": 1, "avg": 2, "stochastic": 3, "stochastic_abs": 4}

LAUNCHES = [0]
_BYTES = [None]   # traffic override of the call being traced (gathers touch a slice of their source)
TRACE = None      # set to a list to record (name, start event, end event, tensor bytes, flops, shapes) per call
_FLOPS = [0.0]


def _dt(t):
    if t.dtype == torch.float32:
        return 0
    if t.dtype == torch.bfloat16:
        return 1
    if t.dtype == torch.uint8:
        return 2
    if t.dtype == torch.int32:
        return 3
    raise TypeError("unsupported dtype %s" % t.dtype)


def _p(t):
    return c_void_p(0 if t is None else t.data_ptr())


_TLS = threading.local()
_BOUND = {"index": None, "generation": 0}    # the device this process computes on; bumped when it changes


def bind_device(index):
    """The device this process computes on (set by B200Device); worker threads bind to it on their first launch."""
    if _BOUND["index"] != int(index):
        _BOUND["index"] = int(index)
        _BOUND["generation"] += 1          # every thread re-binds


def _stream():
    # The C entry points use the calling thread's current CUDA device / context (driver-API tensor-map encoding
    # fails with CUDA_ERROR_INVALID_CONTEXT in a thread that never touched the runtime).  A workflow's thread
    # pool starts threads that may reach a launch before any torch op: bind them once.
    if getattr(_TLS, "generation", -1) != _BOUND["generation"]:
        idx = _BOUND["index"]
        torch.cuda.set_device(torch.cuda.current_device() if idx is None else idx)
        _TLS.generation = _BOUND["generation"]
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _ld(t):
    """Leading dimension (elements) of a 2-D row-major, possibly pitched, tensor."""
    assert t.dim() == 2 and (t.stride(1) == 1 or t.shape[1] == 1), "need unit inner stride, got %s" % (t.stride(),)
    return max(t.stride(0), t.shape[1])


def pitched_empty(rows, cols, dtype, device, align=8):
    """[rows, cols] view whose row pitch is a multiple of ``align`` elements (TMA needs 16 B)."""
    pitch = (cols + align - 1) // align * align
    buf = torch.zeros(rows, pitch, dtype=dtype, device=device)
    return buf[:, :cols]


def set_direct_epilogue(on):
    """Test hook: force the GEMM's per-thread store epilogue instead of the smem-staged TMA store / reduce-add."""
    _lib.check(_lib.lib().vb_gemm_set_direct_epilogue(c_int(int(bool(on)))), "vb_gemm_set_direct_epilogue")


def set_dynamic_schedule(on):
    """GEMM tile scheduling: True = cluster-launch-control work stealing, False = static persistent round-robin."""
    _lib.check(_lib.lib().vb_gemm_set_dynamic_schedule(c_int(int(bool(on)))), "vb_gemm_set_dynamic_schedule")


def set_pdl(on):
    """Programmatic dependent launch for every kernel of the library (csrc/common.cuh ``launch_k``): the next kernel of
    a stream is scheduled while the current one drains and waits at its first global access.  Returns the old value."""
    return bool(_lib.lib().vb_set_pdl(c_int(int(bool(on)))))


def set_pair_mode(mode):
    """Test hook: 0 = the cost model picks, 1 = single-CTA tiles only, 2 = CTA-pair (cta_group::2) tiles whenever
    the shape allows."""
    _lib.check(_lib.lib().vb_gemm_set_pair_mode(c_int(int(mode))), "vb_gemm_set_pair_mode")


def bias_act(acc, out, bias=None, act=ACT_LINEAR, z=None, deriv=0, zero_acc=True):
    """out = act(acc + bias) [* f'(z)]; ``acc`` is the fp32 result of a split-K GEMM and is zeroed again on the way."""
    assert acc.dtype == torch.float32 and acc.shape == out.shape
    rc = _lib.lib().vb_bias_act_2d(_p(acc), c_longlong(_ld(acc)), _p(bias), c_int(act), _p(z),
                                   c_longlong(_ld(z) if z is not None else 0), c_int(deriv), _p(out), c_int(_dt(out)),
                                   c_longlong(_ld(out)), c_int(acc.shape[0]), c_int(acc.shape[1]),
                                   c_int(int(zero_acc)), _stream())
    _lib.check(rc, "vb_bias_act_2d")
    LAUNCHES[0] += 1
    return out


def gemm_plan(M, N, K, b_mn=False, can_split=True, block_n=0):
    """(tile width, split-K factor, uses CTA pairs) the GEMM scheduler picks for this problem."""
    bn, split, pair = c_int(0), c_int(0), c_int(0)
    _lib.check(_lib.lib().vb_gemm_plan(c_int(M), c_int(N), c_int(K), c_int(int(b_mn)), c_int(int(can_split)),
                                       c_int(block_n), ctypes.byref(bn), ctypes.byref(split), ctypes.byref(pair)),
               "vb_gemm_plan")
    return bn.value, split.value, bool(pair.value)


def gemm(a, b, out, *, a_mn=False, b_mn=False, bias=None, act=ACT_LINEAR, accumulate=False, z=None, deriv=0,
         alpha=1.0, split_k=1, block_n=0, b_peers=None):
    """out[M,N] = epilogue(sum_k A(m,k) B(n,k)); see csrc/gemm_tcgen05.cu.

    ``a`` is [M,K] (K-major) or [K,M] (``a_mn``); ``b`` is [N,K] or [K,N] (``b_mn``); bf16.
    ``out`` is bf16 or fp32 [M,N].  ``z``/``deriv`` multiply by f'(z) (activation backward).
    ``b_peers = (pointers, shard_rows, own_rank)``: the weight matrix ``b`` is row-sharded over the replicas at
    ``pointers`` (one per data-parallel rank, same shape/pitch); every tile is TMA-loaded from its owner.
    """
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb, "inner dimensions differ: %d vs %d" % (K, Kb)
    assert out.shape[0] == M and out.shape[1] == N, "out shape %s != (%d,%d)" % (tuple(out.shape), M, N)
    out_fp32 = 1 if out.dtype == torch.float32 else 0
    _FLOPS[0] = 2.0 * M * N * K
    L = _lib.lib()
    if b_peers is None:
        rc = L.vb_gemm_bf16(_p(a), c_longlong(_ld(a)), c_int(int(a_mn)), _p(b), c_longlong(_ld(b)), c_int(int(b_mn)),
                            _p(out), c_longlong(_ld(out)), c_int(M), c_int(N), c_int(K), _p(bias), c_int(act),
                            c_int(out_fp32), c_int(int(accumulate)), _p(z), c_longlong(_ld(z) if z is not None else 0),
                            c_int(deriv), c_float(alpha), c_int(split_k), c_int(block_n), _stream())
        _lib.check(rc, "vb_gemm_bf16")
    else:
        ptrs, shard_rows, own = b_peers
        arr = (c_void_p * len(ptrs))(*[c_void_p(int(q)) for q in ptrs])
        rc = L.vb_gemm_bf16_sharded_b(_p(a), c_longlong(_ld(a)), c_int(int(a_mn)), arr, c_int(len(ptrs)),
                                      c_int(int(shard_rows)), c_int(int(own)), c_longlong(_ld(b)), c_int(int(b_mn)),
                                      _p(out), c_longlong(_ld(out)), c_int(M), c_int(N), c_int(K), _p(bias), c_int(act),
                                      c_int(out_fp32), c_int(int(accumulate)), _p(z),
                                      c_longlong(_ld(z) if z is not None else 0), c_int(deriv), c_float(alpha),
                                      c_int(split_k), c_int(block_n), _stream())
        _lib.check(rc, "vb_gemm_bf16_sharded_b")
    LAUNCHES[0] += 1
    return out


def gemm_fused_gd(a, b, w, vel, shadow, *, a_mn=True, b_mn=True, lr, wd=0.0, l1_vs_l2=0.0, moment=0.0, alpha=1.0,
                  block_n=0):
    """Weight-gradient GEMM whose epilogue applies the momentum GD step to ``w`` / ``vel`` (fp32 [M, N]) and refreshes
    the bf16 ``shadow``; the gradient is never stored.  ``a`` / ``b`` as in :func:`gemm`."""
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16
    assert w.dtype == torch.float32 and vel.dtype == torch.float32 and w.dim() == 2 and vel.shape == w.shape
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb and tuple(w.shape) == (M, N) and _ld(vel) == _ld(w)
    assert shadow is None or (shadow.dtype == torch.bfloat16 and tuple(shadow.shape) == (M, N))
    _FLOPS[0] = 2.0 * M * N * K
    rc = _lib.lib().vb_gemm_bf16_fused_gd(
        _p(a), c_longlong(_ld(a)), c_int(int(a_mn)), _p(b), c_longlong(_ld(b)), c_int(int(b_mn)), c_int(M), c_int(N),
        c_int(K), _p(w), c_longlong(_ld(w)), _p(vel), _p(shadow), c_longlong(_ld(shadow) if shadow is not None else 0),
        c_float(lr), c_float(wd), c_float(l1_vs_l2), c_float(moment), c_float(alpha), c_int(block_n), _stream())
    _lib.check(rc, "vb_gemm_bf16_fused_gd")
    LAUNCHES[0] += 1
    return w


def gemm_tf32(a, b, out, *, a_mn=False, b_mn=False, bias=None, act=ACT_LINEAR, accumulate=False, z=None, deriv=0,
              alpha=1.0, split_k=1, block_n=0):
    """``gemm`` with fp32 operands multiplied as TF32 (fp32 accumulate): the "float" precision mode."""
    assert a.dtype == torch.float32 and b.dtype == torch.float32
    if a_mn:
        K, M = a.shape
    else:
        M, K = a.shape
    if b_mn:
        Kb, N = b.shape
    else:
        N, Kb = b.shape
    assert K == Kb and tuple(out.shape) == (M, N)
    assert z is None or z.dtype == torch.float32
    _FLOPS[0] = 2.0 * M * N * K
    rc = _lib.lib().vb_gemm_tf32(_p(a), c_longlong(_ld(a)), c_int(int(a_mn)), _p(b), c_longlong(_ld(b)),
                                 c_int(int(b_mn)), _p(out), c_longlong(_ld(out)), c_int(M), c_int(N), c_int(K),
                                 _p(bias), c_int(act), c_int(1 if out.dtype == torch.float32 else 0),
                                 c_int(int(accumulate)), _p(z), c_longlong(_ld(z) if z is not None else 0),
                                 c_int(deriv), c_float(alpha), c_int(split_k), c_int(block_n), _stream())
    _lib.check(rc, "vb_gemm_tf32")
    LAUNCHES[0] += 1
    return out


def _conv_gemm(mode, t4, ky, kx, sy, sx, low_h, low_w, n_h, n_w, other, other_rows, cg, out, M, N, bias=None,
               act=ACT_LINEAR, accumulate=False, z=None, deriv=0, alpha=1.0, split_k=1, block_n=0):
    assert t4.dtype == torch.bfloat16 and other.dtype == torch.bfloat16 and t4.dim() == 4
    B, H, W, C = t4.shape
    pitch = t4.stride(2)
    _FLOPS[0] = 2.0 * (B * n_h * n_w if mode != 3 else M) * (N if mode != 3 else B * n_h * n_w) * ky * kx * \
        (C if mode != 3 else cg)
    assert t4.stride(3) == 1 and t4.stride(1) == W * pitch and t4.stride(0) == H * W * pitch, \
        "im2col-mode tensor must be a (channel-sliced) contiguous NHWC view"
    rc = _lib.lib().vb_conv_gemm_bf16(
        c_int(mode), _p(t4), c_int(B), c_int(H), c_int(W), c_longlong(pitch), c_int(C), c_int(ky), c_int(kx),
        c_int(sy), c_int(sx), c_int(low_h), c_int(low_w), c_int(n_h), c_int(n_w), _p(other), c_longlong(_ld(other)),
        c_int(other_rows), c_int(cg), _p(out), c_longlong(_ld(out)), c_int(M), c_int(N), _p(bias), c_int(act),
        c_int(1 if out.dtype == torch.float32 else 0), c_int(int(accumulate)), _p(z),
        c_longlong(_ld(z) if z is not None else 0), c_int(deriv), c_float(alpha), c_int(split_k), c_int(block_n),
        _stream())
    _lib.check(rc, "vb_conv_gemm_bf16")
    LAUNCHES[0] += 1
    return out


def conv_forward(x4, w, out2, g, bias=None, act=ACT_LINEAR, block_n=0):
    """Implicit-GEMM convolution: out2[B*OH*OW, n] = act(im2col(x4) @ w^T + bias) with the im2col operand fetched
    by TMA in im2col mode (no column matrix).  ``x4`` [B,H,W,Cg] (may be a channel slice), ``w`` [n, ky*kx*Cg];
    ``g`` is the layer geometry dict (OH, OW, ky, kx, sy, sx, pt, pl)."""
    B, H, W, Cg = x4.shape
    n = w.shape[0]
    assert w.shape[1] == g["ky"] * g["kx"] * Cg and tuple(out2.shape) == (B * g["OH"] * g["OW"], n)
    return _conv_gemm(1, x4, g["ky"], g["kx"], g["sy"], g["sx"], -g["pt"], -g["pl"], g["OH"], g["OW"], w, n, Cg, out2,
                      out2.shape[0], n, bias=bias, act=act, block_n=block_n)


def conv_err_input(dy4, w, dx2, g, z=None, deriv=0, block_n=0):
    """dx2[B*H*W, Cg] = (conv-transpose of dy4 [B,OH,OW,n] with w [n, ky*kx*Cg]) * f'(z); unit stride only.
    ``g`` additionally carries H and W of the layer input."""
    B, OH, OW, n = dy4.shape
    Cg = dx2.shape[1]
    assert g["sx"] == 1 and g["sy"] == 1 and w.shape[0] == n and w.shape[1] == g["ky"] * g["kx"] * Cg
    assert dx2.shape[0] == B * g["H"] * g["W"]
    return _conv_gemm(2, dy4, g["ky"], g["kx"], 1, 1, -(g["ky"] - 1 - g["pt"]), -(g["kx"] - 1 - g["pl"]), g["H"], g["W"],
                      w, n, Cg, dx2, dx2.shape[0], Cg, z=z, deriv=deriv, block_n=block_n)


def _conv_gemm_grouped(mode, groups, t4, ky, kx, sy, sx, low_h, low_w, n_h, n_w, w, cg, out2, bias=None,
                       act=ACT_LINEAR, z=None, deriv=0, block_n=0):
    B, H, W, C = t4.shape
    pitch = t4.stride(2)
    assert t4.is_contiguous() and out2.dtype == torch.bfloat16
    M, N = out2.shape
    _FLOPS[0] = 2.0 * M * N * ky * kx * (C // groups)
    rc = _lib.lib().vb_conv_gemm_grouped_bf16(
        c_int(mode), c_int(groups), _p(t4), c_int(B), c_int(H), c_int(W), c_longlong(pitch), c_int(C), c_int(ky),
        c_int(kx), c_int(sy), c_int(sx), c_int(low_h), c_int(low_w), c_int(n_h), c_int(n_w), _p(w),
        c_longlong(_ld(w)), c_int(w.shape[0]), c_int(cg), _p(out2), c_longlong(_ld(out2)), c_int(M), c_int(N),
        _p(bias), c_int(act), c_int(0), _p(z), c_longlong(_ld(z) if z is not None else 0), c_int(deriv),
        c_int(block_n), _stream())
    _lib.check(rc, "vb_conv_gemm_grouped_bf16")
    LAUNCHES[0] += 1
    return out2


def grouped_conv_ok(C, n_kernels, groups):
    """Can all groups of the layer run in one implicit-GEMM launch? (per-group widths multiples of 64)"""
    return groups > 1 and C % groups == 0 and n_kernels % groups == 0 and (C // groups) % 64 == 0 \
        and (n_kernels // groups) % 64 == 0


def conv_forward_grouped(x4, w, out2, g, groups, bias=None, act=ACT_LINEAR, block_n=0):
    """All groups of ``conv_forward`` in one launch: full NHWC input, full [n, ky*kx*Cg] weights, full output."""
    return _conv_gemm_grouped(1, groups, x4, g["ky"], g["kx"], g["sy"], g["sx"], -g["pt"], -g["pl"], g["OH"], g["OW"],
                              w, x4.shape[3] // groups, out2, bias=bias, act=act, block_n=block_n)


def conv_err_input_grouped(dy4, w, dx2, g, groups, z=None, deriv=0, block_n=0):
    """All groups of ``conv_err_input`` in one launch."""
    assert g["sx"] == 1 and g["sy"] == 1
    return _conv_gemm_grouped(2, groups, dy4, g["ky"], g["kx"], 1, 1, -(g["ky"] - 1 - g["pt"]),
                              -(g["kx"] - 1 - g["pl"]), g["H"], g["W"], w, dx2.shape[1] // groups, dx2, z=z,
                              deriv=deriv, block_n=block_n)


def conv_halo_preferred(cg, g):
    """Layer shapes on which the halo-tile kernel beats the im2col-mode kernel (measured, profiles/
    conv_microbench_v3.json): few input channels per group with many taps on images of >= 24 rows - AlexNet conv1
    after space-to-depth (+17 %) and conv2 (+15 % forward, +6 % err_input).  On 13x13 images its 128-row tiles waste a
    third of the MMA rows.  ``root.common.engine.conv_halo``: "auto" (default), "on", "off"."""
    from ..config import root, get
    mode = get(root.common.engine.conv_halo, "auto")
    if mode == "off" or g["sx"] != 1 or g["sy"] != 1 or g["OW"] + g["kx"] - 1 > 64:
        return False
    if mode == "on":
        return True
    return cg <= 64 and g["kx"] * g["ky"] >= 9 and min(g["OH"], g.get("H", g["OH"])) >= 24


def conv_halo_set_options(mt=0, resident=-1):
    """Tuning / test hook of the halo-tile kernel: sub-tiles per CTA tile (0 = choose, 1, 2) and resident weights
    (-1 = choose, 0 = never, 1 = whenever they fit)."""
    _lib.check(_lib.lib().vb_conv_halo_set_options(c_int(mt), c_int(resident)), "vb_conv_halo_set_options")


def _conv_halo(mode, t4, ky, kx, low_h, low_w, n_h, n_w, w, cg, out2, bias=None, act=ACT_LINEAR, z=None, deriv=0):
    """Halo-tile implicit GEMM (csrc/conv_halo.cu); returns False when the geometry is not supported."""
    B, H, W, C = t4.shape
    pitch = t4.stride(2)
    if not (t4.stride(3) == 1 and t4.stride(1) == W * pitch and t4.stride(0) == H * W * pitch):
        return False
    N = out2.shape[1]
    _FLOPS[0] = 2.0 * B * n_h * n_w * N * ky * kx * C
    rc = _lib.lib().vb_conv_halo_bf16(
        c_int(mode), _p(t4), c_int(B), c_int(H), c_int(W), c_longlong(pitch), c_int(C), c_int(ky), c_int(kx),
        c_int(low_h), c_int(low_w), c_int(n_h), c_int(n_w), _p(w), c_longlong(_ld(w)), c_int(w.shape[0]), c_int(cg),
        _p(out2), c_longlong(_ld(out2)), c_int(N), _p(bias), c_int(act), _p(z),
        c_longlong(_ld(z) if z is not None else 0), c_int(deriv), _stream())
    if rc == 1:
        return False
    _lib.check(rc, "vb_conv_halo_bf16")
    LAUNCHES[0] += 1
    return True


def conv_forward_halo(x4, w, out2, g, bias=None, act=ACT_LINEAR):
    """``conv_forward`` through the halo-tile kernel (unit stride, OW + kx - 1 <= 64); False if unsupported."""
    if g["sx"] != 1 or g["sy"] != 1 or out2.dtype != torch.bfloat16:
        return False
    return _conv_halo(1, x4, g["ky"], g["kx"], -g["pt"], -g["pl"], g["OH"], g["OW"], w, x4.shape[3], out2, bias=bias,
                      act=act)


def conv_err_input_halo(dy4, w, dx2, g, z=None, deriv=0):
    """``conv_err_input`` through the halo-tile kernel; False if unsupported."""
    if g["sx"] != 1 or g["sy"] != 1 or dx2.dtype != torch.bfloat16:
        return False
    return _conv_halo(2, dy4, g["ky"], g["kx"], -(g["ky"] - 1 - g["pt"]), -(g["kx"] - 1 - g["pl"]), g["H"], g["W"], w,
                      dx2.shape[1], dx2, z=z, deriv=deriv)


def conv_weight_grad(x4, dy2, gw, g, accumulate=True, split_k=0, alpha=1.0, block_n=0):
    """gw[n, ky*kx*Cg] (+)= dy2^T @ im2col(x4) with the im2col operand fetched by TMA (fp32 output)."""
    B, H, W, Cg = x4.shape
    n = dy2.shape[1]
    assert dy2.shape[0] == B * g["OH"] * g["OW"] and tuple(gw.shape) == (n, g["ky"] * g["kx"] * Cg)
    assert gw.dtype == torch.float32
    return _conv_gemm(3, x4, g["ky"], g["kx"], g["sy"], g["sx"], -g["pt"], -g["pl"], g["OH"], g["OW"], dy2, n, Cg, gw,
                      n, g["ky"] * g["kx"] * Cg, accumulate=accumulate, split_k=split_k, alpha=alpha, block_n=block_n)


def space_to_depth(x4, xs, sy, sx):
    """xs[b,h',w',(dy*sx+dx)*C+c] = x4[b,h'*sy+dy,w'*sx+dx,c] (zeros past the image); xs is bf16."""
    B, H, W, C = x4.shape
    assert xs.dtype == torch.bfloat16 and tuple(xs.shape) == (B, (H + sy - 1) // sy, (W + sx - 1) // sx, sy * sx * C)
    assert x4.is_contiguous() and xs.is_contiguous()
    rc = _lib.lib().vb_space_to_depth(_p(x4), c_int(_dt(x4)), _p(xs), c_int(B), c_int(H), c_int(W), c_int(C),
                                      c_int(sy), c_int(sx), _stream())
    _lib.check(rc, "vb_space_to_depth")
    LAUNCHES[0] += 1
    return xs


def gather_cols(src, dst, col_map, accumulate=False):
    """dst[r, j] (+)= src[r, col_map[j]] (0 where col_map[j] < 0); col_map is an int32 device tensor."""
    assert col_map.dtype == torch.int32 and col_map.numel() == dst.shape[1] and src.shape[0] == dst.shape[0]
    rc = _lib.lib().vb_gather_cols(_p(src), c_int(_dt(src)), c_longlong(_ld(src)), _p(dst), c_int(_dt(dst)),
                                   c_longlong(_ld(dst)), _p(col_map), c_int(dst.shape[0]), c_int(dst.shape[1]),
                                   c_int(int(accumulate)), _stream())
    _lib.check(rc, "vb_gather_cols")
    LAUNCHES[0] += 1
    return dst


def gather_minibatch(src, indices, start, count, dst, mean=None, rdisp=None, labels=None, mb_labels=None,
                     mb_indices=None):
    """K1/K2/K4: dst[s] = normalise(cast(src[indices[start+s]])) for s<count else 0."""
    src2 = src.reshape(src.shape[0], -1)
    dst2 = dst.reshape(dst.shape[0], -1) if dst.dim() != 2 else dst
    L = _lib.lib()
    rc = L.vb_gather_minibatch(_p(src2), c_int(_dt(src2)), c_longlong(src2.stride(0)), _p(indices), c_int(start),
                               c_int(count), c_int(dst2.shape[0]), c_int(src2.shape[1]), _p(dst2), c_int(_dt(dst2)),
                               c_longlong(dst2.stride(0)), _p(mean), _p(rdisp), _p(labels), _p(mb_labels),
                               _p(mb_indices), _stream())
    _lib.check(rc, "vb_gather_minibatch")
    LAUNCHES[0] += 2 if (mb_labels is not None or mb_indices is not None) else 1
    _BYTES[0] = count * src2.shape[1] * src2.element_size() + dst2.numel() * dst2.element_size() + (
        2 * 4 * src2.shape[1] if mean is not None else 0)
    return dst


def gather_s2d(src, indices, start, count, xs, H, W, C, sy, sx, mean=None, rdisp=None, labels=None, mb_labels=None,
               mb_indices=None):
    """Minibatch gather fused with the first convolution's space-to-depth transform: ``xs`` [rows, ceil(H/sy),
    ceil(W/sx), sy*sx*C] bf16 receives normalise(cast(src[indices[start+s]])) rearranged; False when the geometry is
    not covered (nothing was launched)."""
    src2 = src.reshape(src.shape[0], -1)
    assert xs.dtype == torch.bfloat16 and xs.is_contiguous() and src2.shape[1] == H * W * C
    rc = _lib.lib().vb_gather_s2d(_p(src2), c_int(_dt(src2)), c_longlong(src2.stride(0)), _p(indices), c_int(start),
                                  c_int(count), c_int(xs.shape[0]), c_int(H), c_int(W), c_int(C), c_int(sy), c_int(sx),
                                  _p(xs), _p(mean), _p(rdisp), _p(labels), _p(mb_labels), _p(mb_indices), _stream())
    if rc == 1:
        return False
    _lib.check(rc, "vb_gather_s2d")
    LAUNCHES[0] += 2 if (mb_labels is not None or mb_indices is not None) else 1
    _BYTES[0] = count * src2.shape[1] * src2.element_size() + xs.numel() * 2 + (
        2 * 4 * src2.shape[1] if mean is not None else 0)
    return True


def mean_disp_normalize(inp, mean, rdisp, out):
    inp2 = inp.reshape(inp.shape[0], -1)
    rc = _lib.lib().vb_mean_disp_normalize(_p(inp2), c_int(_dt(inp2)), _p(mean), _p(rdisp), _p(out), c_int(_dt(out)),
                                           c_int(inp2.shape[0]), c_int(inp2.shape[1]), _stream())
    _lib.check(rc, "vb_mean_disp_normalize")
    LAUNCHES[0] += 1
    return out


def copy_2d(src, dst):
    """dst[:, :] = cast(src) for 2-D (possibly pitched / column-sliced) tensors."""
    rc = _lib.lib().vb_copy_2d(_p(src), c_int(_dt(src)), c_longlong(src.stride(0)), _p(dst), c_int(_dt(dst)),
                               c_longlong(dst.stride(0)), c_int(src.shape[0]), c_int(src.shape[1]), _stream())
    _lib.check(rc, "vb_copy_2d")
    LAUNCHES[0] += 1
    return dst


def activation_forward(act, x, y, factor=1.0):
    rc = _lib.lib().vb_activation_forward(c_int(act), _p(x), _p(y), c_longlong(x.numel()), c_int(_dt(x)),
                                          c_float(factor), _stream())
    _lib.check(rc, "vb_activation_forward")
    LAUNCHES[0] += 1
    return y


def activation_backward(act, err_out, x, y, err_in, factor=1.0):
    rc = _lib.lib().vb_activation_backward(c_int(act), _p(err_out), _p(x), _p(y), _p(err_in),
                                           c_longlong(err_out.numel()), c_int(_dt(err_out)), c_float(factor),
                                           _stream())
    _lib.check(rc, "vb_activation_backward")
    LAUNCHES[0] += 1
    return err_in


def dropout(inp, out, states_in, states_out, ratio):
    n_states = states_in.numel() // 2
    rc = _lib.lib().vb_dropout(_p(inp), _p(out), c_longlong(inp.numel()), c_int(_dt(inp)), _p(states_in),
                               _p(states_out), c_int(n_states), c_float(ratio), _stream())
    _lib.check(rc, "vb_dropout")
    LAUNCHES[0] += 1
    return out


def gd_update(w, grad, vel=None, acc=None, w_bf16=None, *, lr, wd=0.0, l1_vs_l2=0.0, moment=0.0, grad_scale=1.0,
              solver=0, eps=1e-6, rho=0.9, zero_grad=False):
    rc = _lib.lib().vb_gd_update(_p(w), _p(grad), _p(vel), _p(acc), _p(w_bf16), c_longlong(w.numel()), c_float(lr),
                                 c_float(wd), c_float(l1_vs_l2), c_float(moment), c_float(grad_scale), c_int(solver),
                                 c_float(eps), c_float(rho), c_int(int(zero_grad)), _stream())
    _lib.check(rc, "vb_gd_update")
    LAUNCHES[0] += 1


def ortho_grad(w2, grad2, sums, f, by_row=False):
    """grad2[i,j] += f * (sums[j] - w2[i,j])  (``by_row``: sums[i]): the factor_ortho regulariser of the GD units."""
    assert w2.dtype == torch.float32 and grad2.dtype == torch.float32 and w2.shape == grad2.shape
    rc = _lib.lib().vb_ortho_grad(_p(w2), c_longlong(_ld(w2)), _p(grad2), c_longlong(_ld(grad2)), _p(sums),
                                  c_longlong(w2.shape[0]), c_longlong(w2.shape[1]), c_float(f), c_int(int(by_row)),
                                  _stream())
    _lib.check(rc, "vb_ortho_grad")
    LAUNCHES[0] += 1


def cast(inp, out):
    rc = _lib.lib().vb_cast(_p(inp), c_int(_dt(inp)), _p(out), c_int(_dt(out)), c_longlong(inp.numel()), _stream())
    _lib.check(rc, "vb_cast")
    LAUNCHES[0] += 1
    return out


def im2col(x, cols, geom):
    """x [B,H,W,C] bf16 -> cols [B*OH*OW, ld] bf16. geom: dict with OH,OW,ky,kx,sy,sx,pt,pl,groups."""
    B, H, W, C = x.shape
    rc = _lib.lib().vb_im2col(_p(x), _p(cols), c_int(B), c_int(H), c_int(W), c_int(C), c_int(geom["OH"]),
                              c_int(geom["OW"]), c_int(geom["ky"]), c_int(geom["kx"]), c_int(geom["sy"]),
                              c_int(geom["sx"]), c_int(geom["pt"]), c_int(geom["pl"]), c_int(geom.get("groups", 1)),
                              c_longlong(cols.stride(0)), _stream())
    _lib.check(rc, "vb_im2col")
    LAUNCHES[0] += 1
    return cols


def col2im(dcols, dx, geom, z=None, deriv=0):
    B, H, W, C = dx.shape
    rc = _lib.lib().vb_col2im(_p(dcols), _p(dx), c_int(B), c_int(H), c_int(W), c_int(C), c_int(geom["OH"]),
                              c_int(geom["OW"]), c_int(geom["ky"]), c_int(geom["kx"]), c_int(geom["sy"]),
                              c_int(geom["sx"]), c_int(geom["pt"]), c_int(geom["pl"]), c_int(geom.get("groups", 1)),
                              c_longlong(dcols.stride(0)), _p(z), c_int(deriv), _stream())
    _lib.check(rc, "vb_col2im")
    LAUNCHES[0] += 1
    return dx


def pool_forward(kind, x, y, offs, ky, kx, sy, sx, rnd=None):
    B, H, W, C = x.shape
    _, OH, OW, _ = y.shape
    rc = _lib.lib().vb_pool_forward(c_int(kind), _p(x), _p(y), _p(offs), c_int(_dt(x)), c_int(B), c_int(H), c_int(W),
                                    c_int(C), c_int(OH), c_int(OW), c_int(ky), c_int(kx), c_int(sy), c_int(sx),
                                    _p(rnd), _stream())
    _lib.check(rc, "vb_pool_forward")
    LAUNCHES[0] += 1
    return y


def pool_backward(kind, dy, offs, dx, ky, kx, sy, sx, z=None, deriv=0):
    """Pooling backward; ``z``/``deriv`` fold in f'(z) of the activation that produced the pooling input."""
    B, H, W, C = dx.shape
    _, OH, OW, _ = dy.shape
    rc = _lib.lib().vb_pool_backward_fused(c_int(kind), _p(dy), _p(offs), _p(dx), c_int(_dt(dy)), c_int(B), c_int(H),
                                           c_int(W), c_int(C), c_int(OH), c_int(OW), c_int(ky), c_int(kx),
                                           c_int(sy), c_int(sx), _p(z), c_int(deriv), _stream())
    _lib.check(rc, "vb_pool_backward_fused")
    LAUNCHES[0] += 1
    return dx


def pool_forward_local(kind, x, y, offs, offs8, ky, kx, sy, sx):
    """Max/maxabs pooling recording the 1-byte window-local argmax (``offs8``); ``offs`` (int32) optional.
    Returns False when the shape is not supported by this path."""
    B, H, W, C = x.shape
    _, OH, OW, _ = y.shape
    rc = _lib.lib().vb_pool_forward_local(c_int(kind), _p(x), _p(y), _p(offs), _p(offs8), c_int(B), c_int(H), c_int(W),
                                          c_int(C), c_int(OH), c_int(OW), c_int(ky), c_int(kx), c_int(sy), c_int(sx),
                                          _stream())
    if rc == 1:
        return False
    _lib.check(rc, "vb_pool_forward_local")
    LAUNCHES[0] += 1
    return True


def pool_backward_local(kind, dy, offs8, dx, ky, kx, sy, sx, z=None, deriv=0):
    B, H, W, C = dx.shape
    _, OH, OW, _ = dy.shape
    rc = _lib.lib().vb_pool_backward_local(c_int(kind), _p(dy), _p(offs8), _p(dx), c_int(B), c_int(H), c_int(W),
                                           c_int(C), c_int(OH), c_int(OW), c_int(ky), c_int(kx), c_int(sy), c_int(sx),
                                           _p(z), c_int(deriv), _stream())
    if rc == 1:
        return False
    _lib.check(rc, "vb_pool_backward_local")
    LAUNCHES[0] += 1
    return True


def pool_expand_offsets(offs8, offs, in_shape4, ky, kx, sy, sx):
    B, H, W, C = in_shape4
    _, OH, OW, _ = offs8.shape
    rc = _lib.lib().vb_pool_expand_offsets(_p(offs8), _p(offs), c_int(B), c_int(H), c_int(W), c_int(C), c_int(OH),
                                           c_int(OW), c_int(ky), c_int(kx), c_int(sy), c_int(sx), _stream())
    _lib.check(rc, "vb_pool_expand_offsets")
    LAUNCHES[0] += 1
    return offs


def lrn_forward(x, y, n, alpha, beta, k):
    C = x.shape[-1]
    rc = _lib.lib().vb_lrn_forward(_p(x), _p(y), c_int(_dt(x)), c_longlong(x.numel() // C), c_int(C), c_int(n),
                                   c_float(alpha), c_float(beta), c_float(k), _stream())
    _lib.check(rc, "vb_lrn_forward")
    LAUNCHES[0] += 1
    return y


def lrn_backward(x, dy, dx, n, alpha, beta, k, deriv=0, partials=None):
    """LRN backward; ``deriv`` folds in f'(x) of the activation that produced x.  ``partials`` ([rows, C] fp32 scratch):
    the kernel also leaves per-block partial sums of dx's columns there (the bias gradient of the convolution that
    produced x; fold them with :func:`col_sum_finish`) and the number of rows written is returned - or None when the
    shape's kernel cannot produce them (nothing launched).  Without ``partials`` returns dx."""
    C = x.shape[-1]
    if partials is not None:
        if x.dtype != torch.bfloat16:
            return None
        assert partials.dtype == torch.float32 and partials.dim() == 2 and partials.shape[1] == C
        blocks = c_int(0)
        rc = _lib.lib().vb_lrn_backward_colsum(_p(x), _p(dy), _p(dx), c_longlong(x.numel() // C), c_int(C), c_int(n),
                                               c_float(alpha), c_float(beta), c_float(k), c_int(deriv), _p(partials),
                                               c_int(partials.shape[0]), ctypes.byref(blocks), _stream())
        if rc == 1:
            return None
        _lib.check(rc, "vb_lrn_backward_colsum")
        LAUNCHES[0] += 1
        return blocks.value
    rc = _lib.lib().vb_lrn_backward_fused(_p(x), _p(dy), _p(dx), c_int(_dt(x)), c_longlong(x.numel() // C), c_int(C),
                                          c_int(n), c_float(alpha), c_float(beta), c_float(k), c_int(deriv),
                                          _stream())
    _lib.check(rc, "vb_lrn_backward_fused")
    LAUNCHES[0] += 1
    return dx


def lrn_pool_backward(x4, dpool4, offs8, dx4, ky, kx, sy, sx, n, alpha, beta, k, deriv=0):
    """Max-pooling backward fused into LRN backward (csrc/fast_paths.cu); False when the geometry is not covered."""
    B, H, W, C = x4.shape
    OH, OW = dpool4.shape[1], dpool4.shape[2]
    if x4.dtype != torch.bfloat16 or dpool4.dtype != torch.bfloat16 or not (x4.is_contiguous() and dpool4.is_contiguous()):
        return False
    rc = _lib.lib().vb_lrn_pool_backward(_p(x4), _p(dpool4), _p(offs8), _p(dx4), c_int(B), c_int(H), c_int(W), c_int(C),
                                         c_int(OH), c_int(OW), c_int(ky), c_int(kx), c_int(sy), c_int(sx), c_int(n),
                                         c_float(alpha), c_float(beta), c_float(k), c_int(deriv), _stream())
    if rc == 1:
        return False
    _lib.check(rc, "vb_lrn_pool_backward")
    LAUNCHES[0] += 1
    return True


def softmax_eval(s, y, err, labels, max_idx, *, err_scale=1.0, n_err=None, confusion=None, max_err_sum=None,
                 loss_sum=None):
    batch, N = s.shape
    out_dt = _dt(y if y is not None else err)
    rc = _lib.lib().vb_softmax_eval(_p(s), c_int(_dt(s)), c_longlong(_ld(s)), _p(y),
                                    c_longlong(_ld(y) if y is not None else 0), _p(err),
                                    c_longlong(_ld(err) if err is not None else 0), c_int(out_dt), _p(labels),
                                    _p(max_idx), c_int(batch), c_int(N), c_float(err_scale), _p(n_err), _p(confusion),
                                    _p(max_err_sum), _p(loss_sum), _stream())
    _lib.check(rc, "vb_softmax_eval")
    LAUNCHES[0] += 1


def mse_eval(y, target, err, batch_valid, *, err_scale=1.0, metrics=None, mse_per_sample=None):
    batch, N = y.shape
    rc = _lib.lib().vb_mse_eval(_p(y), c_longlong(_ld(y)), _p(target), c_longlong(_ld(target)), _p(err),
                                c_longlong(_ld(err) if err is not None else 0), c_int(_dt(y)), c_int(batch),
                                c_int(batch_valid), c_int(N), c_float(err_scale), _p(metrics), _p(mse_per_sample),
                                _stream())
    _lib.check(rc, "vb_mse_eval")
    LAUNCHES[0] += 1


def row_sum(a, out):
    rc = _lib.lib().vb_row_sum(_p(a), c_int(_dt(a)), c_longlong(_ld(a)), c_int(a.shape[0]), c_int(a.shape[1]),
                               _p(out), _stream())
    _lib.check(rc, "vb_row_sum")
    LAUNCHES[0] += 1
    return out


def col_sum_scratch_elems(rows, cols):
    return int(_lib.lib().vb_col_sum_scratch_elems(c_int(rows), c_int(cols)))


def col_sum(a, out, scratch, alpha=1.0, accumulate=False):
    """out[j] (+)= alpha * sum_i a[i,j]; ``scratch`` fp32 with col_sum_scratch_elems() elements."""
    assert scratch.numel() >= col_sum_scratch_elems(a.shape[0], a.shape[1])
    rc = _lib.lib().vb_col_sum(_p(a), c_int(_dt(a)), c_longlong(_ld(a)), c_int(a.shape[0]), c_int(a.shape[1]),
                               _p(out), c_float(alpha), c_int(int(accumulate)), _p(scratch), _stream())
    _lib.check(rc, "vb_col_sum")
    LAUNCHES[0] += 2
    return out


def col_sum_gd(a, grad, scratch, w, vel, *, lr, wd=0.0, l1_vs_l2=0.0, moment=0.0, alpha=1.0, accumulate=True):
    """``col_sum`` that applies the momentum GD step to ``w`` / ``vel`` (a bias) in its second phase; ``grad`` is the
    accumulation buffer (read when ``accumulate``) and is left zeroed."""
    assert scratch.numel() >= col_sum_scratch_elems(a.shape[0], a.shape[1])
    assert w.dtype == torch.float32 and vel.dtype == torch.float32 and w.numel() == a.shape[1] == vel.numel()
    rc = _lib.lib().vb_col_sum_gd(_p(a), c_int(_dt(a)), c_longlong(_ld(a)), c_int(a.shape[0]), c_int(a.shape[1]),
                                  _p(grad), c_float(alpha), c_int(int(accumulate)), _p(scratch), _p(w), _p(vel),
                                  c_float(lr), c_float(wd), c_float(l1_vs_l2), c_float(moment), _stream())
    _lib.check(rc, "vb_col_sum_gd")
    LAUNCHES[0] += 2
    return w


def col_sum_finish(partials, slabs, grad, w=None, vel=None, *, lr=0.0, wd=0.0, l1_vs_l2=0.0, moment=0.0,
                   accumulate=True):
    """grad (+)= column sums of ``partials[:slabs]``; with ``w`` / ``vel`` the momentum GD step of that bias is applied
    instead and ``grad`` is left zeroed (the second phase of :func:`col_sum_gd` on its own)."""
    assert partials.dtype == torch.float32 and grad.dtype == torch.float32 and partials.shape[1] == grad.numel()
    rc = _lib.lib().vb_col_sum_finish(_p(partials), c_int(slabs), c_int(grad.numel()), _p(grad), c_int(int(accumulate)),
                                      _p(w), _p(vel), c_float(lr), c_float(wd), c_float(l1_vs_l2), c_float(moment),
                                      _stream())
    _lib.check(rc, "vb_col_sum_finish")
    LAUNCHES[0] += 1
    return grad


def xorshift1024star(states, rounds, out):
    n_states = states.numel() // 16
    rc = _lib.lib().vb_xorshift1024star(_p(states), c_int(n_states), c_int(rounds), _p(out), _stream())
    _lib.check(rc, "vb_xorshift1024star")
    LAUNCHES[0] += 1
    return out


def xorshift128plus(states, out):
    n = states.numel() // 2
    rc = _lib.lib().vb_xorshift128plus(_p(states), c_longlong(n), _p(out), _stream())
    _lib.check(rc, "vb_xorshift128plus")
    LAUNCHES[0] += 1
    return out


def fill_random(states, out, kind, amp):
    """kind 0: uniform[-amp, amp); 1: normal(0, amp) clipped at 5 sigma."""
    rc = _lib.lib().vb_fill_random(_p(states), c_int(states.numel() // 2), _p(out), c_longlong(out.numel()),
                                   c_int(kind), c_float(amp), _stream())
    _lib.check(rc, "vb_fill_random")
    LAUNCHES[0] += 1
    return out


# -- optional per-call tracing (utils/roofline.py): CUDA events around every wrapper call ---------------------------
def _traced(fn):
    import functools

    @functools.wraps(fn)
    def wrap(*a, **kw):
        if TRACE is None:
            return fn(*a, **kw)
        tensors = [t for t in list(a) + list(kw.values()) if isinstance(t, torch.Tensor)]
        nbytes = sum(t.numel() * t.element_size() for t in tensors)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        _FLOPS[0] = 0.0
        _BYTES[0] = None
        e0.record()
        r = fn(*a, **kw)
        e1.record()
        if _BYTES[0] is not None:      # the call knows its traffic better than the sizes of its arguments
            nbytes = _BYTES[0]
        TRACE.append([fn.__name__, e0, e1, nbytes, _FLOPS[0], [tuple(t.shape) for t in tensors[:3]]])
        return r
    return wrap


for _name in ("gemm", "gemm_fused_gd", "bias_act", "gemm_tf32", "conv_forward", "conv_forward_grouped", "conv_err_input_grouped", "conv_forward_halo", "conv_err_input_halo", "conv_err_input", "conv_weight_grad", "space_to_depth", "gather_cols",
              "gather_minibatch", "gather_s2d", "lrn_pool_backward", "mean_disp_normalize", "copy_2d", "activation_forward", "activation_backward",
              "dropout", "gd_update", "cast", "im2col", "col2im", "pool_forward", "pool_backward", "pool_forward_local",
              "pool_backward_local", "pool_expand_offsets", "lrn_forward", "lrn_backward", "softmax_eval", "mse_eval",
              "row_sum", "col_sum", "col_sum_gd", "col_sum_finish", "xorshift1024star", "xorshift128plus", "fill_random"):
    globals()[_name] = _traced(globals()[_name])


# -- batched signal kernels (csrc/signal.cu; API in veles_b200/simd) --------------------------------------------------
def _rows(t):
    assert t.dim() == 2 and t.stride(1) == 1, "need [batch, length] rows with unit inner stride"
    return t.shape[0], t.shape[1], t.stride(0)


def sig_convolve(x, h, out, reverse=False):
    """out[b] = x[b] (*) h[b or 0] (full convolution; ``reverse`` = cross-correlation).  fp32 rows."""
    B, X, xp = _rows(x)
    hb, M, hp = _rows(h)
    assert hb in (1, B) and out.shape == (B, X + M - 1)
    rc = _lib.lib().vb_sig_convolve(_p(x), c_longlong(xp), c_int(X), _p(h), c_longlong(hp if hb == B else 0), c_int(M),
                                    c_int(int(reverse)), _p(out), c_longlong(out.stride(0)), c_int(B), _stream())
    _lib.check(rc, "vb_sig_convolve")
    LAUNCHES[0] += 1
    return out


def sig_wavelet(x, lowpass, highpass, hi, lo, ext, level=1, stationary=False):
    """One analysis step over every row of ``x``; ``lowpass`` / ``highpass`` are host float32 arrays."""
    import numpy
    B, n, xp = _rows(x)
    lp = numpy.ascontiguousarray(lowpass, dtype=numpy.float32)
    hp = numpy.ascontiguousarray(highpass, dtype=numpy.float32)
    assert hi.stride(0) == lo.stride(0)
    rc = _lib.lib().vb_sig_wavelet(_p(x), c_longlong(xp), c_int(n), lp.ctypes.data_as(c_void_p),
                                   hp.ctypes.data_as(c_void_p), c_int(lp.size), c_int(level), c_int(int(stationary)),
                                   c_int(int(ext)), _p(hi), _p(lo), c_longlong(hi.stride(0)), c_int(B), _stream())
    _lib.check(rc, "vb_sig_wavelet")
    LAUNCHES[0] += 1
    return hi, lo


def sig_minmax(x, out):
    B, n, xp = _rows(x)
    rc = _lib.lib().vb_sig_minmax_f32(_p(x), c_longlong(xp), c_int(n), c_int(B), _p(out), _stream())
    _lib.check(rc, "vb_sig_minmax_f32")
    LAUNCHES[0] += 1
    return out


def sig_normalize_u8(src, minmax, dst, have_minmax=False):
    """src uint8 [batch, H, W] (row stride free), minmax int32 [batch, 2], dst float32 [batch, H, W] or None."""
    assert src.dim() == 3 and src.stride(2) == 1
    B, H, W = src.shape
    rc = _lib.lib().vb_sig_normalize_u8(
        _p(src), c_longlong(src.stride(0)), c_int(src.stride(1)), c_int(W), c_int(H), c_int(B), _p(minmax),
        c_int(int(have_minmax)), _p(dst), c_longlong(0 if dst is None else dst.stride(0)),
        c_int(W if dst is None else dst.stride(1)), _stream())
    _lib.check(rc, "vb_sig_normalize_u8")
    LAUNCHES[0] += 1 + (0 if have_minmax else 1) - (1 if dst is None else 0)
    return dst


def sig_detect_peaks(x, extremum_type, counts, positions, values, totals):
    B, n, xp = _rows(x)
    rc = _lib.lib().vb_sig_detect_peaks(_p(x), c_longlong(xp), c_int(n), c_int(B), c_int(int(extremum_type)),
                                        _p(counts), _p(positions), _p(values), c_int(positions.shape[1]), _p(totals),
                                        _stream())
    _lib.check(rc, "vb_sig_detect_peaks")
    LAUNCHES[0] += 2
    return totals


def sig_complex_mul(a, b, out, conj_b=False, scale=1.0):
    """Interleaved complex rows (float32 views of complex64): out = a * b (b broadcast when it has one row)."""
    B, n2, ap = _rows(a)
    bb, _, bp = _rows(b)
    rc = _lib.lib().vb_sig_complex_mul(_p(a), c_longlong(ap // 2), _p(b), c_longlong(bp // 2 if bb == B else 0), _p(out),
                                       c_longlong(out.stride(0) // 2), c_int(n2 // 2), c_int(B), c_int(int(conj_b)),
                                       c_float(scale), _stream())
    _lib.check(rc, "vb_sig_complex_mul")
    LAUNCHES[0] += 1
    return out


def sig_convert(src, dst, kind):
    rc = _lib.lib().vb_sig_convert(_p(src), _p(dst), c_longlong(src.numel()), c_int(kind), _stream())
    _lib.check(rc, "vb_sig_convert")
    LAUNCHES[0] += 1
    return dst


# -- recurrent cells (csrc/recurrent.cu) ------------------------------------------------------------------------------
def _opt(t):
    return c_void_p(0) if t is None else _p(t)


def lstm_cell_forward(z, c_prev, c, gates, h_next, out):
    """z [B, 4H] fp32 pre-activations -> gates (saved), c, h into ``h_next`` (bf16 slot or None) and ``out`` rows."""
    B, H = c.shape
    rc = _lib.lib().vb_lstm_cell_forward(
        _p(z), c_longlong(z.stride(0)), _opt(c_prev), _p(c), _p(gates), c_longlong(gates.stride(0)), _opt(h_next),
        c_longlong(0 if h_next is None else h_next.stride(0)), _opt(out), c_int(0 if out is None else _dt(out)),
        c_longlong(0 if out is None else out.stride(0)), c_int(B), c_int(H), _stream())
    _lib.check(rc, "vb_lstm_cell_forward")
    LAUNCHES[0] += 1


def lstm_cell_backward(dh_ext, dh_rec, dc, first, gates, c_prev, c, dz):
    B, H = c.shape
    rc = _lib.lib().vb_lstm_cell_backward(
        _opt(dh_ext), c_int(0 if dh_ext is None else _dt(dh_ext)), c_longlong(0 if dh_ext is None else dh_ext.stride(0)),
        _opt(dh_rec), c_longlong(0 if dh_rec is None else dh_rec.stride(0)), _p(dc), c_int(int(first)), _p(gates),
        c_longlong(gates.stride(0)), _opt(c_prev), _p(c), _p(dz), c_longlong(dz.stride(0)), c_int(B), c_int(H),
        _stream())
    _lib.check(rc, "vb_lstm_cell_backward")
    LAUNCHES[0] += 1


def rnn_cell_forward(z, h_save, h_next, out):
    B, H = h_save.shape
    rc = _lib.lib().vb_rnn_cell_forward(
        _p(z), c_longlong(z.stride(0)), _p(h_save), _opt(h_next), c_longlong(0 if h_next is None else h_next.stride(0)),
        _opt(out), c_int(0 if out is None else _dt(out)), c_longlong(0 if out is None else out.stride(0)), c_int(B),
        c_int(H), _stream())
    _lib.check(rc, "vb_rnn_cell_forward")
    LAUNCHES[0] += 1


def rnn_cell_backward(dh_ext, dh_rec, h_save, dz):
    B, H = h_save.shape
    rc = _lib.lib().vb_rnn_cell_backward(
        _opt(dh_ext), c_int(0 if dh_ext is None else _dt(dh_ext)), c_longlong(0 if dh_ext is None else dh_ext.stride(0)),
        _opt(dh_rec), c_longlong(0 if dh_rec is None else dh_rec.stride(0)), _p(h_save), _p(dz),
        c_longlong(dz.stride(0)), c_int(B), c_int(H), _stream())
    _lib.check(rc, "vb_rnn_cell_backward")
    LAUNCHES[0] += 1
