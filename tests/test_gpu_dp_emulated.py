"""The fused all-reduce + GD kernel of the data-parallel group on ONE GPU: `world` ranks emulated by a single
cooperative launch (blockIdx.y = rank, ``vb_fused_allreduce_gd_emulated``), every rank's slabs living on this GPU.
Checks the flag protocol (two consecutive epochs), the slice arithmetic, the Znicz update math, the bf16 shadow
broadcast, the per-segment options (zero the other parity / always publish the fp32 master) and the sharded master
against a plain fp32 restatement.  (The multimem path and real NVLink need >= 2 GPUs: tests/test_multigpu_dp.py.)"""
import ctypes
from ctypes import c_double, c_float, c_int, c_longlong, c_void_p

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SEG_ZERO_OTHER, SEG_BCAST_MASTER = 1, 2


def _gd(w, g, v, lr, wd, l12, moment, scale):
    g = g * scale + wd * ((1.0 - l12) * w + 0.5 * l12 * np.sign(w))
    v = moment * v - lr * g
    return w + v, v


def _slice(numel, world, rank):
    chunk = (numel + world - 1) // world
    chunk = (chunk + 3) & ~3
    lo = min(numel, chunk * rank)
    return lo, min(numel, lo + chunk)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("sharded", [False, True])
@pytest.mark.parametrize("push", [True, False], ids=["push", "pull"])
def test_fused_allreduce_gd_emulated_ranks(world, sharded, push):
    import torch
    from veles_b200.ops import _lib
    L = _lib.lib()
    L.vb_fused_params_bytes.restype = ctypes.c_longlong
    dev = torch.device("cuda:0")
    torch.manual_seed(world * 2 + int(sharded))
    # three segments in one slab: a weight matrix with a flat bf16 shadow (plainly stored gradient), its bias (no
    # shadow: fp32 master read by the forward pass), and an odd-sized weight matrix whose shadow rows are pitched
    rows2, len2, pitch2 = 37, 50, 56
    segs = [  # offset, numel, shadow_offset, row_len, row_pitch, flags
        [0, 40000, 0, 200, 200, 0],
        [40064, 203, -1, 1, 1, SEG_ZERO_OTHER | SEG_BCAST_MASTER],
        [40320, rows2 * len2, 40000, len2, pitch2, SEG_ZERO_OTHER],
    ]
    hyper = [[0.05, 0.001, 0.2, 0.9], [0.1, 0.0, 0.0, 0.5], [0.02, 0.01, 0.0, 0.0]]
    slab = 40320 + rows2 * len2 + 64
    shadow_elems = 40000 + rows2 * pitch2 + 8
    w0 = torch.randn(slab, device=dev) * 0.3
    ranks = []
    for r in range(world):
        ranks.append(dict(
            grads=[torch.randn(slab, device=dev), torch.randn(slab, device=dev)],
            weights=w0.clone(), vel=torch.zeros(slab, device=dev),
            shadow=torch.full((shadow_elems,), 7.0, device=dev, dtype=torch.bfloat16),
            flags=torch.zeros(64 * 2 * 8, dtype=torch.int32, device=dev),
            epochs=torch.zeros(64, dtype=torch.int32, device=dev),
            counter=torch.zeros(128, dtype=torch.int32, device=dev),
            stage=torch.full((slab,), float("nan"), device=dev),
            status=torch.zeros(1, dtype=torch.int32, device=dev)))
    params = torch.zeros(int(L.vb_fused_params_bytes()) * world + 256, dtype=torch.uint8, device=dev)
    seg_arr = (c_longlong * (6 * len(segs)))(*[v for s in segs for v in s])
    hyp_arr = (c_float * (4 * len(segs)))(*[v for h in hyper for v in h])

    def ptrs(key, idx=None):
        return (c_void_p * world)(*[c_void_p((rk[key] if idx is None else rk[key][idx]).data_ptr()) for rk in ranks])

    scale = 1.0 / world
    # fp32 oracle state
    w_ref = w0.double().cpu().numpy()
    v_ref = np.zeros(slab)
    channel = 3
    for step in range(2):
        parity = 0 if push else (step & 1)      # push: one gradient buffer, cleared in place where flagged
        g_sum = sum(rk["grads"][parity].double().cpu().numpy() for rk in ranks)
        g_before = [rk["grads"][parity].clone() for rk in ranks]
        other_before = [rk["grads"][1 - parity].clone() for rk in ranks]
        rc = L.vb_fused_allreduce_gd_emulated(
            c_int(world), c_int(channel), c_int(len(segs)), seg_arr, hyp_arr, ptrs("grads", parity),
            ptrs("grads", 1 - parity), ptrs("weights"), ptrs("vel"), ptrs("shadow"), ptrs("flags"), ptrs("epochs"),
            ptrs("counter"), ptrs("status"), c_float(scale), c_int(2),
            c_int((2 if sharded else 0) | (16 if push else 0)), c_double(5.0),
            c_void_p(params.data_ptr()), ptrs("stage"), c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, "vb_fused_allreduce_gd_emulated")
        torch.cuda.synchronize()
        for (off, numel, sh_off, row_len, pitch, flags), (lr, wd, l12, mom) in zip(segs, hyper):
            w_new, v_new = _gd(w_ref[off:off + numel], g_sum[off:off + numel], v_ref[off:off + numel], lr, wd, l12,
                               mom, scale)
            for r, rk in enumerate(ranks):
                assert int(rk["status"].item()) == 0
                assert int(rk["epochs"][channel].item()) == step + 1
                lo, hi = _slice(numel, world, r)
                w_r = rk["weights"][off:off + numel].double().cpu().numpy()
                v_r = rk["vel"][off:off + numel].double().cpu().numpy()
                np.testing.assert_allclose(w_r[lo:hi], w_new[lo:hi], rtol=2e-5, atol=2e-6)
                np.testing.assert_allclose(v_r[lo:hi], v_new[lo:hi], rtol=2e-5, atol=2e-6)
                everywhere = (not sharded) or (flags & SEG_BCAST_MASTER)
                if everywhere:      # the whole segment is current on every rank
                    np.testing.assert_allclose(w_r, w_new, rtol=2e-5, atol=2e-6)
                else:               # foreign slices keep the previous step's (own-view) values
                    mask = np.ones(numel, bool)
                    mask[lo:hi] = False
                    assert not np.allclose(w_r[mask], w_new[mask], rtol=1e-6, atol=1e-7)
                # momentum stays with the owner: foreign slices never move
                mask = np.ones(numel, bool)
                mask[lo:hi] = False
                assert np.all(v_r[mask] == 0.0)
                if sh_off >= 0:     # every replica of the bf16 operand copy holds the whole updated segment
                    sh = rk["shadow"].float().cpu().numpy()
                    e = np.arange(numel)
                    got = sh[sh_off + (e // row_len) * pitch + e % row_len]
                    ref = torch.from_numpy(w_new).float().bfloat16().float().numpy()
                    np.testing.assert_allclose(got, ref, rtol=1e-2, atol=1e-6)
                    if pitch > row_len:   # the pitch padding is not written
                        assert sh[sh_off + row_len] == 7.0
                if push:    # accumulating segments restart from zero in place, stored ones are left alone
                    mine = rk["grads"][0][off:off + numel]
                    if flags & SEG_ZERO_OTHER:
                        assert float(mine.abs().max()) == 0.0
                    else:
                        assert torch.equal(mine, g_before[r][off:off + numel])
                    assert torch.equal(rk["grads"][1][off:off + numel], other_before[r][off:off + numel])
                else:
                    other = rk["grads"][1 - parity][off:off + numel]
                    if flags & SEG_ZERO_OTHER:
                        assert float(other.abs().max()) == 0.0
                    else:
                        assert torch.equal(other, other_before[r][off:off + numel])
            # the oracle follows rank-local truth: owners hold the new values
            w_ref[off:off + numel] = w_new
            v_ref[off:off + numel] = v_new
        if sharded:
            # what sync_master() does at class boundaries: every rank adopts the owners' slices
            for (off, numel, *_rest) in segs:
                for r, rk in enumerate(ranks):
                    for q, rq in enumerate(ranks):
                        lo, hi = _slice(numel, world, q)
                        rk["weights"][off + lo:off + hi] = rq["weights"][off + lo:off + hi]
        # fresh gradients for the next step in the buffer that was just cleared / left alone
        for rk in ranks:
            rk["grads"][0 if push else 1 - parity].normal_()
