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


@pytest.mark.parametrize("world", [2, 8])
def test_counter_allreduce_emulated_ranks(world):
    """The class-end counter exchange (``csrc/comm_counters.cu``) with all ranks on one GPU: sums, maxima, the n = 0
    barrier form and slot reuse over several consecutive epochs against numpy."""
    import torch
    from veles_b200.ops import _lib
    L = _lib.lib()
    L.vb_counter_region_bytes.restype = ctypes.c_longlong
    L.vb_counter_params_bytes.restype = ctypes.c_longlong
    dev = torch.device("cuda:0")
    slot = 512
    region = int(L.vb_counter_region_bytes(c_longlong(slot)))
    bases = [torch.zeros(region, dtype=torch.uint8, device=dev) for _ in range(world)]
    epochs = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(world)]
    status = [torch.zeros(1, dtype=torch.int32, device=dev) for _ in range(world)]
    ins = [torch.zeros(slot, dtype=torch.float64, device=dev) for _ in range(world)]
    outs = [torch.full((slot,), -1.0, dtype=torch.float64, device=dev) for _ in range(world)]
    params = torch.zeros(int(L.vb_counter_params_bytes()) * world + 256, dtype=torch.uint8, device=dev)

    def arr(ts):
        return (c_void_p * world)(*[c_void_p(t.data_ptr()) for t in ts])

    rng = np.random.default_rng(world)
    for step, (n, op) in enumerate([(3, 0), (slot, 0), (0, 0), (100, 1), (7, 0)]):
        vals = rng.integers(-1000, 1000, size=(world, max(n, 1))).astype(np.float64) + 0.25
        for r in range(world):
            ins[r][:max(n, 1)] = torch.from_numpy(vals[r]).to(dev)
        rc = L.vb_counter_allreduce_emulated(c_int(world), c_int(n), c_int(op), c_longlong(slot), arr(ins), arr(outs),
                                             arr(bases), arr(epochs), arr(status), c_double(5.0),
                                             c_void_p(params.data_ptr()),
                                             c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, "vb_counter_allreduce_emulated")
        torch.cuda.synchronize()
        ref = vals[:, :n].max(axis=0) if op == 1 else vals[:, :n].sum(axis=0)
        for r in range(world):
            assert int(status[r].item()) == 0 and int(epochs[r].item()) == step + 1
            got = outs[r][:n].cpu().numpy()
            np.testing.assert_array_equal(got, ref)          # same order of additions on every rank: same bits
    # more values than the slot holds are refused before any launch
    rc = L.vb_counter_allreduce_emulated(c_int(world), c_int(slot + 1), c_int(0), c_longlong(slot), arr(ins), arr(outs),
                                         arr(bases), arr(epochs), arr(status), c_double(5.0),
                                         c_void_p(params.data_ptr()), c_void_p(0))
    assert rc != 0


def test_counter_allreduce_times_out_when_a_rank_is_missing():
    """A rank whose peers never arrive sets ``status`` and leaves its epoch untouched (the host raises)."""
    import torch
    from veles_b200.ops import _lib
    L = _lib.lib()
    L.vb_counter_region_bytes.restype = ctypes.c_longlong
    dev = torch.device("cuda:0")
    slot = 64
    bases = [torch.zeros(int(L.vb_counter_region_bytes(c_longlong(slot))), dtype=torch.uint8, device=dev)
             for _ in range(2)]
    epoch = torch.zeros(1, dtype=torch.int32, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    vin = torch.ones(slot, dtype=torch.float64, device=dev)
    vout = torch.zeros(slot, dtype=torch.float64, device=dev)
    peers = (c_void_p * 8)(*([c_void_p(b.data_ptr()) for b in bases] + [c_void_p(0)] * 6))
    rc = L.vb_counter_allreduce(c_int(2), c_int(0), c_int(4), c_int(0), c_longlong(slot), c_void_p(vin.data_ptr()),
                                c_void_p(vout.data_ptr()), c_void_p(bases[0].data_ptr()), peers,
                                c_void_p(epoch.data_ptr()), c_void_p(status.data_ptr()), c_double(0.05),
                                c_void_p(torch.cuda.current_stream().cuda_stream))
    _lib.check(rc, "vb_counter_allreduce")
    torch.cuda.synchronize()
    assert int(status.item()) == 1001 and int(epoch.item()) == 0


@pytest.mark.parametrize("world", [2, 8])
def test_gather_slices_reassembles_sharded_state(world):
    """``vb_gather_slices``: every rank copies each peer's slice of every segment (the slice arithmetic of the update
    kernel) - odd sizes, unaligned segment offsets, segments shorter than the world."""
    import torch
    from veles_b200.ops import _lib
    L = _lib.lib()
    dev = torch.device("cuda:0")
    segs = [(0, 40000), (40064, 203), (40321, 1850), (42200, 3)]
    slab = 42240
    truth = torch.randn(slab, device=dev)
    slabs = []
    for r in range(world):
        t = torch.full((slab,), float(-1 - r), device=dev)
        for off, numel in segs:
            lo, hi = _slice(numel, world, r)
            t[off + lo:off + hi] = truth[off + lo:off + hi]       # each rank is current on its own slices only
        slabs.append(t)
    seg_arr = (c_longlong * (2 * len(segs)))(*[v for sg in segs for v in sg])
    peers = (c_void_p * 8)(*([c_void_p(t.data_ptr()) for t in slabs] + [c_void_p(0)] * (8 - world)))
    before = [t.clone() for t in slabs]
    for r in range(world):
        rc = L.vb_gather_slices(c_int(world), c_int(r), c_int(len(segs)), seg_arr, c_void_p(slabs[r].data_ptr()),
                                peers, c_int(4), c_void_p(torch.cuda.current_stream().cuda_stream))
        _lib.check(rc, "vb_gather_slices")
    torch.cuda.synchronize()
    for r in range(world):
        covered = torch.zeros(slab, dtype=torch.bool, device=dev)
        for off, numel in segs:
            assert torch.equal(slabs[r][off:off + numel], truth[off:off + numel])
            covered[off:off + numel] = True
        assert torch.equal(slabs[r][~covered], before[r][~covered])     # nothing outside the segments is touched
