"""End-to-end device runs of StandardWorkflow against the numpy oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _build(device, layers, loader_cfg, loss="softmax", epochs=3, seed=7, loader="synthetic_fullbatch"):
    from veles_b200 import prng
    from veles_b200.config import root
    from veles_b200.dummy import DummyLauncher
    from veles_b200.znicz.standard_workflow import StandardWorkflow
    import veles_b200.loader.synthetic  # noqa: F401
    root.common.disable.snapshotting = True
    prng.get().seed(seed)
    launcher = DummyLauncher()
    wf = StandardWorkflow(launcher, layers=layers, loader_name=loader, loader_config=loader_cfg,
                          loss_function=loss, decision_config={"max_epochs": epochs})
    wf.initialize(device=device)
    return launcher, wf


MLP = [{"type": "all2all_tanh", "->": {"output_sample_shape": 100},
        "<-": {"learning_rate": 0.05, "gradient_moment": 0.9}},
       {"type": "softmax", "<-": {"learning_rate": 0.05, "gradient_moment": 0.9}}]
CFG = {"minibatch_size": 100, "n_train": 2000, "n_valid": 500, "sample_shape": (784,), "noise": 2.0}


def test_mlp_trains_on_device_and_matches_numpy_oracle():
    from veles_b200.backends import NumpyDevice, B200Device
    l0, ref = _build(NumpyDevice(), MLP, CFG)
    ref.run()
    l1, wf = _build(B200Device(), MLP, CFG)
    wf.run()
    h_ref = ref.decision.epoch_metrics_history
    h_dev = wf.decision.epoch_metrics_history
    assert len(h_ref) == len(h_dev)
    # identical data order and init => the curves must agree up to bf16 rounding
    assert h_dev[-1]["n_err_pt"][1] <= 2.0
    for a, b in zip(h_ref, h_dev):
        assert abs(a["n_err_pt"][1] - b["n_err_pt"][1]) <= 3.0
        assert abs(a["loss"][1] - b["loss"][1]) <= 0.05 + 0.1 * abs(a["loss"][1])
    wf.forwards[0].weights.map_read()
    ref.forwards[0].weights.map_read()
    w_dev, w_ref = wf.forwards[0].weights.mem, ref.forwards[0].weights.mem
    assert np.abs(w_dev - w_ref).max() < 0.05
    l0.stop()
    l1.stop()


def test_first_iteration_gradients_match_oracle():
    """One minibatch, no update: err_input / gradients of every layer vs the numpy backend."""
    from veles_b200.backends import NumpyDevice, B200Device
    layers = [{"type": "all2all_tanh", "->": {"output_sample_shape": 64}, "<-": {"learning_rate": 0.0}},
              {"type": "all2all_str", "->": {"output_sample_shape": 48}, "<-": {"learning_rate": 0.0}},
              {"type": "all2all_sigmoid", "->": {"output_sample_shape": 40}, "<-": {"learning_rate": 0.0}},
              {"type": "softmax", "<-": {"learning_rate": 0.0}}]
    cfg = dict(CFG, sample_shape=(100,), n_valid=0, minibatch_size=64)
    res = []
    for dev in (NumpyDevice(), B200Device()):
        l, wf = _build(dev, layers, cfg, epochs=1)
        for gd in wf.gds:
            gd.apply_gradient = False
        wf.loader.run()
        for f in wf.forwards:
            f.run()
        wf.evaluator.run()
        for gd in reversed(wf.gds):
            gd.run()
        out = {}
        for i, gd in enumerate(wf.gds):
            gd.gradient_weights.map_read()
            out["gw%d" % i] = gd.gradient_weights.mem.copy()
            gd.gradient_bias.map_read()
            out["gb%d" % i] = gd.gradient_bias.mem.copy()
        wf.forwards[-1].output.map_read()
        out["y"] = wf.forwards[-1].output.mem.copy()
        res.append(out)
        l.stop()
    for k in res[0]:
        a, b = res[0][k], res[1][k]
        denom = np.abs(a).max() + 1e-9
        assert np.abs(a - b).max() / denom < 0.12, k  # four bf16 layers deep


def test_float_precision_mode_runs_tf32_gemms_and_tracks_the_oracle_closely():
    """root.common.engine.precision_type = "float": fp32 activations + tcgen05 kind::tf32 GEMMs.  The weight
    trajectory must follow the numpy (fp32) oracle markedly closer than the default bf16 mode does."""
    from veles_b200.backends import NumpyDevice, B200Device
    from veles_b200.config import root
    l0, ref = _build(NumpyDevice(), MLP, CFG, epochs=2)
    ref.run()
    ref.forwards[0].weights.map_read()
    w_ref = ref.forwards[0].weights.mem.copy()
    errs = {}
    for mode in ("float", "bf16"):
        root.common.engine.precision_type = mode
        try:
            l1, wf = _build(B200Device(), MLP, CFG, epochs=2)
            assert bool(getattr(wf.forwards[0], "tf32_", False)) == (mode == "float")
            wf.run()
            wf.forwards[0].weights.map_read()
            errs[mode] = float(np.abs(wf.forwards[0].weights.mem - w_ref).max())
            assert wf.decision.epoch_metrics_history[-1]["n_err_pt"][1] <= 2.0
            l1.stop()
        finally:
            root.common.engine.precision_type = "bf16"
    assert errs["float"] < 0.02 and errs["float"] < errs["bf16"], errs
    l0.stop()


def test_partial_minibatches_match_numpy_oracle():
    """Dataset sizes that do not divide by the minibatch size: padded rows must not leak into errors / gradients."""
    from veles_b200.backends import NumpyDevice, B200Device
    cfg = dict(CFG, n_train=230, n_valid=70, minibatch_size=100)
    l0, ref = _build(NumpyDevice(), MLP, cfg, epochs=2)
    ref.run()
    l1, wf = _build(B200Device(), MLP, cfg, epochs=2)
    wf.run()
    for a, b in zip(ref.decision.epoch_metrics_history, wf.decision.epoch_metrics_history):
        assert a["n_err"] is None or abs(a["n_err_pt"][1] - b["n_err_pt"][1]) <= 5.0
        assert abs(a["loss"][1] - b["loss"][1]) <= 0.05 + 0.1 * abs(a["loss"][1])
        assert abs(a["loss"][2] - b["loss"][2]) <= 0.05 + 0.1 * abs(a["loss"][2])
    wf.forwards[0].weights.map_read()
    ref.forwards[0].weights.map_read()
    assert np.abs(wf.forwards[0].weights.mem - ref.forwards[0].weights.mem).max() < 0.05
    l0.stop()
    l1.stop()


def test_snapshot_from_device_resumes_on_device(tmp_path):
    """Pickle a workflow trained on the GPU (device tensors must not leak into the pickle), restore it, keep training on
    the GPU: the resumed run must continue from the restored weights, loader position and epoch counter."""
    import pickle
    from veles_b200.backends import B200Device
    from veles_b200.snapshotter import SnapshotterToFile
    l1, wf = _build(B200Device(), MLP, CFG, epochs=1)
    wf.run()
    snap = SnapshotterToFile(wf, directory=str(tmp_path), prefix="dev", compression="gz", time_interval=0)
    snap.suffix = "e1"
    snap.export()
    raw = open(snap.destination, "rb").read()
    assert b"torch" not in __import__("gzip").decompress(raw)      # host arrays only
    wf.forwards[0].weights.map_read()
    w_before = wf.forwards[0].weights.mem.copy()
    err_before = wf.decision.epoch_metrics_history[-1]["n_err_pt"][1]
    n_before = len(wf.decision.epoch_metrics_history)
    epoch_before = wf.decision.epoch_metrics_history[-1]["epoch"]
    l1.stop()
    back = SnapshotterToFile.import_(snap.destination)
    assert back.restored_from_snapshot
    assert np.array_equal(back.forwards[0].weights.mem, w_before)
    from veles_b200.dummy import DummyLauncher
    launcher = DummyLauncher()
    back.workflow = launcher
    back.decision.max_epochs = 3
    back.decision.complete <<= False
    back.initialize(device=B200Device())
    back.run()
    hist = back.decision.epoch_metrics_history
    assert hist[-1]["epoch"] == 3 and len(hist) == n_before + (3 - epoch_before)   # history restored, then extended
    assert [h["epoch"] for h in hist] == sorted(set(h["epoch"] for h in hist))
    assert hist[-1]["n_err_pt"][1] <= err_before + 1.0
    back.forwards[0].weights.map_read()
    assert np.abs(back.forwards[0].weights.mem - w_before).max() > 0
    launcher.stop()


def test_lstm_workflow_on_device_follows_numpy_oracle():
    """LSTM -> softmax trained through StandardWorkflow: device curve vs the numpy oracle (same data, same init)."""
    from veles_b200.backends import NumpyDevice, B200Device
    layers = [{"type": "lstm", "->": {"hidden_size": 32, "return_sequences": False},
               "<-": {"learning_rate": 0.05, "gradient_moment": 0.9}},
              {"type": "softmax", "<-": {"learning_rate": 0.05, "gradient_moment": 0.9}}]
    cfg = {"minibatch_size": 64, "n_train": 640, "n_valid": 128, "sample_shape": (6, 16), "noise": 0.5}
    l0, ref = _build(NumpyDevice(), layers, cfg, epochs=4)
    ref.run()
    l1, wf = _build(B200Device(), layers, cfg, epochs=4)
    wf.run()
    h_ref, h_dev = ref.decision.epoch_metrics_history, wf.decision.epoch_metrics_history
    assert len(h_ref) == len(h_dev)
    assert h_dev[-1]["loss"][2] < h_dev[1]["loss"][2]
    for a, b in zip(h_ref, h_dev):
        assert abs(a["loss"][1] - b["loss"][1]) <= 0.05 + 0.1 * abs(a["loss"][1]), (a, b)
    wf.forwards[0].weights.map_read()
    ref.forwards[0].weights.map_read()
    assert np.abs(wf.forwards[0].weights.mem - ref.forwards[0].weights.mem).max() < 0.05
    l0.stop()
    l1.stop()


def test_workflow_run_replays_cuda_graphs_and_matches_the_eager_walk():
    """``wf.run()`` (the scheduler of the CLI path) executes the steady-state forward and backward unit sequences as
    replayed CUDA graphs; the trajectory equals the one of the unit-by-unit walk."""
    from veles_b200.backends import B200Device
    from veles_b200.config import root
    conv = [{"type": "conv_str", "->": {"n_kernels": 16, "kx": 3, "ky": 3, "padding": (1, 1, 1, 1)},
             "<-": {"learning_rate": 0.02, "gradient_moment": 0.9}},
            {"type": "max_pooling", "->": {"kx": 2, "ky": 2, "sliding": (2, 2)}},
            {"type": "all2all_tanh", "->": {"output_sample_shape": 32}, "<-": {"learning_rate": 0.02}},
            {"type": "softmax", "<-": {"learning_rate": 0.02}}]
    cfg = {"minibatch_size": 64, "n_train": 640, "n_valid": 128, "sample_shape": (8, 8, 8), "noise": 1.0}
    results = []
    for replay in (True, False):
        root.common.engine.graph_replay = replay
        try:
            l, wf = _build(B200Device(), conv, cfg, epochs=3)
            chains = getattr(wf, "graph_chains_", [])
            assert bool(chains) == replay
            wf.run()
            if replay:
                assert [c.name for c in chains] == ["forward", "backward"]
                assert all(c.enabled for c in chains)
                assert chains[0].replays > 20 and chains[1].replays > 10, [(c.replays, c.captures) for c in chains]
                assert chains[1].captures <= 2
            wf.forwards[0].weights.map_read()
            wf.forwards[-1].weights.map_read()
            results.append((wf.decision.epoch_metrics_history, wf.forwards[0].weights.mem.copy(),
                            wf.forwards[-1].weights.mem.copy()))
            l.stop()
        finally:
            root.common.engine.graph_replay = True
    (h1, a1, b1), (h0, a0, b0) = results
    assert len(h1) == len(h0) >= 3
    # same kernels, same order; split-K reduce-adds land in a different order from run to run, hence not bit-equal
    for x, y in zip(h1, h0):
        assert all(abs(a - b) <= 2 for a, b in zip(x["n_err"], y["n_err"]))
    assert np.abs(a1 - a0).max() <= 1e-3 * max(1.0, np.abs(a0).max())
    assert np.abs(b1 - b0).max() <= 1e-3 * max(1.0, np.abs(b0).max())


@pytest.mark.parametrize("streaming", [False, True])
def test_input_gather_fused_with_space_to_depth_matches_the_two_kernel_path(streaming):
    """AlexNet's first layer reads its input through a space-to-depth transform; when it is the only reader of the
    loader's minibatch the loader gathers straight into that layout (``vb_gather_s2d``).  Same losses and weights as
    with the NHWC minibatch + separate transform, for the resident data set and for minibatches fed from the host."""
    import torch
    from veles_b200.backends import B200Device
    from veles_b200.config import root
    from veles_b200.trainer import Trainer
    rng = np.random.default_rng(5)
    fed = [(rng.integers(0, 256, size=(8, 227, 227, 3), dtype=np.uint8), rng.integers(0, 1000, size=8).astype(np.int32))
           for _ in range(3)]
    out = []
    for fuse in (True, False):
        root.common.engine.fuse_input_gather = fuse
        try:
            tr = Trainer("alexnet", minibatch_size=8, device=B200Device(), use_graph=False, streaming=streaming,
                         n_train=32, layers_kwargs={"dropout": 0.0})
            conv1 = tr.wf.forwards[0]
            sink = tr.wf.loader.__dict__.get("minibatch_sink_")
            assert (sink is not None) == fuse and (conv1.s2d_.get("sink") is not None) == fuse
            losses = []
            for data, labels in fed:
                if streaming:
                    losses.append(tr.step(torch.from_numpy(data).pin_memory(), torch.from_numpy(labels).pin_memory()))
                else:
                    tr.step_resident()
                    losses.append(tr.read_loss())
            assert sink is None or sink["fed"]
            conv1.weights.map_read()
            tr.wf.forwards[-1].weights.map_read()
            out.append((losses, conv1.weights.mem.copy(), tr.wf.forwards[-1].weights.mem.copy()))
            tr.close()
        finally:
            root.common.engine.fuse_input_gather = True
    (l1, a1, b1), (l0, a0, b0) = out
    assert np.allclose(l1, l0, rtol=2e-3, atol=1e-4), (l1, l0)
    assert np.abs(a1 - a0).max() <= 2e-3 * max(1e-3, np.abs(a0).max())
    assert np.abs(b1 - b0).max() <= 2e-3 * max(1e-3, np.abs(b0).max())


def test_bias_gradient_summed_by_lrn_backward_matches_the_column_sum_pass():
    """conv -> LRN: the LRN backward kernel accumulates the convolution's bias gradient while it writes that layer's
    err_output (StandardWorkflow._plan_bias_sum_fusion).  Same biases, momenta and losses as with the separate
    column-sum pass."""
    from veles_b200.backends import B200Device
    from veles_b200.config import root
    from veles_b200.trainer import Trainer
    out = []
    for fuse in (True, False):
        root.common.engine.fuse_bias_sum = fuse
        try:
            tr = Trainer("alexnet", minibatch_size=8, device=B200Device(), use_graph=False, n_train=32,
                         layers_kwargs={"dropout": 0.0})
            planned = [g.name for g in tr.wf.gds if getattr(g, "bias_sum_for", None) is not None]
            assert (len(planned) == 2) == fuse, planned          # norm1 and norm2
            losses = []
            for _ in range(4):
                tr.step_resident()
                losses.append(tr.read_loss())
            state = []
            for f, g in zip(tr.wf.forwards[:4], tr.wf.gds[:4]):
                if getattr(f, "HAS_WEIGHTS", False):
                    f.bias.map_read()
                    g.gradient_bias_with_moment.map_read()
                    state.append((f.bias.mem.copy(), g.gradient_bias_with_moment.mem.copy()))
            out.append((losses, state))
            tr.close()
        finally:
            root.common.engine.fuse_bias_sum = True
    (l1, s1), (l0, s0) = out
    assert np.allclose(l1, l0, rtol=2e-3, atol=1e-4), (l1, l0)
    assert len(s1) == len(s0) >= 2
    for (b1, v1), (b0, v0) in zip(s1, s0):
        assert np.abs(v0).max() > 0
        assert np.abs(b1 - b0).max() <= 1e-2 * np.abs(b0 - b0.mean()).max() + 1e-2 * np.abs(v0).max()
        assert np.abs(v1 - v0).max() <= 2e-2 * np.abs(v0).max()


def test_full_alexnet_pass_matches_numpy_oracle():
    """The whole flagship network once, every kernel of the benchmark at its real shape (227x227x3 input, grouped
    convolutions, LRN, overlapping pooling, 4096-wide fully connected layers; batch 4) against the numpy backend:
    * forward: the output of every layer;
    * backward, layer by layer: each device unit is handed the ORACLE's err_output and must reproduce the oracle's
      err_input and weight / bias gradients (no accumulation of differences from layer to layer).  Max-pooling units
      are compared by gradient mass only: this random-init network has bias-dominated activations, so bf16 rounding
      makes many window maxima tie and a quarter of the gradient is routed to a different (equally valid) position."""
    from veles_b200.backends import NumpyDevice, B200Device
    from veles_b200.trainer import Trainer
    from veles_b200.znicz.gd_pooling import GDPooling

    def build(dev):
        tr = Trainer("alexnet", minibatch_size=4, device=dev, use_graph=False, n_train=8, seed=11,
                     layers_kwargs={"dropout": 0.0})
        for gd in tr.wf.gds:
            gd.apply_gradient = False
        tr.wf.loader.run()
        for f in tr.wf.forwards:
            f.run()
        tr.wf.evaluator.run()
        return tr

    def grab(arr):
        arr.map_read()
        return np.array(arr.mem, dtype=np.float32).copy()

    # oracle pass
    tr = build(NumpyDevice())
    wf = tr.wf
    ref = {"y": [grab(f.output).reshape(4, -1) for f in wf.forwards], "eo": {}, "ei": {}, "gw": {}, "gb": {}}
    for i in range(len(wf.gds) - 1, -1, -1):
        gd = wf.gds[i]
        ref["eo"][i] = grab(gd.err_output)
        gd.run()
        if i > 0 and gd.err_input is not None and gd.err_input:
            ref["ei"][i] = grab(gd.err_input)
        if getattr(wf.forwards[i], "HAS_WEIGHTS", False):
            ref["gw"][i] = grab(gd.gradient_weights)
            ref["gb"][i] = grab(gd.gradient_bias)
    tr.close()

    def fro(a, b):
        return float(np.linalg.norm((a - b).ravel()) / (np.linalg.norm(a.ravel()) + 1e-20))

    tr = build(B200Device())
    wf = tr.wf
    errors = {}
    for i, f in enumerate(wf.forwards):
        y = grab(f.output).reshape(4, -1)
        assert np.isfinite(y).all()
        errors["y%02d_%s" % (i, type(f).__name__)] = float(np.abs(y - ref["y"][i]).max() / (np.abs(ref["y"][i]).max() + 1e-12))
    top = len(wf.gds) - 1
    for i in range(top, -1, -1):
        gd = wf.gds[i]
        if i < top:                                   # the top unit keeps the evaluator's err_output
            gd.err_output.map_invalidate()
            gd.err_output.mem[...] = ref["eo"][i].reshape(gd.err_output.mem.shape)
        gd.run()
        name = "%02d_%s" % (i, type(gd).__name__)
        if i in ref["ei"]:
            ei = grab(gd.err_input)
            assert np.isfinite(ei).all(), name
            if isinstance(gd, GDPooling):
                errors["ei" + name + "_mass"] = abs(float(np.linalg.norm(ei)) / float(np.linalg.norm(ref["ei"][i])) - 1.0)
            else:
                errors["ei" + name] = fro(ref["ei"][i].ravel(), ei.ravel())
        if i in ref["gw"]:
            errors["gw" + name] = fro(ref["gw"][i].ravel(), grab(gd.gradient_weights).ravel())
            errors["gb" + name] = fro(ref["gb"][i].ravel(), grab(gd.gradient_bias).ravel())
    tr.close()
    import json, os
    if os.path.isdir("gpurun_out"):
        with open("gpurun_out/alexnet_oracle_errors.json", "w") as fdbg:
            json.dump(errors, fdbg, indent=1)
    assert sum(k.startswith("gw") for k in errors) == 8
    fwd = {k: v for k, v in errors.items() if k.startswith("y")}
    assert max(fwd.values()) < 0.02, fwd            # bf16 through eight layers: half a per cent of each tensor's maximum
    # relative Frobenius error per tensor.  Most land at 0.2-0.4 % (bf16 operands); units that multiply by a strict-ReLU
    # derivative reach 2-4 %: an activation that is a rounding error away from zero passes its gradient on one side
    # and blocks it on the other, and a fraction p of such elements costs sqrt(p)
    bad = {k: v for k, v in errors.items() if not k.startswith("y") and v > 0.06}
    assert not bad, bad
    tight = [v for k, v in errors.items() if k.startswith("gw")]
    assert sorted(tight)[len(tight) // 2] < 0.005, tight


def test_device_tensor_reflects_pending_host_writes():
    """``Array.devmem`` flushes a pending host write: a device unit that reads its input without an explicit unmap()
    (LRN, dropout, activations...) still sees what a host-side producer wrote."""
    from veles_b200.backends import B200Device
    from veles_b200.memory import Array
    dev = B200Device()
    a = Array(np.arange(12, dtype=np.float32).reshape(3, 4))
    a.initialize(dev)
    assert float(a.devmem.sum()) == 66.0
    a.map_write()
    a.mem[...] = 1.0
    assert float(a.devmem.sum()) == 12.0            # no unmap() in between
    a.devmem.mul_(2.0)
    a.mark_device_written()
    a.map_read()
    assert float(a.mem.sum()) == 24.0
    a.map_invalidate()
    a.mem[...] = 3.0
    assert float(a.devmem[0, 0]) == 3.0


def test_pipelined_step_on_device_returns_each_loss_one_call_late():
    """Trainer.step_pipelined on the device: the loss of step i is copied to pinned memory asynchronously and returned by
    call i + 1 (flush_loss for the last); same values as the synchronous step().  (Same model / feeding as the
    streaming case of the input-gather test above.)"""
    import torch
    from veles_b200.backends import B200Device
    from veles_b200.trainer import Trainer
    rng = np.random.default_rng(6)
    fed = [(torch.from_numpy(rng.integers(0, 256, size=(8, 227, 227, 3), dtype=np.uint8)).pin_memory(),
            torch.from_numpy(rng.integers(0, 1000, size=8).astype(np.int32)).pin_memory()) for _ in range(4)]
    seqs = []
    for pipelined in (False, True):
        tr = Trainer("alexnet", minibatch_size=8, device=B200Device(), use_graph=False, streaming=True, n_train=32,
                     layers_kwargs={"dropout": 0.0})
        if pipelined:
            got = [tr.step_pipelined(d, l) for d, l in fed]
            assert got[0] is None
            got = got[1:] + [tr.flush_loss()]
            assert tr.flush_loss() is None
        else:
            got = [tr.step(d, l) for d, l in fed]
        assert all(np.isfinite(v) and v > 0 for v in got)
        seqs.append(got)
        tr.close()
    assert np.allclose(seqs[0], seqs[1], rtol=2e-3), seqs
