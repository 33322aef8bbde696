"""Extra Znicz layers: cutter, channel split/merge, deconv (gradient check + conv adjointness), depooling,
zero_filter, resizable / rprop all2all, Kohonen, RBM, LR policies, mcdnnic topology parsing."""
import numpy as np
import pytest

from veles_b200.backends import NumpyDevice
from veles_b200.dummy import DummyWorkflow
from veles_b200.memory import Array
from veles_b200.znicz.standard_workflow import LayerRegistry, parse_mcdnnic_topology


def _fwd(tpe, x, **kw):
    wf = DummyWorkflow()
    u = LayerRegistry.forwards()[tpe](wf, **kw)
    u.input = Array(np.asarray(x, dtype=np.float32))
    return wf, u


def _gd(wf, tpe, fwd, eo, **kw):
    g = LayerRegistry.gds()[tpe](wf, forward=fwd, apply_gradient=False, **kw)
    g.input, g.output = fwd.input, fwd.output
    if getattr(fwd, "HAS_WEIGHTS", False):
        g.weights, g.bias = fwd.weights, fwd.bias
    g.err_output = Array(np.asarray(eo, dtype=np.float32))
    assert not g.initialize(device=NumpyDevice())
    g.run()
    return g


def test_cutter_and_backward():
    x = np.arange(2 * 6 * 7 * 3).reshape(2, 6, 7, 3)
    wf, f = _fwd("cutter", x, padding=(1, 2, 3, 1))
    assert not f.initialize(device=NumpyDevice())
    f.run()
    assert f.output.shape == (2, 3, 3, 3) and np.array_equal(f.output.mem, x[:, 2:5, 1:4, :])
    g = _gd(wf, "cutter", f, np.ones(f.output.shape))
    assert g.err_input.mem.sum() == f.output.size and g.err_input.mem[:, 0].sum() == 0


def test_channel_splitter_merger_roundtrip():
    x = np.random.rand(2, 4, 5, 3)
    wf, s = _fwd("channel_splitter", x)
    s.initialize(device=NumpyDevice())
    s.run()
    assert s.output.shape == (2, 3, 4, 5)
    m = LayerRegistry.forwards()["channel_merger"](wf)
    m.input = s.output
    m.initialize(device=NumpyDevice())
    m.run()
    assert np.allclose(m.output.mem, x)
    g = _gd(wf, "channel_splitter", s, np.random.rand(2, 3, 4, 5))
    assert g.err_input.shape == (2, 4, 5, 3)


def test_deconv_is_adjoint_of_conv_and_gradients_check():
    rng = np.random.RandomState(0)
    x = rng.rand(2, 9, 8, 3).astype(np.float32)
    wf, conv = _fwd("conv", x, n_kernels=4, kx=3, ky=3, padding=(1, 1, 1, 1), sliding=(2, 2), include_bias=False)
    conv.initialize(device=NumpyDevice())
    conv.run()
    y = rng.rand(*conv.output.shape).astype(np.float32)
    dec = LayerRegistry.forwards()["deconv"](wf, n_kernels=4, kx=3, ky=3, padding=(1, 1, 1, 1), sliding=(2, 2),
                                             output_shape_source=conv.input)
    dec.input = Array(y)
    dec.weights = conv.weights                      # tied weights
    assert not dec.initialize(device=NumpyDevice())
    dec.run()
    # <conv(x), y> == <x, deconv(y)>
    assert abs((conv.output.mem * y).sum() - (x * dec.output.mem).sum()) < 1e-2
    R = rng.rand(*dec.output.shape).astype(np.float32)
    g = _gd(wf, "deconv", dec, R)

    def loss():
        dec.run()
        return float((dec.output.mem.astype(np.float64) * R).sum())
    for arr, ana in ((dec.input.mem, g.err_input.mem), (dec.weights.mem, g.gradient_weights.mem)):
        flat = arr.reshape(-1)
        for i in rng.choice(flat.size, 6, replace=False):
            saved = flat[i]
            flat[i] = saved + 1e-2
            up = loss()
            flat[i] = saved - 1e-2
            dn = loss()
            flat[i] = saved
            assert abs((up - dn) / 2e-2 - ana.reshape(-1)[i]) < 2e-2 * max(1.0, abs(ana.reshape(-1)[i]))


def test_depooling_inverts_max_pooling_positions():
    x = np.random.permutation(2 * 6 * 6 * 2).reshape(2, 6, 6, 2).astype(np.float32)
    wf, pool = _fwd("max_pooling", x, kx=2, ky=2, sliding=(2, 2))
    pool.initialize(device=NumpyDevice())
    pool.run()
    dep = LayerRegistry.forwards()["depooling"](wf)
    dep.input = pool.output
    dep.output_offset = pool.input_offset
    dep.output_shape_source = pool.input
    assert not dep.initialize(device=NumpyDevice())
    dep.run()
    nz = dep.output.mem != 0
    assert nz.sum() == pool.output.size and np.array_equal(dep.output.mem[nz], x[nz])
    g = _gd(wf, "depooling", dep, x)
    assert np.array_equal(g.err_input.mem, pool.output.mem)


def test_zero_filter_keeps_block_diagonal_weights():
    wf, fc = _fwd("all2all", np.random.rand(3, 8), output_sample_shape=4)
    fc.initialize(device=NumpyDevice())
    z = LayerRegistry.forwards()["zero_filter"](wf, grouping=2, target=fc)
    z.input = fc.output
    assert not z.initialize(device=NumpyDevice())
    z.run()
    w = fc.weights.mem
    assert (w[:2, 4:] == 0).all() and (w[2:, :4] == 0).all() and (w[:2, :4] != 0).any()


def test_resizable_all2all_keeps_old_neurons():
    wf, fc = _fwd("all2all_resizable", np.random.rand(3, 5), output_sample_shape=4)
    fc.initialize(device=NumpyDevice())
    old = fc.weights.mem.copy()
    assert fc.resize(6) and not fc.resize(6)
    fc.initialize(device=NumpyDevice())
    fc.run()
    assert fc.output.shape == (3, 6) and np.array_equal(fc.weights.mem[:4], old)


def test_rprop_reduces_loss():
    rng = np.random.RandomState(1)
    x = rng.rand(20, 6).astype(np.float32)
    target = x.dot(rng.rand(6, 3)).astype(np.float32)
    wf, fc = _fwd("rprop_all2all", x, output_sample_shape=3)
    fc.initialize(device=NumpyDevice())
    g = LayerRegistry.gds()["rprop_all2all"](wf, forward=fc, need_err_input=False)
    g.input, g.output, g.weights, g.bias = fc.input, fc.output, fc.weights, fc.bias
    g.err_output = Array(np.zeros((20, 3), dtype=np.float32))
    g.initialize(device=NumpyDevice())
    losses = []
    for _ in range(60):
        fc.run()
        err = fc.output.mem - target
        losses.append(float((err ** 2).mean()))
        g.err_output.mem[...] = err / 20
        g.run()
    assert losses[-1] < 0.05 * losses[0]


def test_kohonen_and_rbm_learn():
    from veles_b200.znicz.kohonen import KohonenTrainer, KohonenForward
    from veles_b200.znicz.rbm import RBM
    rng = np.random.RandomState(2)
    centers = rng.rand(4, 5) * 4
    x = (centers[rng.randint(0, 4, 200)] + 0.05 * rng.randn(200, 5)).astype(np.float32)
    wf = DummyWorkflow()
    tr = KohonenTrainer(wf, shape=(3, 3))
    tr.input = Array(x)
    tr.initialize(device=NumpyDevice())
    for _ in range(80):
        tr.run()
    fw = KohonenForward(wf)
    fw.input, fw.weights = tr.input, tr.weights
    fw.initialize(device=NumpyDevice())
    fw.run()
    w = tr.weights.mem[fw.output.mem]
    assert np.sqrt(((x - w) ** 2).sum(1)).mean() < 0.6
    rbm = RBM(wf, n_hidden=16, learning_rate=0.2)
    rbm.input = Array((rng.rand(100, 12) < 0.5).astype(np.float32) * np.array([1] * 6 + [0] * 6, dtype=np.float32))
    rbm.initialize(device=NumpyDevice())
    errs = []
    for _ in range(150):
        rbm.run()
        errs.append(rbm.reconstruction_error)
    assert errs[-1] < errs[0] * 0.7


def test_lr_policies_and_mcdnnic():
    from veles_b200.znicz.lr_adjust import ExpPolicy, StepExpPolicy, InvAdjustPolicy, ArbitraryStepPolicy, FixedAdjustPolicy
    assert FixedAdjustPolicy(0.1)(50) == 0.1
    assert abs(ExpPolicy(0.1, gamma=0.5)(2) - 0.025) < 1e-12
    assert abs(StepExpPolicy(0.1, gamma=0.1, step=10)(25) - 0.001) < 1e-12
    assert abs(InvAdjustPolicy(0.1, gamma=1.0, pow_ratio=1.0)(9) - 0.01) < 1e-12
    p = ArbitraryStepPolicy(lrs_with_lengths=[(0.1, 3), (0.01, 2)])
    assert [p(i) for i in range(7)] == [0.1, 0.1, 0.1, 0.01, 0.01, 0.01, 0.01]
    layers = parse_mcdnnic_topology("12x256x256-32C4-MP2-64C4-MP3-32N-4N")
    assert [l["type"] for l in layers] == ["conv_tanh", "max_pooling", "conv_tanh", "max_pooling", "all2all_tanh", "softmax"]
    assert layers[0]["->"]["n_kernels"] == 32 and layers[3]["->"]["sliding"] == (3, 3)
    assert layers[-1]["->"]["output_sample_shape"] == 4


def test_recurrent_layers_train_in_standard_workflow():
    """rnn / lstm layer types are reachable from a StandardWorkflow config and learn a sequence task (numpy oracle)."""
    from veles_b200 import prng
    from veles_b200.config import root
    from veles_b200.dummy import DummyLauncher
    from veles_b200.backends import NumpyDevice
    from veles_b200.znicz.standard_workflow import StandardWorkflow
    import veles_b200.loader.synthetic  # noqa: F401
    root.common.disable.snapshotting = True
    cfg = {"minibatch_size": 50, "n_train": 500, "n_valid": 100, "sample_shape": (6, 8), "noise": 0.5}
    for tpe in ("lstm", "rnn"):
        prng.get().seed(3)
        layers = [{"type": tpe, "->": {"hidden_size": 16, "return_sequences": False},
                   "<-": {"learning_rate": 0.05, "gradient_moment": 0.9}},
                  {"type": "softmax", "<-": {"learning_rate": 0.05, "gradient_moment": 0.9}}]
        launcher = DummyLauncher()
        wf = StandardWorkflow(launcher, layers=layers, loader_name="synthetic_fullbatch", loader_config=cfg,
                              loss_function="softmax", decision_config={"max_epochs": 12})
        wf.initialize(device=NumpyDevice())
        assert wf.forwards[0].weights.shape == ((4 if tpe == "lstm" else 1) * 16, 8 + 16)
        wf.run()
        hist = wf.decision.epoch_metrics_history
        assert hist[-1]["n_err_pt"][1] < 0.75 * hist[0]["n_err_pt"][1], (tpe, hist)
        assert hist[-1]["loss"][2] < hist[1]["loss"][2]
        launcher.stop()


def test_stacked_recurrent_layers_and_sequence_outputs():
    import numpy as np
    from veles_b200.backends import NumpyDevice
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.memory import Array
    from veles_b200.znicz.recurrent import LSTM, RNN, GDLSTM, GDRNN
    rng = np.random.RandomState(0)
    wf = DummyWorkflow()
    x = rng.uniform(-1, 1, (4, 5, 3)).astype(np.float32)
    l1 = LSTM(wf, hidden_size=6)
    l1.input = Array(x)
    assert not l1.initialize(device=NumpyDevice())
    assert np.allclose(l1.bias.mem[6:12] - 1.0, l1.bias.mem[6:12] - 1.0) and l1.bias.mem[6:12].mean() > 0.5
    l1.run()
    assert l1.output.shape == (4, 5, 6)
    l2 = RNN(wf, hidden_size=2, return_sequences=False)
    l2.input = l1.output
    assert not l2.initialize(device=NumpyDevice())
    l2.run()
    assert l2.output.shape == (4, 2) and np.all(np.abs(l2.output.mem) <= 1)
    g2 = GDRNN(wf, forward=l2, learning_rate=0.1, weights_decay=0.0)
    g2.input, g2.output, g2.weights, g2.bias = l2.input, l2.output, l2.weights, l2.bias
    g2.err_output = Array(np.ones((4, 2), np.float32))
    assert not g2.initialize(device=NumpyDevice())
    w_before = l2.weights.mem.copy()
    g2.run()
    assert g2.err_input.shape == (4, 5, 6) and np.abs(l2.weights.mem - w_before).max() > 0
    g1 = GDLSTM(wf, forward=l1, learning_rate=0.1, weights_decay=0.0, need_err_input=False)
    g1.input, g1.output, g1.weights, g1.bias = l1.input, l1.output, l1.weights, l1.bias
    g1.err_output = g2.err_input
    assert not g1.initialize(device=NumpyDevice())
    g1.run()
    assert g1.gradient_weights.shape == (24, 9)
    # a causal layer: output at step t must not depend on inputs after t
    x2 = x.copy()
    x2[:, 3:, :] += 1.0
    l1b = LSTM(wf, hidden_size=6)
    l1b.input = Array(x2)
    l1b.initialize(device=NumpyDevice())
    l1b.weights.mem[...] = l1.weights.mem
    l1b.bias.mem[...] = l1.bias.mem
    l1.run()
    l1b.run()
    assert np.allclose(l1.output.mem[:, :3], l1b.output.mem[:, :3]) and \
        not np.allclose(l1.output.mem[:, 3:], l1b.output.mem[:, 3:])


def test_recurrent_workflow_snapshot_roundtrip(tmp_path):
    """A workflow with an LSTM layer pickles (cell-state caches are not part of the snapshot) and resumes training."""
    import numpy as np
    from veles_b200 import prng
    from veles_b200.backends import NumpyDevice
    from veles_b200.config import root
    from veles_b200.dummy import DummyLauncher
    from veles_b200.models import build_workflow
    from veles_b200.snapshotter import SnapshotterToFile
    root.common.disable.snapshotting = True
    prng.get().seed(5)
    wf = build_workflow("lstm_classifier", minibatch_size=32, n_train=128, n_valid=32, max_epochs=1)
    wf.initialize(device=NumpyDevice())
    wf.run()
    snap = SnapshotterToFile(wf, directory=str(tmp_path), prefix="seq", compression="gz", time_interval=0)
    snap.suffix = "e1"
    snap.export()
    w_before = wf.forwards[0].weights.mem.copy()
    wf.launcher.stop()
    back = SnapshotterToFile.import_(snap.destination)
    assert back.restored_from_snapshot and back.forwards[0].np_cache_ is None
    assert np.array_equal(back.forwards[0].weights.mem, w_before)
    assert back.forwards[0].hidden_size == 128 and not back.forwards[0].return_sequences
    launcher = DummyLauncher()
    back.workflow = launcher
    back.decision.max_epochs = 2
    back.decision.complete <<= False
    back.initialize(device=NumpyDevice())
    back.run()
    assert back.decision.epoch_metrics_history[-1]["epoch"] == 2
    assert np.abs(back.forwards[0].weights.mem - w_before).max() > 0
    launcher.stop()


def _one_gd_step(gd_kwargs, device=None, steps=2):
    """Train a 12-8-4 MLP for ``steps`` minibatches; returns the first layer's weights."""
    from veles_b200 import prng
    from veles_b200.backends import NumpyDevice
    from veles_b200.config import root
    from veles_b200.dummy import DummyLauncher
    from veles_b200.znicz.standard_workflow import StandardWorkflow
    import veles_b200.loader.synthetic  # noqa: F401
    root.common.disable.snapshotting = True
    prng.get().seed(3)
    layers = [{"type": "all2all_tanh", "->": {"output_sample_shape": 8}, "<-": dict(gd_kwargs)},
              {"type": "softmax", "<-": {"learning_rate": 0.0}}]
    wf = StandardWorkflow(DummyLauncher(), layers=layers, loader_name="synthetic_fullbatch",
                          loader_config={"minibatch_size": 16, "n_train": 64, "n_valid": 0, "sample_shape": (12,),
                                         "n_classes": 4},
                          loss_function="softmax", decision_config={"max_epochs": 1})
    wf.initialize(device=device or NumpyDevice())
    w0 = None
    for _ in range(steps):
        wf.loader.run()
        for f in wf.forwards:
            f.run()
        wf.evaluator.run()
        if w0 is None:
            wf.forwards[0].weights.map_read()
            w0 = wf.forwards[0].weights.mem.copy()
        for gd in reversed(wf.gds):
            gd.run()
    wf.forwards[0].weights.map_read()
    w = wf.forwards[0].weights.mem.copy()
    gd0 = wf.gds[0]
    gd0.gradient_weights_with_moment.map_read()
    vel = gd0.gradient_weights_with_moment.mem.copy()
    wf.launcher.stop()
    return w0, w, vel


def test_gd_variant_switches_and_factor_ortho_have_semantics():
    import numpy as np
    import pytest
    base = {"learning_rate": 0.1, "gradient_moment": 0.9, "weights_decay": 0.0}
    w0, w_full, v_full = _one_gd_step(base)
    # no momentum term: two plain SGD steps, the velocity holds the last step only
    _, w_nomom, v_nomom = _one_gd_step(dict(base, variant_moment_gradient=False))
    _, w_sgd, _ = _one_gd_step(dict(base, gradient_moment=0.0))
    np.testing.assert_allclose(w_nomom, w_sgd, rtol=1e-6, atol=1e-7)
    assert np.abs(w_nomom - w_full).max() > 1e-5
    # no gradient step: with zero initial velocity the weights never move
    _, w_coast, v_coast = _one_gd_step(dict(base, variant_gradient=False))
    np.testing.assert_array_equal(w_coast, w0)
    assert not v_coast.any()
    # factor_ortho: the first step differs from the plain one by exactly -lr * f / n * (column sums - w)
    f = 0.5
    w0o, w_ortho, _ = _one_gd_step(dict(base, factor_ortho=f, gradient_moment=0.0), steps=1)
    _, w_plain, _ = _one_gd_step(dict(base, gradient_moment=0.0), steps=1)
    expect = w_plain - 0.1 * f / w0o.shape[0] * (w0o.sum(axis=0, keepdims=True) - w0o)
    np.testing.assert_allclose(w_ortho, expect, rtol=1e-5, atol=1e-6)
    # the undocumented option is refused, not ignored
    with pytest.raises(ValueError, match="fast_learning_rate"):
        _one_gd_step(dict(base, fast_learning_rate=0.5), steps=1)


@pytest.mark.gpu
def test_gd_options_on_device_match_the_numpy_oracle():
    import numpy as np
    from veles_b200.backends import B200Device
    for kw in ({"learning_rate": 0.1, "gradient_moment": 0.9, "factor_ortho": 0.5},
               {"learning_rate": 0.1, "gradient_moment": 0.9, "variant_moment_gradient": False},
               {"learning_rate": 0.1, "gradient_moment": 0.9, "variant_gradient": False}):
        _, w_ref, v_ref = _one_gd_step(kw, steps=3)
        _, w_dev, v_dev = _one_gd_step(kw, device=B200Device(), steps=3)
        assert np.abs(w_dev - w_ref).max() < 3e-2 * max(np.abs(w_ref).max(), 1e-3), kw


def test_bias_sum_fusion_plan_pairs_lrn_units_with_their_convolutions():
    """StandardWorkflow plans conv -> LRN pairs whose LRN backward unit also produces the convolution's bias-gradient
    sums (device path); the plan is inert on the numpy backend and can be switched off."""
    from veles_b200.backends import NumpyDevice
    from veles_b200.config import root
    from veles_b200.trainer import Trainer
    from veles_b200.znicz.gd_conv import GradientDescentConv
    from veles_b200.znicz.normalization import LRNormalizerBackward
    for fuse in (True, False):
        root.common.engine.fuse_bias_sum = fuse
        try:
            tr = Trainer("alexnet", minibatch_size=2, device=NumpyDevice(), use_graph=False, n_train=4,
                         layers_kwargs={"dropout": 0.0})
            gds = tr.wf.gds
            pairs = [(i, gds.index(g.bias_sum_for)) for i, g in enumerate(gds) if getattr(g, "bias_sum_for", None) is not None]
            assert pairs == ([(1, 0), (4, 3)] if fuse else [])
            for i, j in pairs:
                assert isinstance(gds[i], LRNormalizerBackward) and isinstance(gds[j], GradientDescentConv)
                assert gds[i].prev_activation and gds[j].derivative_applied      # the sums are of the masked gradient
            tr.step_resident()                                                    # the oracle path ignores the plan
            assert tr.wf.loader.__dict__.get("minibatch_sink_") is None           # (and gathers into minibatch_data)
            tr.close()
        finally:
            root.common.engine.fuse_bias_sum = True
