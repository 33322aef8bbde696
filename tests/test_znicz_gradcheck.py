"""Numeric gradient checks of every forward/backward pair on the numpy oracle backend
(the reference ships NumDiff for exactly this — veles/numpy_ext.py:116-132 — its Znicz tests are absent)."""
import numpy as np
import pytest

from veles_b200.backends import NumpyDevice
from veles_b200.dummy import DummyWorkflow
from veles_b200.memory import Array
from veles_b200.znicz.standard_workflow import LayerRegistry


def _pair(tpe, x, fwd_kw=None, prev_activation=0):
    wf = DummyWorkflow()
    fwd = LayerRegistry.forwards()[tpe](wf, **(fwd_kw or {}))
    fwd.input = Array(x.astype(np.float32))
    dev = NumpyDevice()
    assert not fwd.initialize(device=dev)
    fwd.run()
    gd = LayerRegistry.gds()[tpe](wf, forward=fwd, learning_rate=0.0, weights_decay=0.0, apply_gradient=False,
                                  prev_activation=prev_activation)
    gd.input, gd.output = fwd.input, fwd.output
    if getattr(fwd, "HAS_WEIGHTS", False):
        gd.weights, gd.bias = fwd.weights, fwd.bias
    return wf, fwd, gd, dev


def _loss(fwd, R):
    fwd.run()
    return float((fwd.output.mem.astype(np.float64) * R).sum())


def _numdiff(arr, fwd, R, idxs, h=1e-3):
    out = []
    flat = arr.reshape(-1)
    for i in idxs:
        saved = flat[i]
        vals = []
        for d in (2 * h, h, -h, -2 * h):
            flat[i] = saved + d
            vals.append(_loss(fwd, R))
        flat[i] = saved
        out.append((-vals[0] + 8 * vals[1] - 8 * vals[2] + vals[3]) / (12 * h))
    return np.array(out)


CASES = [
    ("all2all", (5, 12), {"output_sample_shape": 7}),
    ("all2all_tanh", (5, 12), {"output_sample_shape": 7}),
    ("all2all_relu", (5, 12), {"output_sample_shape": 7}),
    ("all2all_sigmoid", (5, 12), {"output_sample_shape": 7}),
    ("all2all_str", (5, 12), {"output_sample_shape": 7}),
    ("conv", (2, 7, 8, 3), {"n_kernels": 4, "kx": 3, "ky": 3, "padding": (1, 1, 1, 1), "sliding": (1, 1)}),
    ("conv_tanh", (2, 9, 8, 4), {"n_kernels": 6, "kx": 3, "ky": 2, "padding": (1, 0, 2, 1), "sliding": (2, 1),
                                 "grouping": 2}),
    ("conv_sigmoid", (2, 11, 11, 2), {"n_kernels": 3, "kx": 5, "ky": 5, "padding": (0, 0, 0, 0), "sliding": (3, 3)}),
    ("max_pooling", (2, 7, 7, 3), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("maxabs_pooling", (2, 6, 7, 3), {"kx": 2, "ky": 2, "sliding": (2, 2)}),
    ("avg_pooling", (2, 7, 6, 3), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("norm", (2, 3, 3, 9), {"alpha": 0.05, "beta": 0.75, "n": 5, "k": 2}),
    ("activation_tanh", (4, 10), {}),
    ("activation_sigmoid", (4, 10), {}),
    ("activation_relu", (4, 10), {}),
    ("activation_log", (4, 10), {}),
    ("activation_tanhlog", (4, 10), {}),
    ("activation_sincos", (4, 10), {}),
    ("activation_mul", (4, 10), {"factor": 0.3}),
    ("rnn", (3, 5, 4), {"hidden_size": 6}),
    ("rnn", (3, 5, 4), {"hidden_size": 6, "return_sequences": False}),
    ("lstm", (3, 5, 4), {"hidden_size": 6}),
    ("lstm", (2, 4, 3), {"hidden_size": 5, "return_sequences": False}),
    ("lstm", (3, 20), {"hidden_size": 4, "sequence_length": 5}),
]


@pytest.mark.parametrize("tpe,shape,kw", CASES, ids=[c[0] + str(i) for i, c in enumerate(CASES)])
def test_backward_matches_numeric_derivative(tpe, shape, kw):
    rng = np.random.RandomState(hash(tpe) % 1000)
    x = rng.uniform(-1.5, 1.5, size=shape)
    if "pooling" in tpe:  # keep argmax away from ties / the finite-difference step
        x = rng.permutation(x.size).reshape(shape) * 0.05 - 1.0
    wf, fwd, gd, dev = _pair(tpe, x, kw)
    R = rng.uniform(-1, 1, size=fwd.output.shape)
    gd.err_output = Array(R.astype(np.float32))
    assert not gd.initialize(device=dev)
    gd.run()
    idxs = rng.choice(x.size, size=min(12, x.size), replace=False)
    num = _numdiff(fwd.input.mem, fwd, R, idxs)
    ana = gd.err_input.mem.reshape(-1)[idxs]
    assert np.allclose(ana, num, atol=3e-3, rtol=2e-2), (ana, num)
    if getattr(fwd, "HAS_WEIGHTS", False):
        widx = rng.choice(fwd.weights.size, size=min(12, fwd.weights.size), replace=False)
        numw = _numdiff(fwd.weights.mem, fwd, R, widx)
        assert np.allclose(gd.gradient_weights.mem.reshape(-1)[widx], numw, atol=3e-3, rtol=2e-2)
        bidx = np.arange(min(4, fwd.bias.size))
        numb = _numdiff(fwd.bias.mem, fwd, R, bidx)
        assert np.allclose(gd.gradient_bias.mem.reshape(-1)[bidx], numb, atol=3e-3, rtol=2e-2)


def test_fused_previous_derivative_equals_two_step_backward():
    """gd[i] with prev_activation == separate derivative pass in gd[i-1] (the fusion protocol)."""
    rng = np.random.RandomState(5)
    x = np.tanh(rng.uniform(-1, 1, size=(6, 10))) * 1.7159  # looks like a tanh layer output
    for tpe, kw in (("all2all", {"output_sample_shape": 5}),
                    ("max_pooling", {"kx": 2, "ky": 2, "sliding": (2, 2)})):
        xx = x if tpe == "all2all" else rng.uniform(0.1, 1.5, size=(2, 4, 4, 3))
        wf, fwd, gd, dev = _pair(tpe, xx, kw, prev_activation=1)
        R = rng.uniform(-1, 1, size=fwd.output.shape).astype(np.float32)
        gd.err_output = Array(R.copy())
        gd.initialize(device=dev)
        gd.run()
        fused = gd.err_input.mem.copy()
        wf2, fwd2, gd2, _ = _pair(tpe, xx, kw, prev_activation=0)
        if getattr(fwd, "HAS_WEIGHTS", False):
            fwd2.weights.mem[...] = fwd.weights.mem
            fwd2.bias.mem[...] = fwd.bias.mem
            fwd2.run()
        gd2.err_output = Array(R.copy())
        gd2.initialize(device=dev)
        gd2.run()
        from veles_b200.znicz.nn_units import act_deriv_numpy
        two_step = gd2.err_input.mem * act_deriv_numpy(1, xx.astype(np.float32))
        assert np.allclose(fused, two_step, atol=1e-5)


def test_softmax_evaluator_gradient_is_cross_entropy_derivative():
    from veles_b200.znicz.all2all import All2AllSoftmax
    from veles_b200.znicz.evaluator import EvaluatorSoftmax
    rng = np.random.RandomState(1)
    wf = DummyWorkflow()
    x = rng.uniform(-1, 1, size=(6, 9)).astype(np.float32)
    sm = All2AllSoftmax(wf, output_sample_shape=4)
    sm.input = Array(x)
    sm.initialize(device=NumpyDevice())
    sm.run()
    ev = EvaluatorSoftmax(wf, softmax_unit=sm)
    labels = np.array([0, 3, 1, 2, -1, -1], dtype=np.int32)
    ev.output, ev.labels, ev.batch_size, ev.max_idx = sm.output, Array(labels), 4, sm.max_idx
    ev.initialize(device=NumpyDevice())
    ev.run()

    def ce():
        sm.run()
        p = sm.output.mem
        return float(-np.log(p[np.arange(4), labels[:4]]).sum() / 4)

    # d CE / d bias_j = sum over valid rows of err_output[:, j]
    num = []
    for j in range(4):
        saved = sm.bias.mem[j]
        sm.bias.mem[j] = saved + 1e-3
        up = ce()
        sm.bias.mem[j] = saved - 1e-3
        dn = ce()
        sm.bias.mem[j] = saved
        num.append((up - dn) / 2e-3)
    assert np.allclose(ev.err_output.mem.sum(axis=0), num, atol=1e-3)
    assert (ev.err_output.mem[4:] == 0).all()
    assert ev.n_err.mem[1] == 4


@pytest.mark.parametrize("solver", ["momentum", "adagrad", "adadelta"])
def test_gd_update_rule(solver):
    rng = np.random.RandomState(2)
    x = rng.uniform(-1, 1, size=(8, 6))
    wf, fwd, gd, dev = _pair("all2all", x, {"output_sample_shape": 3})
    gd.apply_gradient = True
    gd.learning_rate, gd.weights_decay, gd.l1_vs_l2, gd.gradient_moment = 0.1, 0.01, 0.25, 0.9
    gd.learning_rate_bias = 0.2
    gd.solvers = {solver}
    gd.err_output = Array(rng.uniform(-1, 1, size=(8, 3)).astype(np.float32))
    gd.initialize(device=dev)
    w0 = fwd.weights.mem.copy()
    gd.run()
    g = gd.err_output.mem.T.dot(x) + 0.01 * (0.75 * w0 + 0.125 * np.sign(w0))
    if solver == "momentum":
        assert np.allclose(fwd.weights.mem, w0 - 0.1 * g, atol=1e-5)
        w1 = fwd.weights.mem.copy()
        gd.run()
        g2 = gd.err_output.mem.T.dot(x) + 0.01 * (0.75 * w1 + 0.125 * np.sign(w1))
        assert np.allclose(fwd.weights.mem, w1 + 0.9 * (-0.1 * g) - 0.1 * g2, atol=1e-5)
    elif solver == "adagrad":
        assert np.allclose(fwd.weights.mem, w0 - 0.1 * g / np.sqrt(g * g + 1e-6), atol=1e-5)
    else:
        d = -np.sqrt(1e-6) / np.sqrt(0.1 * g * g + 1e-6) * g
        assert np.allclose(fwd.weights.mem, w0 + d, atol=1e-5)


def test_dropout_train_vs_eval():
    from veles_b200.znicz.dropout import DropoutForward, DropoutBackward
    from veles_b200.loader.base import TRAIN, VALID
    wf = DummyWorkflow()
    x = np.ones((50, 40), dtype=np.float32)
    f = DropoutForward(wf, dropout_ratio=0.4)
    f.input = Array(x)
    f.minibatch_class = TRAIN
    f.initialize(device=NumpyDevice())
    f.run()
    kept = f.output.mem != 0
    assert 0.5 < kept.mean() < 0.7
    assert np.allclose(f.output.mem[kept], 1 / 0.6)
    b = DropoutBackward(wf, forward=f)
    b.input, b.output = f.input, f.output
    b.err_output = Array(np.full_like(x, 2.0))
    b.initialize(device=NumpyDevice())
    b.run()
    assert np.allclose(b.err_input.mem, 2.0 * f.mask.mem)
    f.minibatch_class = VALID
    f.run()
    assert np.array_equal(f.output.mem, x)
