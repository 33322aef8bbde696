"""Every Znicz layer pair: sm_100a device path vs the numpy oracle on identical inputs."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [
    ("all2all_tanh", (64, 100), {"output_sample_shape": 50}),
    ("all2all_str", (33, 257), {"output_sample_shape": 130}),
    ("all2all_sigmoid", (128, 784), {"output_sample_shape": 100}),
    ("all2all_relu", (16, 24), {"output_sample_shape": 8}),
    ("conv_str", (4, 32, 32, 3), {"n_kernels": 32, "kx": 5, "ky": 5, "padding": (2, 2, 2, 2), "sliding": (1, 1)}),
    ("conv_tanh", (3, 15, 15, 16), {"n_kernels": 32, "kx": 3, "ky": 3, "padding": (1, 1, 1, 1), "sliding": (2, 2),
                                    "grouping": 2}),
    ("conv", (2, 27, 27, 3), {"n_kernels": 8, "kx": 11, "ky": 11, "padding": (0, 0, 0, 0), "sliding": (4, 4)}),
    ("conv_str", (3, 37, 30, 3), {"n_kernels": 16, "kx": 7, "ky": 5, "padding": (0, 0, 0, 0), "sliding": (2, 4)}),
    ("conv_str", (2, 13, 13, 64), {"n_kernels": 64, "kx": 3, "ky": 3, "padding": (1, 1, 1, 1), "sliding": (1, 1)}),
    ("conv_tanh", (2, 14, 14, 32), {"n_kernels": 48, "kx": 5, "ky": 5, "padding": (2, 2, 2, 2), "sliding": (1, 1),
                                    "grouping": 2}),
    ("conv_sigmoid", (3, 12, 10, 5), {"n_kernels": 20, "kx": 3, "ky": 4, "padding": (1, 2, 0, 1), "sliding": (1, 2)}),
    ("all2all_tanh", (7, 33), {"output_sample_shape": 19}),
    ("max_pooling", (3, 9, 9, 5), {"kx": 2, "ky": 2, "sliding": (2, 2)}),
    ("max_pooling", (4, 15, 15, 32), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("avg_pooling", (4, 16, 16, 32), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("maxabs_pooling", (4, 8, 8, 16), {"kx": 2, "ky": 2, "sliding": (2, 2)}),
    ("cutter", (3, 10, 12, 8), {"padding": (2, 1, 3, 2)}),
    ("cutter", (2, 9, 7, 5), {"padding": (1, 0, 0, 2)}),
    ("all2all_resizable", (16, 40), {"output_sample_shape": 24}),
    ("norm", (4, 6, 6, 96), {"alpha": 0.0001, "beta": 0.75, "n": 5, "k": 1}),
    ("norm", (3, 5, 7, 256), {"alpha": 0.0001, "beta": 0.75, "n": 5, "k": 2}),
    ("norm", (5, 3, 3, 40), {"alpha": 0.01, "beta": 0.5, "n": 3, "k": 1}),
    ("norm", (2, 9, 4, 8), {"alpha": 0.01, "beta": 0.75, "n": 7, "k": 1}),
    ("norm", (2, 4, 4, 384), {"alpha": 0.001, "beta": 0.75, "n": 5, "k": 2}),
    ("max_pooling", (3, 16, 14, 24), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("max_pooling", (2, 13, 13, 256), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("maxabs_pooling", (3, 11, 15, 16), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("max_pooling", (2, 5, 3, 8), {"kx": 3, "ky": 3, "sliding": (2, 2)}),
    ("activation_tanh", (8, 40), {}),
    ("activation_str", (8, 40), {}),
    ("activation_log", (8, 40), {}),
    ("rnn", (16, 6, 24), {"hidden_size": 32}),
    ("rnn", (5, 7, 10), {"hidden_size": 12, "return_sequences": False}),
    ("lstm", (16, 6, 24), {"hidden_size": 32}),
    ("lstm", (8, 5, 13), {"hidden_size": 20}),
    ("lstm", (32, 4, 64), {"hidden_size": 64, "return_sequences": False}),
    ("lstm", (4, 36), {"hidden_size": 8, "sequence_length": 6}),
    ("all2all_tanh", (64, 96), {"output_sample_shape": 48, "weights_transposed": True}),
    ("all2all_str", (33, 257), {"output_sample_shape": 130, "weights_transposed": True}),
    ("softmax", (32, 100), {"output_sample_shape": 10, "weights_transposed": True}),
]


def _run(tpe, x, kw, device, R, weights=None, prev_activation=0):
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.memory import Array
    from veles_b200.znicz.standard_workflow import LayerRegistry
    wf = DummyWorkflow()
    inp = Array(x.astype(np.float32))
    if device.exists:
        inp.dev_dtype = device.compute_dtype
        inp.initialize(device)
    fwd = LayerRegistry.forwards()[tpe](wf, **kw)
    fwd.input = inp
    assert not fwd.initialize(device=device)
    if weights is not None:
        fwd.weights.map_write()
        fwd.weights.mem[...] = weights[0]
        fwd.bias.map_write()
        fwd.bias.mem[...] = weights[1]
        fwd.weights.unmap()
        fwd.bias.unmap()
        if device.exists:
            fwd.refresh_shadow()
    fwd.run()
    gd = LayerRegistry.gds()[tpe](wf, forward=fwd, learning_rate=0.0, weights_decay=0.0, apply_gradient=False,
                                  prev_activation=prev_activation)
    gd.input, gd.output = fwd.input, fwd.output
    if getattr(fwd, "HAS_WEIGHTS", False):
        gd.weights, gd.bias = fwd.weights, fwd.bias
    eo = Array((R if R is not None else np.ones(fwd.output.shape)).astype(np.float32))
    if device.exists:
        eo.dev_dtype = device.compute_dtype
        eo.initialize(device)
    gd.err_output = eo
    assert not gd.initialize(device=device)
    gd.run()
    res = {}
    fwd.output.map_read()
    res["output"] = fwd.output.mem.copy()
    gd.err_input.map_read()
    res["err_input"] = gd.err_input.mem.copy()
    if getattr(fwd, "input_offset", None) is not None and fwd.input_offset:
        fwd.input_offset.map_read()     # lazily expanded from the 1-byte records on the device path
        res["offs"] = fwd.input_offset.mem.copy()
    if getattr(fwd, "HAS_WEIGHTS", False):
        gd.gradient_weights.map_read()
        gd.gradient_bias.map_read()
        res["gw"] = gd.gradient_weights.mem.copy()
        res["gb"] = gd.gradient_bias.mem.copy()
        fwd.weights.map_read()
        fwd.bias.map_read()
        res["w"] = (fwd.weights.mem.copy(), fwd.bias.mem.copy())
    return res


@pytest.mark.parametrize("tpe,shape,kw", CASES, ids=[c[0] + str(i) for i, c in enumerate(CASES)])
@pytest.mark.parametrize("prev_act", [0, 4])
def test_layer_device_vs_numpy(tpe, shape, kw, prev_act):
    from veles_b200.backends import NumpyDevice, B200Device
    from veles_b200 import prng
    if prev_act and (tpe.startswith("activation") or tpe in ("rnn", "lstm")):
        pytest.skip("this unit does not fuse a previous derivative")
    rng = np.random.RandomState(abs(hash(tpe)) % 1000)
    x = rng.uniform(-1, 1, size=shape)
    if "pooling" in tpe:
        x = rng.permutation(x.size).reshape(shape) / x.size * 2 - 1  # no ties
    x = x.astype(np.float32)
    import torch
    x = torch.from_numpy(x).bfloat16().float().numpy()  # both backends see bf16-exact inputs
    prng.get().seed(11)
    ref = _run(tpe, x, kw, NumpyDevice(), None, prev_activation=prev_act)
    R = rng.uniform(-1, 1, size=ref["output"].shape).astype(np.float32)
    R = torch.from_numpy(R).bfloat16().float().numpy()
    w = ref.get("w")
    if w is not None:
        w = (torch.from_numpy(w[0]).bfloat16().float().numpy(), w[1])
    ref = _run(tpe, x, kw, NumpyDevice(), R, weights=w, prev_activation=prev_act)
    dev = _run(tpe, x, kw, B200Device(), R, weights=w, prev_activation=prev_act)
    if "offs" in ref:
        assert np.array_equal(ref["offs"], dev["offs"])
    for key in ("output", "err_input", "gw", "gb"):
        if key not in ref:
            continue
        a, b = ref[key], dev[key]
        assert a.shape == b.shape, key
        scale = np.abs(a).max() + 1e-6
        err = np.abs(a - b).max() / scale
        assert err < 2.5e-2, (key, err)


def test_dropout_device_statistics_and_backward_mask():
    import torch
    from veles_b200.backends import B200Device
    from veles_b200.dummy import DummyWorkflow
    from veles_b200.memory import Array
    from veles_b200.znicz.dropout import DropoutForward, DropoutBackward
    dev = B200Device()
    wf = DummyWorkflow()
    x = np.ones((256, 4096), dtype=np.float32)
    inp = Array(x)
    inp.dev_dtype = torch.bfloat16
    inp.initialize(dev)
    f = DropoutForward(wf, dropout_ratio=0.5)
    f.input = inp
    f.initialize(device=dev)
    f.run()
    f.output.map_read()
    kept = f.output.mem != 0
    assert 0.49 < kept.mean() < 0.51
    assert np.allclose(f.output.mem[kept], 2.0)
    b = DropoutBackward(wf, forward=f)
    b.input, b.output = f.input, f.output
    eo = Array(np.full_like(x, 3.0))
    eo.dev_dtype = torch.bfloat16
    eo.initialize(dev)
    b.err_output = eo
    b.initialize(device=dev)
    b.run()
    b.err_input.map_read()
    assert np.array_equal(b.err_input.mem != 0, kept)
    assert np.allclose(b.err_input.mem[kept], 6.0)


def test_cifar_convnet_trains_on_device():
    from veles_b200 import prng
    from veles_b200.backends import B200Device
    from veles_b200.config import root
    from veles_b200.models import build_workflow
    root.common.disable.snapshotting = True
    prng.get().seed(3)
    wf = build_workflow("cifar_convnet", minibatch_size=100, n_train=2000, n_valid=500, max_epochs=4,
                        layers_kwargs={"lr": 0.01, "caffe_init": False},
                        loader_extra={"noise": 1.0, "normalization_type": "mean_disp"})
    wf.initialize(device=B200Device())
    wf.run()
    hist = wf.decision.epoch_metrics_history
    assert hist[-1]["loss"][1] < hist[0]["loss"][1] - 0.3, hist
    assert hist[-1]["n_err_pt"][1] < 60.0, hist
    wf.launcher.stop()


def test_smoke_entry_point():
    import __graft_entry__
    __graft_entry__.smoke()
