"""Layer lists (StandardWorkflow ``layers`` format) and factories for the benchmark models."""
from __future__ import annotations


def _gd(lr, wd=0.0005, moment=0.9, lr_bias=None, wd_bias=0.0):
    return {"learning_rate": lr, "learning_rate_bias": lr_bias if lr_bias is not None else 2 * lr,
            "weights_decay": wd, "weights_decay_bias": wd_bias, "gradient_moment": moment,
            "gradient_moment_bias": moment}


def mnist_mlp_layers(lr=0.03):
    return [{"type": "all2all_tanh", "name": "fc_tanh1", "->": {"output_sample_shape": 100}, "<-": _gd(lr, 0.0)},
            {"type": "softmax", "name": "fc_softmax2", "->": {"output_sample_shape": 10}, "<-": _gd(lr, 0.0)}]


def cifar_convnet_layers(lr=0.001, n_classes=10, caffe_init=True):
    """Caffe "cifar10_quick"-style net as shipped in Znicz samples (conv 32-32-64, 5x5 pad 2).
    ``caffe_init=False`` uses the framework's fan-based default initialisation (trains faster)."""
    def conv(n, std):
        fwd = {"n_kernels": n, "kx": 5, "ky": 5, "padding": (2, 2, 2, 2), "sliding": (1, 1)}
        if caffe_init:
            fwd.update({"weights_filling": "gaussian", "weights_stddev": std, "bias_filling": "constant",
                        "bias_stddev": 0})
        return {"type": "conv_str", "->": fwd, "<-": _gd(lr)}

    def fc(tpe, n):
        fwd = {"output_sample_shape": n}
        if caffe_init:
            fwd.update({"weights_filling": "gaussian", "weights_stddev": 0.05})
        return {"type": tpe, "->": fwd, "<-": _gd(lr)}
    return [conv(32, 0.01),
            {"type": "max_pooling", "->": {"kx": 3, "ky": 3, "sliding": (2, 2)}},
            conv(32, 0.01),
            {"type": "avg_pooling", "->": {"kx": 3, "ky": 3, "sliding": (2, 2)}},
            conv(64, 0.01),
            {"type": "avg_pooling", "->": {"kx": 3, "ky": 3, "sliding": (2, 2)}},
            fc("all2all_str", 64), fc("softmax", n_classes)]


def alexnet_layers(lr=0.01, n_classes=1000, dropout=0.5):
    """AlexNet (Caffe reference layout: 96-256-384-384-256 kernels, groups 2 on conv2/4/5, LRN after
    conv1/conv2, 3x3/2 max pooling, fc 4096-4096-1000 with dropout)."""
    def conv(n, k, pad, stride=1, group=1, std=0.01, bias=0.0):
        return {"type": "conv_str", "->": {"n_kernels": n, "kx": k, "ky": k, "padding": (pad,) * 4,
                                           "sliding": (stride, stride), "grouping": group,
                                           "weights_filling": "gaussian", "weights_stddev": std,
                                           "bias_filling": "constant", "bias_stddev": bias}, "<-": _gd(lr)}
    pool = {"type": "max_pooling", "->": {"kx": 3, "ky": 3, "sliding": (2, 2)}}
    norm = {"type": "norm", "->": {"alpha": 0.0001, "beta": 0.75, "n": 5, "k": 1}}

    def fc(n, std=0.005, bias=0.1):
        return {"type": "all2all_str", "->": {"output_sample_shape": n, "weights_filling": "gaussian",
                                              "weights_stddev": std, "bias_filling": "constant",
                                              "bias_stddev": bias}, "<-": _gd(lr)}
    return [conv(96, 11, 0, 4), dict(norm), dict(pool),
            conv(256, 5, 2, 1, 2, bias=0.1), dict(norm), dict(pool),
            conv(384, 3, 1), conv(384, 3, 1, 1, 2, bias=0.1), conv(256, 3, 1, 1, 2, bias=0.1), dict(pool),
            fc(4096), {"type": "dropout", "->": {"dropout_ratio": dropout}},
            fc(4096), {"type": "dropout", "->": {"dropout_ratio": dropout}},
            {"type": "softmax", "->": {"output_sample_shape": n_classes, "weights_filling": "gaussian",
                                       "weights_stddev": 0.01, "bias_filling": "constant", "bias_stddev": 0},
             "<-": _gd(lr)}]


def wide_mlp_ae_layers(width=4096, depth=4, in_features=4096, lr=0.001):
    layers = [{"type": "all2all_tanh", "->": {"output_sample_shape": width}, "<-": _gd(lr, 0.0)}
              for _ in range(depth - 1)]
    layers.append({"type": "all2all", "->": {"output_sample_shape": in_features}, "<-": _gd(lr, 0.0)})
    return layers


def lstm_classifier_layers(hidden=128, n_classes=10, lr=0.02, stacked=False):
    """Sequence classifier: LSTM over [T, D] samples (optionally two stacked layers), last state -> softmax."""
    layers = []
    if stacked:
        layers.append({"type": "lstm", "->": {"hidden_size": hidden}, "<-": _gd(lr, 0.0)})
    layers.append({"type": "lstm", "->": {"hidden_size": hidden, "return_sequences": False}, "<-": _gd(lr, 0.0)})
    layers.append({"type": "softmax", "->": {"output_sample_shape": n_classes}, "<-": _gd(lr, 0.0)})
    return layers


def mnist_lstm_layers(hidden=64, lr=0.02):
    """MNIST-shape images read row by row: a 28-step LSTM over 28 pixels, last state -> softmax."""
    return [{"type": "lstm", "->": {"hidden_size": hidden, "return_sequences": False, "sequence_length": 28},
             "<-": _gd(lr, 0.0)},
            {"type": "softmax", "->": {"output_sample_shape": 10}, "<-": _gd(lr, 0.0)}]


MODEL_CONFIGS = {
    "mnist_lstm": dict(layers=mnist_lstm_layers, loader_name="mnist_loader", loss_function="softmax",
                       sample_shape=(784,), n_classes=10, minibatch=100),
    "lstm_classifier": dict(layers=lstm_classifier_layers, loader_name="synthetic_fullbatch", loss_function="softmax",
                            sample_shape=(16, 64), n_classes=10, minibatch=128),
    "mnist_mlp": dict(layers=mnist_mlp_layers, loader_name="mnist_loader", loss_function="softmax",
                      sample_shape=(784,), n_classes=10, minibatch=100),
    "cifar_convnet": dict(layers=cifar_convnet_layers, loader_name="cifar_loader", loss_function="softmax",
                          sample_shape=(32, 32, 3), n_classes=10, minibatch=256),
    "alexnet": dict(layers=alexnet_layers, loader_name="imagenet_loader", loss_function="softmax",
                    sample_shape=(227, 227, 3), n_classes=1000, minibatch=256),
    "wide_mlp_ae": dict(layers=wide_mlp_ae_layers, loader_name="synthetic_mse", loss_function="mse",
                        sample_shape=(4096,), n_classes=0, minibatch=1024),
}


def build_workflow(model, launcher=None, minibatch_size=None, n_train=None, n_valid=0, max_epochs=None,
                   shard_rank=0, shard_world=1, loader_extra=None, layers_kwargs=None, **wf_kwargs):
    """Create (not initialize) a StandardWorkflow for one of MODEL_CONFIGS."""
    from ..dummy import DummyLauncher
    from ..znicz.standard_workflow import StandardWorkflow
    from ..loader import synthetic  # noqa: F401  (registers the loaders)
    cfg = MODEL_CONFIGS[model]
    mb = minibatch_size or cfg["minibatch"]
    loader_cfg = {"minibatch_size": mb, "n_train": n_train or 4 * mb, "n_valid": n_valid,
                  "sample_shape": cfg["sample_shape"], "shard_rank": shard_rank, "shard_world": shard_world}
    if cfg["n_classes"]:
        loader_cfg["n_classes"] = cfg["n_classes"]
    loader_cfg.update(loader_extra or {})
    launcher = launcher or DummyLauncher()
    wf = StandardWorkflow(launcher, layers=cfg["layers"](**(layers_kwargs or {})), loader_name=cfg["loader_name"],
                          loader_config=loader_cfg, loss_function=cfg["loss_function"],
                          decision_config={"max_epochs": max_epochs}, **wf_kwargs)
    return wf


def smoke_step():
    """One tiny forward+backward+update of the CIFAR convnet on cuda:0 (driver smoke test)."""
    import torch
    from ..backends import B200Device
    from ..config import root
    root.common.disable.snapshotting = True
    wf = build_workflow("cifar_convnet", minibatch_size=32, n_train=64, max_epochs=1)
    wf.initialize(device=B200Device(index=0))
    wf.loader.run()
    for f in wf.forwards:
        f.run()
    wf.evaluator.run()
    for gd in reversed(wf.gds):
        gd.run()
    torch.cuda.synchronize()
    wf.forwards[-1].output.map_read()
    out = wf.forwards[-1].output.mem
    assert out.shape[0] == 32 and abs(float(out[0].sum()) - 1.0) < 1e-2, "softmax output is not normalised"
    wf.launcher.stop()
    return wf
