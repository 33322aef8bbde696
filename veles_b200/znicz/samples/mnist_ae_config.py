from veles_b200.config import root
from veles_b200.models.configs import wide_mlp_ae_layers

root.mnist_ae.update({
    "loss_function": "mse",
    "loader_name": "synthetic_mse",
    "loader": {"minibatch_size": 100, "n_train": 2000, "n_valid": 400, "sample_shape": (64,)},
    "decision": {"max_epochs": 5, "fail_iterations": 20},
    "snapshotter": {"prefix": "mnist_ae", "time_interval": 0},
    "layers": wide_mlp_ae_layers(width=32, depth=3, in_features=64, lr=0.02),
})
