from veles_b200.config import root
from veles_b200.models.configs import alexnet_layers

root.alexnet.update({
    "loss_function": "softmax",
    "loader_name": "imagenet_loader",
    "loader": {"minibatch_size": 256, "n_train": 2048, "n_valid": 256, "normalization_type": "mean_disp"},
    "decision": {"max_epochs": 2, "fail_iterations": 20},
    "snapshotter": {"prefix": "alexnet", "time_interval": 0},
    "layers": alexnet_layers(lr=0.01),
})
