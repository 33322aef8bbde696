from veles_b200.config import root
from veles_b200.genetics import Range
from veles_b200.models.configs import cifar_convnet_layers

root.cifar.update({
    "loss_function": "softmax",
    "loader_name": "cifar_loader",
    "loader": {"minibatch_size": 100, "n_train": 2000, "n_valid": 500, "noise": 1.0,
               "normalization_type": "mean_disp"},
    "decision": {"max_epochs": 4, "fail_iterations": 20},
    "snapshotter": {"prefix": "cifar", "time_interval": 0},
    "layers": cifar_convnet_layers(lr=Range(0.01, 0.0005, 0.05), caffe_init=False),
})
