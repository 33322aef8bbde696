from veles_b200.config import root
from veles_b200.genetics import Range

root.mnist.update({
    "loss_function": "softmax",
    "loader_name": "mnist_loader",
    "loader": {"minibatch_size": Range(100, 20, 500), "n_train": 3000, "n_valid": 500, "noise": 1.5},
    "decision": {"max_epochs": 3, "fail_iterations": 20},
    "snapshotter": {"prefix": "mnist", "time_interval": 0, "compression": "gz"},
    "layers": [{"type": "all2all_tanh", "name": "fc_tanh1", "->": {"output_sample_shape": 100},
                "<-": {"learning_rate": Range(0.03, 0.001, 0.3), "weights_decay": 0.0, "gradient_moment": 0.9}},
               {"type": "softmax", "name": "fc_softmax2", "->": {"output_sample_shape": 10},
                "<-": {"learning_rate": Range(0.03, 0.001, 0.3), "weights_decay": 0.0, "gradient_moment": 0.9}}],
})
