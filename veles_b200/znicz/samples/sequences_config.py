from veles_b200.config import root
from veles_b200.genetics import Range

root.sequences.update({
    "loss_function": "softmax",
    "loader_name": "synthetic_fullbatch",
    "loader": {"minibatch_size": 64, "n_train": 1280, "n_valid": 256, "sample_shape": (12, 16), "n_classes": 10,
               "noise": 0.5},
    "decision": {"max_epochs": 8, "fail_iterations": 20},
    "snapshotter": {"prefix": "sequences", "time_interval": 0, "compression": "gz"},
    "layers": [{"type": "lstm", "name": "lstm1", "->": {"hidden_size": Range(32, 8, 128), "return_sequences": False},
                "<-": {"learning_rate": Range(0.05, 0.001, 0.3), "weights_decay": 0.0, "gradient_moment": 0.9}},
               {"type": "softmax", "name": "fc_softmax2", "->": {"output_sample_shape": 10},
                "<-": {"learning_rate": Range(0.05, 0.001, 0.3), "weights_decay": 0.0, "gradient_moment": 0.9}}],
})
