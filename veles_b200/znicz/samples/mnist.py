"""MNIST-shape fully connected sample (784-100-10, All2AllTanh + All2AllSoftmax)."""
from veles_b200.config import root
from veles_b200.znicz.standard_workflow import StandardWorkflow
import veles_b200.loader.synthetic  # noqa: F401


class MnistWorkflow(StandardWorkflow):
    """Model created for digits recognition (synthetic MNIST-shape data)."""


def run(load, main):
    cfg = root.mnist
    load(MnistWorkflow, layers=cfg.layers, loader_name=cfg.loader_name, loader_config=cfg.loader,
         loss_function=cfg.loss_function, decision_config=cfg.decision, snapshotter_config=cfg.snapshotter)
    main()
