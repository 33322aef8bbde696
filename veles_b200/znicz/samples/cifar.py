"""CIFAR-10-shape convolutional sample (Caffe cifar10_quick topology, Conv/MaxPooling/GDConv)."""
from veles_b200.config import root
from veles_b200.znicz.standard_workflow import StandardWorkflow
import veles_b200.loader.synthetic  # noqa: F401


class CifarWorkflow(StandardWorkflow):
    """Convolutional network on 32x32x3 images."""


def run(load, main):
    cfg = root.cifar
    load(CifarWorkflow, layers=cfg.layers, loader_name=cfg.loader_name, loader_config=cfg.loader,
         loss_function=cfg.loss_function, decision_config=cfg.decision, snapshotter_config=cfg.snapshotter)
    main()
