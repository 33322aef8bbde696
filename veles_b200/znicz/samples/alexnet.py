"""AlexNet on ImageNet-shape (227x227x3) synthetic data."""
from veles_b200.config import root
from veles_b200.znicz.standard_workflow import StandardWorkflow
import veles_b200.loader.synthetic  # noqa: F401


class AlexNetWorkflow(StandardWorkflow):
    """AlexNet (96-256-384-384-256 conv, 4096-4096-1000 fc)."""


def run(load, main):
    cfg = root.alexnet
    load(AlexNetWorkflow, layers=cfg.layers, loader_name=cfg.loader_name, loader_config=cfg.loader,
         loss_function=cfg.loss_function, decision_config=cfg.decision, snapshotter_config=cfg.snapshotter)
    main()
