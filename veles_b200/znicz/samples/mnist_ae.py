"""Wide MLP autoencoder (All2All stack, MSE loss) — the gradient all-reduce stress model."""
from veles_b200.config import root
from veles_b200.znicz.standard_workflow import StandardWorkflow
import veles_b200.loader.synthetic  # noqa: F401


class AutoencoderWorkflow(StandardWorkflow):
    """Autoencoder trained with EvaluatorMSE / DecisionMSE."""


def run(load, main):
    cfg = root.mnist_ae
    load(AutoencoderWorkflow, layers=cfg.layers, loader_name=cfg.loader_name, loader_config=cfg.loader,
         loss_function=cfg.loss_function, decision_config=cfg.decision, snapshotter_config=cfg.snapshotter)
    main()
