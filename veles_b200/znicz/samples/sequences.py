"""Sequence classification sample: LSTM over [T, D] samples, last state -> softmax (recurrent layers with BPTT;
the reference lists RNN / LSTM workflows as "created, not tested": docs/source/manualrst_veles_algorithms.rst:108-135)."""
from veles_b200.config import root
from veles_b200.znicz.standard_workflow import StandardWorkflow
import veles_b200.loader.synthetic  # noqa: F401


class SequencesWorkflow(StandardWorkflow):
    """Classifies synthetic sequences (class-dependent mean trajectory + noise)."""


def run(load, main):
    cfg = root.sequences
    load(SequencesWorkflow, layers=cfg.layers, loader_name=cfg.loader_name, loader_config=cfg.loader,
         loss_function=cfg.loss_function, decision_config=cfg.decision, snapshotter_config=cfg.snapshotter)
    main()
