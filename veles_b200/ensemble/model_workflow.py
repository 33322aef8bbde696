"""Ensemble training: N models, each on a random ``train_ratio`` subset of the train set
(reference: ``veles/ensemble/model_workflow.py:50-152``).  Per-model result = the model's result JSON
(id, log_id, Snapshot, EvaluationFitness, ...)."""
from __future__ import annotations

import sys

from .base_workflow import EnsembleManagerBase, EnsembleWorkflowBase


class EnsembleModelManager(EnsembleManagerBase):
    """Trains the N members of an ensemble as separate `veles` runs (one per GPU at a time), each on a random
    subset of the train set, and collects their result files."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(EnsembleModelManager, self).__init__(workflow, **kwargs)
        self._size = int(kwargs.get("size", 2))
        self.train_ratio = float(kwargs.get("train_ratio", 1.0))

    @property
    def size(self):
        return self._size

    def command(self, index, result_path):
        cmd = [sys.executable, "-m", "veles_b200"] + self._base_argv() + [
            "-s", "--no-logo", "--result-file", result_path, "--train-ratio", str(self.train_ratio),
            "-r", "%08x" % (0x5eed0000 + index), self.model_path, self.config_file or ""] + self._overrides() + [
            "root.common.ensemble.model_index=%d" % index, "root.common.ensemble.size=%d" % self.size]
        return cmd


class EnsembleModelWorkflow(EnsembleWorkflowBase):
    """start -> Repeater -> EnsembleModelManager -> end: the workflow behind `--ensemble-train`."""
    MANAGER = EnsembleModelManager


def run(load, main, **kwargs):
    model = kwargs.pop("model")
    kwargs["model"] = getattr(model, "__file__", model)
    load(EnsembleModelWorkflow, **kwargs)
    main()
