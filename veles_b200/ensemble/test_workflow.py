"""Ensemble evaluation: replay each member's snapshot in ``--test`` mode
(reference: ``veles/ensemble/test_workflow.py:50-109``).  Input = the result file written by
``--ensemble-train`` (list of per-model results with ``Snapshot`` paths)."""
from __future__ import annotations

import json
import sys

from .base_workflow import EnsembleManagerBase, EnsembleWorkflowBase


class EnsembleTestManager(EnsembleManagerBase):
    """Replays the snapshot of every ensemble member in evaluation mode and collects the outputs."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        super(EnsembleTestManager, self).__init__(workflow, **kwargs)
        self.input_file = kwargs["input_file"]
        with open(self.input_file) as f:
            data = json.load(f)
        self.models = [m for m in data.get("models", data if isinstance(data, list) else []) if m and "Snapshot" in m]

    @property
    def size(self):
        return len(self.models)

    def command(self, index, result_path):
        return [sys.executable, "-m", "veles_b200"] + self._base_argv() + [
            "-s", "--no-logo", "--test", "--result-file", result_path, "-w", self.models[index]["Snapshot"],
            self.model_path, self.config_file or ""] + self._overrides()


class EnsembleTestWorkflow(EnsembleWorkflowBase):
    """start -> Repeater -> EnsembleTestManager -> end: the workflow behind `--ensemble-test`."""
    MANAGER = EnsembleTestManager


def run(load, main, **kwargs):
    model = kwargs.pop("model")
    kwargs["model"] = getattr(model, "__file__", model)
    load(EnsembleTestWorkflow, **kwargs)
    main()
