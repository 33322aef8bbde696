"""Shared machinery of the ensemble train / test workflows.

Parity target: ``veles/ensemble/base_workflow.py:59-176``: a manager unit runs the model N times as
``python -m veles`` subprocesses (``:135-152``), collects each run's ``--result-file`` JSON, is
distributable (one model per slave job, ``:104-122``) and re-queues jobs of dropped slaves (``:124-127``).
B200 addition: models are evaluated ``parallel`` at a time, one per GPU (CUDA_VISIBLE_DEVICES).
"""
from __future__ import annotations

import json
import os
import re
import subprocess
import sys
import tempfile

from ..launcher import filter_argv
from ..mutable import Bool
from ..plumbing import Repeater
from ..units import Unit
from ..workflow import Workflow, NoMoreJobs


class EnsembleManagerBase(Unit):
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "TRAINER")
        super(EnsembleManagerBase, self).__init__(workflow, **kwargs)
        self.model_path = kwargs["model"]
        self.config_file = kwargs.get("config_file") or ""
        self.argv = list(kwargs.get("argv", []))
        self.result_file = kwargs.get("result_file")
        self.parallel = int(kwargs.get("parallel", 0))
        self.extra_env = dict(kwargs.get("env", {}))
        self.complete = Bool(False)
        self.results = []
        self._pending = {}

    @property
    def size(self):
        raise NotImplementedError

    def initialize(self, **kwargs):
        if len(self.results) != self.size:
            self.results = [None] * self.size
        if not self.parallel:
            try:
                import torch
                self.parallel = max(1, torch.cuda.device_count())
            except Exception:  # noqa: BLE001
                self.parallel = 1

    def _base_argv(self):
        argv = filter_argv(self.argv, "--ensemble-train=", "--ensemble-test=", "--result-file=", "-b", "--background",
                           "-l=", "--listen-address=", "-m=", "--master-address=", "-n=", "--nodes=", "-w=",
                           "--snapshot=", "--train-ratio=", "-t", "--test")
        keep = []
        with_value = {"-v", "--verbosity", "-a", "--backend", "-d", "--device", "-r", "--random-seed", "-p",
                      "--matplotlib-backend", "--debug", "-f", "--log-file", "--job-timeout"}
        it = iter(argv)
        for a in it:
            if a in with_value:
                keep += [a, next(it, "")]
            elif a.startswith("-") and a != "-":
                keep.append(a)
        return keep

    def _overrides(self):
        """The ``root.x.y=value`` positionals of this run's command line: every member trains under them too."""
        return [a for a in self.argv if re.match(r"^root\.[\w.]+=", a)]

    def command(self, index, result_path):
        raise NotImplementedError

    def launch(self, index, slot=0):
        tmp = tempfile.mkdtemp(prefix="veles_b200_ens_")
        result_path = os.path.join(tmp, "result.json")
        env = dict(os.environ)
        # an evaluation is an independent single-process training run: it must not join the rank group this process
        # may itself belong to (the CLI turns data parallel when it sees the torch.distributed variables)
        for key in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT",
                    "VELES_HEARTBEAT_FILE", "VELES_RANK_DEATH_STEP", "VELES_RANK_DEATH_PROBABILITY"):
            env.pop(key, None)
        env.update(self.extra_env)
        env["PYTHONPATH"] = os.pathsep.join(filter(None, [os.path.dirname(os.path.dirname(os.path.dirname(
            os.path.abspath(__file__)))), env.get("PYTHONPATH", "")]))
        if self.parallel > 1:
            env["CUDA_VISIBLE_DEVICES"] = str(slot)
        cmd = self.command(index, result_path)
        self.debug("Model %d: %s", index, " ".join(cmd))
        return subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE), result_path

    def collect(self, index, proc, result_path):
        _, err = proc.communicate()
        if proc.returncode != 0 or not os.path.exists(result_path):
            self.warning("Model %d failed (exit %s): %s", index, proc.returncode,
                         (err or b"").decode(errors="replace")[-500:])
            self.results[index] = {"error": proc.returncode}
            return
        with open(result_path) as f:
            self.results[index] = json.load(f)

    def run(self):
        if self.is_slave:
            return
        todo = [i for i, r in enumerate(self.results) if r is None]
        width = max(1, self.parallel)
        for k in range(0, len(todo), width):
            batch = [(i,) + self.launch(i, slot) for slot, i in enumerate(todo[k:k + width])]
            for i, proc, path in batch:
                self.collect(i, proc, path)
        self.complete <<= True

    # job-level distribution: one model per job
    def generate_data_for_slave(self, slave):
        for i, r in enumerate(self.results):
            if r is None and i not in self._pending.values():
                self._pending[getattr(slave, "id", slave)] = i
                return {"index": i}
        if not self._pending:
            self.complete <<= True
            raise NoMoreJobs()
        self.has_data_for_slave = False
        return None

    def apply_data_from_master(self, data):
        self.job_ = data

    def generate_data_for_master(self):
        job = getattr(self, "job_", None)
        if not job:
            return None
        proc, path = self.launch(job["index"])
        self.collect(job["index"], proc, path)
        return {"index": job["index"], "result": self.results[job["index"]]}

    def apply_data_from_slave(self, data, slave):
        if data:
            self.results[data["index"]] = data["result"]
            self._pending.pop(getattr(slave, "id", slave), None)
            self.has_data_for_slave = True

    def drop_slave(self, slave):
        self._pending.pop(getattr(slave, "id", slave), None)
        self.has_data_for_slave = True

    def get_metric_names(self):
        return {"models"}

    def get_metric_values(self):
        return {"models": self.results}


class EnsembleWorkflowBase(Workflow):
    hide_from_registry = True
    MANAGER = None

    def __init__(self, workflow, **kwargs):
        manager_keys = ("model", "config_file", "argv", "size", "train_ratio", "input_file", "parallel", "env")
        super(EnsembleWorkflowBase, self).__init__(
            workflow, **{k: v for k, v in kwargs.items() if k not in manager_keys})
        self.repeater = Repeater(self)
        self.repeater.link_from(self.start_point)
        self.ensemble = type(self).MANAGER(self, **kwargs)
        self.ensemble.link_from(self.repeater)
        self.repeater.link_from(self.ensemble)
        self.repeater.gate_block = self.ensemble.complete
        self.end_point.link_from(self.ensemble)
        self.end_point.gate_block = ~self.ensemble.complete
        if kwargs.get("result_file"):
            self.result_file = kwargs["result_file"]

    @property
    def computing_power(self):
        return 1
