"""Genetic optimisation of any model's ``Range`` hyper-parameters.

Parity target: ``veles/genetics/optimization_workflow.py:70-339``: ``GeneticsOptimizer`` evaluates a
chromosome by writing the config with gene values substituted to a pickle and running
``python -m veles --result-file ... <workflow> <config.pickle>`` as a subprocess, then reads
``EvaluationFitness`` from the result JSON (:223-286); the unit is distributable — a master hands
(chromosome index, genes) to slaves (:186-216); the best configuration is written to
``<model>_best_config.py`` (:157-170).  ``OptimizationWorkflow``: start -> Repeater -> optimizer ->
(Repeater | EndPoint gated by ``optimizer.complete``) (:298-334).

B200 addition: ``parallel_evaluations`` evaluates several chromosomes at once, one subprocess per
GPU (``CUDA_VISIBLE_DEVICES=i``) — "one model per B200".
"""
from __future__ import annotations

import json
import os
import pickle
import subprocess
import sys
import tempfile
import threading

from ..config import root, Config
from ..launcher import filter_argv
from ..mutable import Bool
from ..plumbing import Repeater
from ..units import Unit
from ..workflow import Workflow, NoMoreJobs
from .config import Tuneable, Range, collect_tuneables, fix_config
from .core import Chromosome, Population


class ConfigChromosome(Chromosome):
    def __init__(self, *args, **kwargs):
        self.optimizer_ = kwargs.pop("optimizer", None)
        super(ConfigChromosome, self).__init__(*args, **kwargs)
        self.result = None

    def init_unpickled(self):
        super(ConfigChromosome, self).init_unpickled()
        self.__dict__.setdefault("optimizer_", None)

    def evaluate(self):
        self.optimizer_.evaluate(self)


class GeneticsOptimizer(Unit):
    """Drives a Population whose fitness = EvaluationFitness of a full model training run."""
    hide_from_registry = False

    def __init__(self, workflow, **kwargs):
        kwargs.setdefault("view_group", "TRAINER")
        super(GeneticsOptimizer, self).__init__(workflow, **kwargs)
        self.model_path = kwargs["model"]            # path to the workflow file
        self.config_file = kwargs.get("config_file")
        self.size = kwargs.get("size", 10)
        self.generations = kwargs.get("generations")
        self.argv = list(kwargs.get("argv", []))
        self.parallel_evaluations = int(kwargs.get("parallel_evaluations", 0))
        self.extra_env = dict(kwargs.get("env", {}))
        self.complete = Bool(False)
        self.population = None
        self.tuneables = []
        self._chromosome_index = 0
        self.best_config_path = None
        self.pending_ = {}

    def init_unpickled(self):
        super(GeneticsOptimizer, self).init_unpickled()
        self.pending_ = {}
        self.eval_lock_ = threading.Lock()

    def initialize(self, **kwargs):
        self.tuneables = collect_tuneables(root)
        if not self.tuneables:
            raise ValueError("There are no tunable parameters (genetics.Range) in the configuration")
        for t in self.tuneables:
            self.info("Tuning %s %s", t.full_name, t.details())
        if self.population is None:
            self.population = Population(
                lambda **kw: ConfigChromosome(optimizer=self, **kw), len(self.tuneables),
                [t.min_value for t in self.tuneables], [t.max_value for t in self.tuneables], self.size,
                max_generations=self.generations)
            # seed the population with the defaults so the search never does worse than the hand-made config
            self.population.chromosomes[0].numeric = [float(t.default) for t in self.tuneables]
        for c in self.population.chromosomes:
            c.optimizer_ = self
        if not self.parallel_evaluations:
            try:
                import torch
                self.parallel_evaluations = max(1, torch.cuda.device_count())
            except Exception:  # noqa: BLE001
                self.parallel_evaluations = 1

    # -- evaluation of one chromosome ----------------------------------------------------------------------------------
    def _config_for(self, genes):
        """Config dict with gene values written into a deep copy of the non-common config subtrees."""
        with self.eval_lock_:
            saved = [(t.addr[0], t.addr[1], t) for t in self.tuneables]
            try:
                for t, g in zip(self.tuneables, genes):
                    t.set(g)
                cfg = {name: node.to_dict() for name, node in root.__content__.items()
                       if name != "common" and isinstance(node, Config)}
                cfg = pickle.loads(pickle.dumps(cfg))
            finally:
                for container, key, t in saved:
                    container[key] = t
        return cfg

    def _command(self, cfg_path, result_path):
        argv = filter_argv(self.argv, "--optimize=", "--result-file=", "-b", "--background", "-w=", "--snapshot=",
                           "-l=", "--listen-address=", "-m=", "--master-address=", "-n=", "--nodes=")
        # strip the positional workflow/config/overrides: they are re-added below
        keep = []
        skip_values = {"-v", "--verbosity", "-a", "--backend", "-d", "--device", "-r", "--random-seed", "-p",
                       "--matplotlib-backend", "--debug", "-f", "--log-file", "--train-ratio", "--job-timeout"}
        it = iter(argv)
        for a in it:
            if a in skip_values:
                keep += [a, next(it, "")]
            elif a.startswith("-") and a != "-":
                keep.append(a)
        return [sys.executable, "-m", "veles_b200"] + keep + ["-s", "--no-logo", "--result-file", result_path,
                                                              self.model_path, cfg_path]

    def launch(self, chromo, slot=0):
        tmpdir = tempfile.mkdtemp(prefix="veles_b200_gen_")
        cfg_path = os.path.join(tmpdir, "config.pickle")
        result_path = os.path.join(tmpdir, "result.json")
        with open(cfg_path, "wb") as f:
            pickle.dump(self._config_for(chromo.numeric), f)
        env = dict(os.environ)
        # an evaluation is an independent single-process training run: it must not join the rank group this process
        # may itself belong to (the CLI turns data parallel when it sees the torch.distributed variables)
        for key in ("RANK", "LOCAL_RANK", "WORLD_SIZE", "LOCAL_WORLD_SIZE", "MASTER_ADDR", "MASTER_PORT",
                    "VELES_HEARTBEAT_FILE", "VELES_RANK_DEATH_STEP", "VELES_RANK_DEATH_PROBABILITY"):
            env.pop(key, None)
        env.update(self.extra_env)
        env["PYTHONPATH"] = os.pathsep.join(filter(None, [os.path.dirname(os.path.dirname(os.path.dirname(
            os.path.abspath(__file__)))), env.get("PYTHONPATH", "")]))
        if self.parallel_evaluations > 1:
            env["CUDA_VISIBLE_DEVICES"] = str(slot)
        cmd = self._command(cfg_path, result_path)
        self.debug("Evaluating %s: %s", chromo, " ".join(cmd))
        proc = subprocess.Popen(cmd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
        return proc, result_path

    def collect(self, chromo, proc, result_path):
        _, err = proc.communicate()
        if proc.returncode != 0 or not os.path.exists(result_path):
            self.warning("Evaluation failed (exit %s): %s", proc.returncode, (err or b"").decode(errors="replace")[-500:])
            chromo.fitness = -1e30
            return
        with open(result_path) as f:
            chromo.result = json.load(f)
        chromo.fitness = float(chromo.result["EvaluationFitness"])

    def evaluate(self, chromo):
        proc, path = self.launch(chromo)
        self.collect(chromo, proc, path)

    def evaluate_generation(self):
        todo = [c for c in self.population if c.fitness is None]
        width = max(1, self.parallel_evaluations)
        for i in range(0, len(todo), width):
            batch = [(c,) + self.launch(c, slot) for slot, c in enumerate(todo[i:i + width])]
            for c, proc, path in batch:
                self.collect(c, proc, path)

    def run(self):
        if self.is_slave:
            return
        self.evaluate_generation()
        done = self.population.update()
        self.write_best_config()
        if done:
            self.complete <<= True

    def write_best_config(self):
        best = self.population.history[-1]["genes"]
        base = os.path.splitext(os.path.basename(self.model_path))[0]
        path = os.path.join(os.path.dirname(os.path.abspath(self.model_path)), base + "_best_config.py")
        lines = ["# generated by veles_b200.genetics: best configuration so far (fitness %.6g)"
                 % self.population.history[-1]["best"], "from veles_b200.config import root", ""]
        for t, g in zip(self.tuneables, best):
            cast = type(t.default)
            val = cast(round(g)) if cast is int else cast(g)
            lines.append("# %s = %r" % (t.full_name, val))
        cfg = self._config_for(best)
        for name, tree in cfg.items():
            lines.append("root.%s.update(%r)" % (name, tree))
        try:
            with open(path, "w") as f:
                f.write("\n".join(lines) + "\n")
            self.best_config_path = path
        except OSError:
            pass

    # -- job-level master/slave: one chromosome evaluation per slave job -----------------------------------------------------
    def generate_data_for_slave(self, slave):
        for i, c in enumerate(self.population):
            if c.fitness is None and i not in self.pending_.values():
                self.pending_[getattr(slave, "id", slave)] = i
                return {"index": i, "genes": list(c.numeric)}
        if self.pending_:
            self.has_data_for_slave = False
            return None
        if self.population.update():
            self.complete <<= True
            raise NoMoreJobs()
        self.write_best_config()
        return self.generate_data_for_slave(slave)

    def apply_data_from_master(self, data):
        self.job_ = data

    def generate_data_for_master(self):
        job = getattr(self, "job_", None)
        if not job:
            return None
        c = ConfigChromosome(optimizer=self, size=len(job["genes"]), min_values=[t.min_value for t in self.tuneables],
                             max_values=[t.max_value for t in self.tuneables], numeric=job["genes"])
        self.evaluate(c)
        return {"index": job["index"], "fitness": c.fitness, "result": c.result}

    def apply_data_from_slave(self, data, slave):
        if not data:
            return
        c = self.population[data["index"]]
        c.fitness, c.result = data["fitness"], data.get("result")
        self.pending_.pop(getattr(slave, "id", slave), None)
        self.has_data_for_slave = True

    def drop_slave(self, slave):
        self.pending_.pop(getattr(slave, "id", slave), None)
        self.has_data_for_slave = True

    def get_metric_names(self):
        return {"EvaluationFitness", "BestGenes", "Generations"}

    def get_metric_values(self):
        hist = self.population.history if self.population else []
        return {"EvaluationFitness": hist[-1]["best"] if hist else None,
                "BestGenes": dict(zip([t.full_name for t in self.tuneables], hist[-1]["genes"])) if hist else {},
                "Generations": len(hist)}


class OptimizationWorkflow(Workflow):
    """start -> Repeater -> GeneticsOptimizer -> (Repeater | EndPoint)."""
    hide_from_registry = True

    def __init__(self, workflow, **kwargs):
        optimizer_keys = ("model", "config_file", "argv", "size", "generations", "parallel", "env")
        super(OptimizationWorkflow, self).__init__(
            workflow, **{k: v for k, v in kwargs.items() if k not in optimizer_keys})
        self.repeater = Repeater(self)
        self.repeater.link_from(self.start_point)
        self.optimizer = GeneticsOptimizer(self, **kwargs)
        self.optimizer.link_from(self.repeater)
        self.repeater.link_from(self.optimizer)
        self.repeater.gate_block = self.optimizer.complete
        self.end_point.link_from(self.optimizer)
        self.end_point.gate_block = ~self.optimizer.complete

    @property
    def computing_power(self):
        return 1


def run(load, main, **kwargs):
    model = kwargs.pop("model")
    kwargs["model"] = getattr(model, "__file__", model)
    load(OptimizationWorkflow, **kwargs)
    main()
