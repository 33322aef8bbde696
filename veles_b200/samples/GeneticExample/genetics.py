"""Smallest possible use of the genetic optimizer (reference sample: ``veles/samples/GeneticExample/genetics.py``).

The workflow is ONE unit that evaluates a fitness from two configuration values; run it with
``python -m veles_b200 --optimize 20:10 veles_b200/samples/GeneticExample/genetics.py -`` and the optimizer
(``veles_b200/genetics``) evolves ``root.test.x`` / ``root.test.y`` inside the ranges of ``genetics_config.py``.
The objective here is Himmelblau's function (four equal minima at f = 0, e.g. x = 3, y = 2), negated because the
optimizer maximises ``EvaluationFitness``."""
from veles_b200.config import root
from veles_b200.result_provider import IResultProvider
from veles_b200.units import Unit
from veles_b200.workflow import Workflow


class Fitness(Unit, IResultProvider):
    """Publishes -himmelblau(x, y) as the run's fitness."""

    def __init__(self, workflow, **kwargs):
        super(Fitness, self).__init__(workflow, **kwargs)
        self.fitness = 0.0

    def initialize(self, **kwargs):
        pass

    def run(self):
        x, y = float(root.test.x), float(root.test.y)
        self.fitness = -((x * x + y - 11.0) ** 2 + (x + y * y - 7.0) ** 2)

    def get_metric_names(self):
        return {"EvaluationFitness"}

    def get_metric_values(self):
        return {"EvaluationFitness": self.fitness}


class FitnessWorkflow(Workflow):
    """start -> fitness -> end: one evaluation per run."""

    def __init__(self, workflow, **kwargs):
        super(FitnessWorkflow, self).__init__(workflow, **kwargs)
        self.fitness = Fitness(self)
        self.fitness.link_from(self.start_point)
        self.end_point.link_from(self.fitness)


def run(load, main):
    load(FitnessWorkflow)
    main()
