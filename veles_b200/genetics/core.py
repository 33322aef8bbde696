"""Real-coded genetic algorithm.

Parity target: ``veles/genetics/core.py:58-830``: chromosomes of bounded real genes (optional
Gray-code binary representation), roulette-wheel selection (:578-594), uniform / arithmetic /
geometric crossover pipeline (:427-429, 672-785), gaussian + uniform ("altering") mutation
(:310-368, 434-450), elitist truncation (:507-512), stop on max generations or no improvement
(:801-830).  Higher fitness is better.
"""
from __future__ import annotations

import numpy

from .. import prng
from ..distributable import Pickleable


def gray_encode(n):
    return n ^ (n >> 1)


def gray_decode(g):
    n = 0
    while g:
        n ^= g
        g >>= 1
    return n


class Chromosome(Pickleable):
    """A vector of genes with per-gene bounds; ``fitness`` is None until evaluated."""

    def __init__(self, size=0, min_values=None, max_values=None, accuracy=None, codes=None, binary=None,
                 numeric=None, rand=None, **kwargs):
        super(Chromosome, self).__init__(**kwargs)
        self.rand = rand or prng.get()
        self.size = size
        self.min_values = list(min_values) if min_values is not None else []
        self.max_values = list(max_values) if max_values is not None else []
        self.accuracy = accuracy
        self.fitness = None
        if numeric is not None:
            self.numeric = [float(v) for v in numeric]
        else:
            self.numeric = [float(self.rand.uniform(lo, hi)) for lo, hi in zip(self.min_values, self.max_values)]
        self.binary_bits = kwargs.get("binary_bits", 16)

    def copy(self):
        c = type(self).__new__(type(self))
        c.__dict__.update({k: v for k, v in self.__dict__.items()})
        c.numeric = list(self.numeric)
        c.min_values = list(self.min_values)
        c.max_values = list(self.max_values)
        return c

    # -- binary (Gray) view used by the bit-flip mutation / binary crossover ------------------------------------
    def to_binary(self):
        bits = []
        top = (1 << self.binary_bits) - 1
        for v, lo, hi in zip(self.numeric, self.min_values, self.max_values):
            q = 0 if hi == lo else int(round((v - lo) / (hi - lo) * top))
            g = gray_encode(q)
            bits.extend((g >> i) & 1 for i in reversed(range(self.binary_bits)))
        return bits

    def from_binary(self, bits):
        top = (1 << self.binary_bits) - 1
        out = []
        for gi, (lo, hi) in enumerate(zip(self.min_values, self.max_values)):
            chunk = bits[gi * self.binary_bits:(gi + 1) * self.binary_bits]
            g = 0
            for b in chunk:
                g = (g << 1) | int(b)
            out.append(lo + (hi - lo) * gray_decode(g) / top)
        self.numeric = out
        self.fitness = None

    def clip(self):
        self.numeric = [min(max(v, lo), hi) for v, lo, hi in zip(self.numeric, self.min_values, self.max_values)]

    # -- mutations ------------------------------------------------------------------------------------------------
    def mutate_gaussian(self, n_points, probability, sigma_scale=0.1):
        for _ in range(n_points):
            if self.rand.random_sample() >= probability or not self.numeric:
                continue
            i = int(self.rand.randint(0, len(self.numeric)))
            span = self.max_values[i] - self.min_values[i]
            self.numeric[i] += float(self.rand.normal(0.0, sigma_scale * span if span else 1.0))
        self.clip()
        self.fitness = None

    def mutate_uniform(self, n_points, probability):
        for _ in range(n_points):
            if self.rand.random_sample() >= probability or not self.numeric:
                continue
            i = int(self.rand.randint(0, len(self.numeric)))
            self.numeric[i] = float(self.rand.uniform(self.min_values[i], self.max_values[i]))
        self.fitness = None

    def mutate_binary_point(self, n_points, probability):
        bits = self.to_binary()
        for _ in range(n_points):
            if self.rand.random_sample() < probability and bits:
                i = int(self.rand.randint(0, len(bits)))
                bits[i] ^= 1
        self.from_binary(bits)

    def evaluate(self):
        raise NotImplementedError("Chromosome.evaluate() must be provided by the application")

    def __repr__(self):
        return "<Chromosome fitness=%s genes=%s>" % (self.fitness, ["%.4g" % v for v in self.numeric])


class Population(Pickleable):
    """A set of chromosomes evolving towards higher fitness."""

    def __init__(self, chromosome_factory, optimization_size, min_values, max_values, population_size,
                 accuracy=0.00001, rand=None, max_generations=None, crossing_attempts=10, **kwargs):
        super(Population, self).__init__(**kwargs)
        self.rand = rand or prng.get()
        self.size = population_size
        self.chromosome_factory = chromosome_factory
        self.optimization_size = optimization_size
        self.min_values = list(min_values)
        self.max_values = list(max_values)
        self.accuracy = accuracy
        self.max_generations = max_generations
        self.crossing_attempts = crossing_attempts
        self.generation = 0
        self.chromosomes = []
        self.fitness = None
        self.average_fit = None
        self.best_fit = None
        self.worst_fit = None
        self.median_fit = None
        self.prev_best = None
        self.improved_generations_ago = 0
        self.no_improvement_limit = kwargs.get("no_improvement_limit", 15)
        self.roulette_select_size = kwargs.get("roulette_select_size", 0.75)
        self.crossings = kwargs.get("crossings", {"uniform": 1.0, "arithmetic": 0.9, "geometric": 0.9})
        self.mutations = kwargs.get("mutations", {"gaussian": {"points": 3, "probability": 0.35},
                                                  "uniform": {"points": 3, "probability": 0.15},
                                                  "binary_point": {"points": 1, "probability": 0.0}})
        self.elite = kwargs.get("elite", 2)
        self.history = []
        for _ in range(self.size):
            self.add(self.new(self.optimization_size))

    def new(self, size):
        return self.chromosome_factory(size=size, min_values=self.min_values, max_values=self.max_values,
                                       accuracy=self.accuracy, rand=self.rand)

    def add(self, chromo):
        self.chromosomes.append(chromo)

    def __len__(self):
        return len(self.chromosomes)

    def __getitem__(self, i):
        return self.chromosomes[i]

    def __iter__(self):
        return iter(self.chromosomes)

    def evaluate(self, callback=None):
        for i, c in enumerate(self.chromosomes):
            if c.fitness is None:
                c.evaluate()
                if callback is not None:
                    callback(i, c)

    # -- selection / crossover ---------------------------------------------------------------------------------------
    def select_roulette(self):
        """Fitness-proportional choice of parents (fitness shifted to be positive)."""
        fits = numpy.array([c.fitness for c in self.chromosomes], dtype=numpy.float64)
        shifted = fits - fits.min() + 1e-12 + 0.05 * (numpy.ptp(fits) or 1.0)
        probs = shifted / shifted.sum()
        n = max(2, int(round(self.roulette_select_size * len(self.chromosomes))))
        idx = self.rand.choice(len(self.chromosomes), size=n, replace=True, p=probs)
        return [self.chromosomes[int(i)] for i in idx]

    def cross_uniform(self, a, b):
        child = a.copy()
        child.numeric = [x if self.rand.random_sample() < 0.5 else y for x, y in zip(a.numeric, b.numeric)]
        child.fitness = None
        return child

    def cross_arithmetic(self, a, b):
        child = a.copy()
        t = float(self.rand.random_sample())
        child.numeric = [t * x + (1.0 - t) * y for x, y in zip(a.numeric, b.numeric)]
        child.fitness = None
        return child

    def cross_geometric(self, a, b):
        child = a.copy()
        t = float(self.rand.random_sample())
        out = []
        for x, y, lo in zip(a.numeric, b.numeric, self.min_values):
            sx, sy = x - lo + 1e-12, y - lo + 1e-12   # shift into the positive domain
            out.append(lo + (sx ** t) * (sy ** (1.0 - t)) - 1e-12)
        child.numeric = out
        child.clip()
        child.fitness = None
        return child

    def cross(self, parents):
        children = []
        names = list(self.crossings)
        for _ in range(self.crossing_attempts * len(parents)):
            if len(children) >= self.size - self.elite:
                break
            a = parents[int(self.rand.randint(0, len(parents)))]
            b = parents[int(self.rand.randint(0, len(parents)))]
            if a is b:
                continue
            name = names[int(self.rand.randint(0, len(names)))]
            if self.rand.random_sample() < self.crossings[name]:
                children.append(getattr(self, "cross_" + name)(a, b))
        return children

    def mutate(self, chromos):
        for c in chromos:
            for name, p in self.mutations.items():
                if p.get("probability", 0) > 0:
                    getattr(c, "mutate_" + name)(p.get("points", 1), p["probability"])

    # -- one generation -----------------------------------------------------------------------------------------------------
    def sort(self):
        self.chromosomes.sort(key=lambda c: c.fitness, reverse=True)

    def update(self):
        """All chromosomes are evaluated: record statistics and breed the next generation.
        Returns True when the evolution should stop."""
        self.sort()
        fits = [c.fitness for c in self.chromosomes]
        self.best_fit, self.worst_fit = fits[0], fits[-1]
        self.average_fit = float(numpy.mean(fits))
        self.median_fit = float(numpy.median(fits))
        self.history.append({"generation": self.generation, "best": self.best_fit, "average": self.average_fit,
                             "genes": list(self.chromosomes[0].numeric)})
        if self.prev_best is None or self.best_fit > self.prev_best + abs(self.prev_best) * 1e-9:
            self.prev_best = self.best_fit
            self.improved_generations_ago = 0
        else:
            self.improved_generations_ago += 1
        self.generation += 1
        self.on_generation_changed()
        if self.max_generations is not None and self.generation >= self.max_generations:
            return True
        if self.improved_generations_ago >= self.no_improvement_limit:
            return True
        elite = [c for c in self.chromosomes[:self.elite]]
        parents = self.select_roulette()
        children = self.cross(parents)
        self.mutate(children)
        while len(children) < self.size - len(elite):
            children.append(self.new(self.optimization_size))
        self.chromosomes = elite + children[:self.size - len(elite)]
        return False

    def on_generation_changed(self):
        self.info("Generation %d: best fitness %.6g, average %.6g", self.generation, self.best_fit, self.average_fit)

    def evolve(self):
        """Blocking evolution loop (chromosomes evaluate themselves)."""
        while True:
            self.evaluate()
            if self.update():
                break
        return self.chromosomes[0]
