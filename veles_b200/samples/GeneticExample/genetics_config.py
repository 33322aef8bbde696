"""Search space of the GeneticExample sample: two tunable values (default, minimum, maximum)."""
from veles_b200.config import root
from veles_b200.genetics import Range

root.test.update({
    "x": Range(0.0, -5.0, 5.0),
    "y": Range(0.0, -5.0, 5.0),
})
