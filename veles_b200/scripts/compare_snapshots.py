"""Diff two snapshots of the same workflow, Array by Array (reference: ``veles/scripts/compare_snapshots.py:82-146``):
for every Array attribute of every unit the average relative difference (2 |a - b| / |a + b| over the elements whose
sum is not zero), the average and the maximum absolute difference, as a table sorted by the chosen keys.

    python -m veles_b200.scripts.compare_snapshots [-q] [-s KEY [KEY ...]] first.pickle.gz second.pickle.gz

Sort keys: ``dep`` (position of the unit in dependency order), ``unit``, ``attr``, ``avgreldiff``, ``avgdiff``,
``maxdiff``; default ``dep avgreldiff``.  Snapshots of different workflows (class or checksum) are refused."""
from __future__ import annotations

import argparse
import logging
import sys

import numpy

SORT_CHOICES = ("dep", "unit", "attr", "avgreldiff", "avgdiff", "maxdiff")
SORT_INDEX = {k: i for i, k in enumerate(SORT_CHOICES)}


def _units(workflow):
    order = getattr(workflow, "units_in_dependency_order", None)
    return list(order() if callable(order) else order) if order is not None else list(workflow.units)


def array_diffs(first, second):
    """Yield (dep index, unit name, attribute, avg relative diff, avg diff, max diff) for every Array both have;
    attributes missing or differently shaped on one side yield NaN statistics and a note in the attribute name."""
    from ..memory import Array
    second_by_name = {}
    for u in _units(second):
        second_by_name.setdefault(u.name, u)
    for index, fu in enumerate(_units(first)):
        su = second_by_name.get(fu.name)
        for key, fv in sorted(fu.__dict__.items()):
            if not isinstance(fv, Array) or fv.mem is None:
                continue
            sv = getattr(su, key, None) if su is not None else None
            if not isinstance(sv, Array) or sv.mem is None:
                yield index, fu.name, key.lstrip("_") + " (only in first)", numpy.nan, numpy.nan, numpy.nan
                continue
            fv.map_read()
            sv.map_read()
            a, b = numpy.asarray(fv.mem, dtype=numpy.float64), numpy.asarray(sv.mem, dtype=numpy.float64)
            if a.shape != b.shape:
                yield (index, fu.name, "%s (shape %s vs %s)" % (key.lstrip("_"), a.shape, b.shape), numpy.nan, numpy.nan,
                       numpy.nan)
                continue
            if a.size == 0:
                continue
            diff = a - b
            total = a + b
            nz = total != 0
            if nz.any():
                avg_rel = float(numpy.mean(numpy.abs(2.0 * diff[nz] / total[nz])))
            else:
                avg_rel = float(bool(diff.any()))
            yield (index, fu.name, key.lstrip("_"), avg_rel, float(numpy.mean(numpy.abs(diff))),
                   float(numpy.max(numpy.abs(diff))))


def sort_diffs(diffs, sorting):
    def key(rec):
        return tuple((rec[SORT_INDEX[k]] if not (isinstance(rec[SORT_INDEX[k]], float) and rec[SORT_INDEX[k]] !=
                                                   rec[SORT_INDEX[k]]) else float("inf")) for k in sorting)
    return sorted(diffs, key=key)


def format_table(diffs):
    head = ("Unit", "Attribute", "Average Relative Diff", "Average Diff", "Max Diff")
    rows = [(d[1], d[2]) + tuple("-" if v != v else "%.6g" % v for v in d[3:]) for d in diffs]
    widths = [max(len(str(r[i])) for r in [head] + rows) for i in range(5)]
    sep = "+" + "+".join("-" * (w + 2) for w in widths) + "+"
    fmt = "| " + " | ".join("%%-%ds" % w if i < 2 else "%%%ds" % w for i, w in enumerate(widths)) + " |"
    return "\n".join([sep, fmt % head, sep] + [fmt % r for r in rows] + [sep])


def compare(first, second, sorting=("dep", "avgreldiff")):
    if type(first).__name__ != type(second).__name__ or \
            getattr(first, "checksum", None) != getattr(second, "checksum", None):
        raise ValueError("Cannot compare different workflows (%s %s vs %s %s)" % (
            type(first).__name__, getattr(first, "checksum", None), type(second).__name__,
            getattr(second, "checksum", None)))
    return sort_diffs(array_diffs(first, second), sorting)


def main(argv=None):
    from ..snapshotter import SnapshotterToFile
    ap = argparse.ArgumentParser(prog="compare_snapshots", description="Compare two snapshots of one workflow.")
    ap.add_argument("-q", "--quiet", action="store_true", help="do not print progress logs")
    ap.add_argument("-s", "--sort", choices=SORT_CHOICES, nargs="+", default=["dep", "avgreldiff"],
                    help="sort by these fields, in order")
    ap.add_argument("first")
    ap.add_argument("second")
    argv = sys.argv[1:] if argv is None else list(argv)
    if len([a for a in argv if not a.startswith("-")]) < 2:
        ap.print_usage()
        print(__doc__.split("\n\n")[0])
        return 1
    args = ap.parse_args(argv)
    log = logging.getLogger("compare_snapshots")
    if not args.quiet:
        logging.basicConfig(level=logging.INFO)
    log.info("Loading snapshots...")
    try:
        first, second = (SnapshotterToFile.import_(p) for p in (args.first, args.second))
    except Exception as e:  # noqa: BLE001
        log.critical("Failed to load a snapshot: %s", e)
        return 2
    log.info("Comparing...")
    try:
        diffs = compare(first, second, args.sort)
    except ValueError as e:
        log.critical("%s", e)
        return 3
    print(format_table(diffs))
    return 0


if __name__ == "__main__":
    sys.exit(main())
