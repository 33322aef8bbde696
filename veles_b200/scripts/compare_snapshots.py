"""Diff two workflow snapshots: a table of every Array (weights, bias, momentum, ...) with max |a-b|
(reference: ``veles/scripts/compare_snapshots.py:82-146``).
    python -m veles_b200.scripts.compare_snapshots first.pickle.gz second.pickle.gz"""
from __future__ import annotations

import sys

import numpy


def arrays_of(workflow):
    from ..memory import Array
    out = {}
    for unit in workflow.units:
        for k, v in sorted(unit.__dict__.items()):
            if isinstance(v, Array) and v:
                out["%s.%s" % (unit.name, k.lstrip("_"))] = v.mem
    return out


def compare(wf_a, wf_b):
    a, b = arrays_of(wf_a), arrays_of(wf_b)
    rows = []
    for key in sorted(set(a) | set(b)):
        if key not in a or key not in b:
            rows.append((key, "only in %s" % ("first" if key in a else "second"), None))
        elif a[key].shape != b[key].shape:
            rows.append((key, "shape %s vs %s" % (a[key].shape, b[key].shape), None))
        else:
            diff = float(numpy.abs(a[key].astype(numpy.float64) - b[key].astype(numpy.float64)).max()) if a[key].size else 0.0
            rows.append((key, "x".join(map(str, a[key].shape)), diff))
    return rows


def main(argv=None):
    from ..snapshotter import SnapshotterToFile
    argv = argv if argv is not None else sys.argv[1:]
    if len(argv) != 2:
        print(__doc__)
        return 1
    rows = compare(SnapshotterToFile.import_(argv[0]), SnapshotterToFile.import_(argv[1]))
    width = max(len(r[0]) for r in rows) if rows else 10
    for key, shape, diff in rows:
        print("%-*s  %-18s %s" % (width, key, shape, "" if diff is None else "%.6g" % diff))
    return 0


if __name__ == "__main__":
    sys.exit(main())
