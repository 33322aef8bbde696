"""Turn an `ncu --metrics gpu__time_duration.sum --csv` log into a per-kernel table and one step's launch sequence."""
import collections
import csv
import sys


def load(path):
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    seq = []
    for row in csv.DictReader(lines):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        unit = row["Metric Unit"]
        v = v / 1000 if unit == "ns" else (v * 1000 if unit == "ms" else v)
        seq.append((row["Kernel Name"], v, row.get("Grid Size", "")))
    return seq


def main(path, step_marker="gather_rows"):
    seq = load(path)
    starts = [i for i, s in enumerate(seq) if step_marker in s[0]]
    if len(starts) >= 3:
        a, b = starts[1], starts[2]
    else:
        a, b = 0, len(seq)
    step = seq[a:b]
    agg = collections.OrderedDict()
    for n, v, g in step:
        e = agg.setdefault(n.split("(")[0][:64], [0, 0.0])
        e[0] += 1
        e[1] += v
    tot = sum(e[1] for e in agg.values())
    print("one step: %d launches, %.1f us" % (len(step), tot))
    print("| kernel | launches | us | share |\n|---|---|---|---|")
    for kname, e in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("| %s | %d | %.1f | %.1f%% |" % (kname, e[0], e[1], 100 * e[1] / tot))
    if "--seq" in sys.argv:
        for n, v, g in step:
            print("%-64s %8.1f %s" % (n.split("(")[0][:64], v, g))


if __name__ == "__main__":
    main(sys.argv[1])
