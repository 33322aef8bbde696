"""python scripts/roofline_report.py [--model alexnet] [--out profiles/roofline_alexnet.md]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="alexnet")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    from veles_b200.backends import B200Device
    from veles_b200.trainer import Trainer
    from veles_b200.utils import roofline
    tr = Trainer(a.model, minibatch_size=a.batch or None, device=B200Device(), use_graph=False)
    rows = roofline.trace_step(tr)
    text = roofline.report(rows, roofline.load_peaks(), "%s training step, batch %d" % (a.model, tr.minibatch_size))
    out = a.out or os.path.join("gpurun_out", "roofline_%s.md" % a.model)
    os.makedirs(os.path.dirname(out) or ".", exist_ok=True)
    with open(out, "w") as f:
        f.write(text)
    print(text[-1500:])
    tr.close()
