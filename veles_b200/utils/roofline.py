"""Per-launch roofline report: every kernel wrapper call of one training step is bracketed by CUDA events
(``ops.kernels.TRACE``) and compared with the measured peaks of the box (``MEASURED_PEAKS.json``: copy bandwidth
and cuBLAS bf16 throughput, written by the driver).

Algorithmic bytes = sizes of the tensors passed to the call (each counted once: what the op must at least read
and write); FLOPs = 2*M*N*K of GEMM-shaped calls (implicit convolutions count their real taps x channels).
"""
from __future__ import annotations

import json
import os

FALLBACK = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "source": "fallback of B200_PROFILING.md"}


def load_peaks(root=None):
    root = root or os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    path = os.path.join(root, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            d = json.load(f)
        d["source"] = "MEASURED_PEAKS.json"
        return d
    return dict(FALLBACK)


def trace_step(trainer, warmup=3):
    """Run ``warmup`` eager steps, then one traced eager step; returns [(name, us, bytes, flops, shapes)]."""
    import torch
    from ..ops import kernels
    for _ in range(warmup):
        trainer.stepper.run(train=True, serve=True)
    torch.cuda.synchronize()
    kernels.TRACE = []
    try:
        # keep the GPU busy while the host enqueues the whole step: the traced kernels and their events then run
        # back to back, without the host's per-call launch latency inside the event brackets
        torch.cuda._sleep(400000000)
        trainer.stepper.run(train=True, serve=True)
        torch.cuda.synchronize()
        rows = [(n, e0.elapsed_time(e1) * 1e3, b, fl, sh) for n, e0, e1, b, fl, sh in kernels.TRACE]
    finally:
        kernels.TRACE = None
    return rows


def report(rows, peaks, title="step"):
    hbm = peaks["hbm_gbs"] * 1e9
    # kernels timed inside a step are compared with the SUSTAINED cuBLAS rate (the burst figure is for a kernel timed
    # alone on a cold chip)
    tf_name = "bf16_tflops_sustained" if peaks.get("bf16_tflops_sustained") else "bf16_tflops"
    tf = peaks[tf_name] * 1e12
    lines = ["# Roofline of one %s (CUDA events around every kernel call)" % title, "",
             "Peaks (of measured): %.1f GB/s copy bandwidth, %.1f TFLOP/s bf16 (%s of %s).  `bound` = the larger of"
             " bytes/BW and FLOPs/peak; `frac` = bound time / measured time.  The host enqueues the traced step behind a"
             " long spin kernel, so the event brackets hold device time only." % (
                 peaks["hbm_gbs"], peaks[tf_name], tf_name, peaks.get("source", "")), "",
             "| # | call | first shapes | us | MB | GFLOP | bound us | frac |", "|---|---|---|---|---|---|---|---|"]
    tot = tot_bound = 0.0
    agg = {}
    for i, (n, us, b, fl, sh) in enumerate(rows):
        bound = max(b / hbm, fl / tf) * 1e6
        tot += us
        tot_bound += bound
        a = agg.setdefault(n, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += us
        a[2] += bound
        lines.append("| %d | %s | %s | %.1f | %.1f | %.2f | %.1f | %.0f%% |" % (
            i, n, " ".join("x".join(map(str, s)) for s in sh), us, b / 1e6, fl / 1e9, bound, 100 * bound / max(us, 1e-3)))
    lines += ["", "Total %.1f us measured (event-bracketed, eager), %.1f us at the roofline -> %.0f%% overall." % (
        tot, tot_bound, 100 * tot_bound / max(tot, 1e-3)), "", "| call | n | us | share | frac of roofline |", "|---|---|---|---|---|"]
    for n, (c, us, bd) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        lines.append("| %s | %d | %.1f | %.1f%% | %.0f%% |" % (n, c, us, 100 * us / tot, 100 * bd / max(us, 1e-3)))
    return "\n".join(lines) + "\n"
