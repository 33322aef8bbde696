#!/usr/bin/env python3
"""Times the batched signal kernels (veles_b200.simd on CUDA tensors) with CUDA events after warm-up; inputs are
larger than L2 or L2 is flushed between iterations.  Prints one JSON object (also written to --out)."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np   # noqa: E402
import torch         # noqa: E402

from veles_b200 import simd   # noqa: E402


def timed(fn, flush, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/signal_microbench.json")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    peaks_path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    hbm = 6546.6
    if os.path.exists(peaks_path):
        try:
            hbm = float(json.load(open(peaks_path)).get("hbm_gbs", hbm))
        except Exception:
            pass
    res = {"hbm_copy_gbps": hbm, "cases": []}

    def add(name, ms, bytes_moved, flops=0.0, **extra):
        row = {"name": name, "ms": round(ms, 4), "GBps": round(bytes_moved / ms / 1e6, 1),
               "pct_of_copy_bw": round(100 * bytes_moved / ms / 1e6 / hbm, 1)}
        if flops:
            row["GFLOPs"] = round(flops / ms / 1e6, 1)
        row.update(extra)
        res["cases"].append(row)

    B, n = 4096, 16384
    x = torch.randn(B, n, device=dev)
    for order in (8, 20, 76):
        ms = timed(lambda: simd.wavelet_apply("daubechies", order, simd.EXT_PERIODIC, x), flush)
        add("dwt_db%d" % order, ms, 2.0 * x.numel() * 4, 4.0 * order * x.numel() / 2, shape=[B, n])
    ms = timed(lambda: simd.stationary_wavelet_apply("daubechies", 8, 3, simd.EXT_PERIODIC, x), flush)
    add("swt_db8_level3", ms, 3.0 * x.numel() * 4, 4.0 * 8 * x.numel(), shape=[B, n])
    for taps in (16, 128, 1024):
        h = torch.randn(taps, device=dev)
        ms = timed(lambda: simd.convolve(x, h, simd.CONV_BRUTE), flush, iters=10)
        add("conv_direct_%dtaps" % taps, ms, 2.0 * x.numel() * 4, 2.0 * taps * x.numel(), shape=[B, n])
    h = torch.randn(4096, device=dev)
    xs = x[:512]
    ms = timed(lambda: simd.convolve(xs, h, simd.CONV_OVERLAP_SAVE), flush, iters=10)
    add("conv_overlap_save_4096taps", ms, 2.0 * xs.numel() * 4, shape=[512, n])
    ms = timed(lambda: simd.convolve(xs, h, simd.CONV_FFT), flush, iters=10)
    add("conv_fft_4096taps", ms, 2.0 * xs.numel() * 4, shape=[512, n])
    ms = timed(lambda: simd.detect_peaks(x, simd.EXTREMUM_BOTH), flush)
    add("detect_peaks_both", ms, 2.0 * x.numel() * 4 + x.numel() * 4, shape=[B, n])
    ms = timed(lambda: simd.minmax1D(x), flush)
    add("minmax1D", ms, 1.0 * x.numel() * 4, shape=[B, n])
    img = torch.randint(0, 255, (2048, 256, 256), dtype=torch.uint8, device=dev)
    ms = timed(lambda: simd.normalize2D(img), flush)
    add("normalize2D_u8", ms, img.numel() * (1 + 1 + 4), shape=list(img.shape))
    i16 = torch.randint(-30000, 30000, (B * n,), dtype=torch.int16, device=dev)
    ms = timed(lambda: simd.int16_to_float(i16), flush)
    add("int16_to_float", ms, i16.numel() * 6, shape=[B * n])
    os.makedirs(os.path.dirname(args.out) or ".", exist_ok=True)
    with open(args.out, "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
