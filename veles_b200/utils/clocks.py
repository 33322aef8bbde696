"""Background nvidia-smi sampler: SM clocks and throttle reasons DURING a timed region
(B200_PROFILING.md "clocks line")."""
from __future__ import annotations

import subprocess
import threading

QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


# NVML clocks-event-reason bits (nvml.h)
_REASON_BITS = {"sw_power_cap": 0x4, "hw_slowdown": 0x8, "sw_thermal_slowdown": 0x20, "hw_thermal_slowdown": 0x40}


class ClockSampler(object):
    """Samples through NVML (pynvml, ~2 ms period: short timed regions still get tens of samples) and falls back to
    a background ``nvidia-smi -lms`` process when the bindings are missing."""

    def __init__(self, gpu_index=0, period_ms=200):
        self.gpu_index = gpu_index
        self.period_ms = period_ms
        self.proc = None
        self.lines = []
        self._thread = None
        self._stop = threading.Event()
        self._nvml = None
        self._samples = []   # (sm_mhz, power_w, reasons bitmask)
        self._smax = None

    def _start_nvml(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(int(self.gpu_index))
            self._smax = float(pynvml.nvmlDeviceGetMaxClockInfo(h, pynvml.NVML_CLOCK_SM))
        except Exception:  # noqa: BLE001
            return False
        self._nvml = pynvml

        def loop():
            import time
            get_reasons = getattr(pynvml, "nvmlDeviceGetCurrentClocksEventReasons",
                                  getattr(pynvml, "nvmlDeviceGetCurrentClocksThrottleReasons", None))
            while not self._stop.is_set():
                try:
                    sm = float(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM))
                    pw = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                    rs = int(get_reasons(h)) if get_reasons else 0
                    self._samples.append((sm, pw, rs))
                except Exception:  # noqa: BLE001
                    pass
                time.sleep(0.002)
        self._thread = threading.Thread(target=loop, daemon=True)
        self._thread.start()
        return True

    def start(self):
        if self._start_nvml():
            return self
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + QUERY, "--format=csv,noheader,nounits",
                 "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return self
        self._thread = threading.Thread(target=self._pump, daemon=True)
        self._thread.start()
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self._nvml is not None:
            self._stop.set()
            self._thread.join(timeout=1)
            return self.summary()
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:  # noqa: BLE001
                self.proc.kill()
        return self.summary()

    def summary(self):
        if self._nvml is not None:
            if not self._samples:
                return {"sm_mhz": None, "sm_max_mhz": self._smax, "reasons": [], "samples": 0, "source": "nvml"}
            sms = sorted(s[0] for s in self._samples)
            bits = 0
            for s in self._samples:
                bits |= s[2]
            return {"sm_mhz": sms[len(sms) // 2], "sm_max_mhz": self._smax,
                    "power_w_max": round(max(s[1] for s in self._samples), 2),
                    "reasons": sorted(n for n, b in _REASON_BITS.items() if bits & b), "samples": len(self._samples),
                    "source": "nvml"}
        sm, smax, power = [], [], []
        reasons = set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                smax.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, val in zip(names, parts[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm_sorted = sorted(sm)
        return {"sm_mhz": sm_sorted[len(sm_sorted) // 2], "sm_max_mhz": max(smax), "power_w_max": max(power),
                "reasons": sorted(reasons), "samples": len(sm)}
