"""Clock sampling for timed regions.

Phase timings themselves are CUDA-event durations reduced with max over ranks (``TrainStep.breakdown``,
``parallel.process_group.max_over_ranks``); the only wall-clock thing here is the nvidia-smi sampler.
"""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from typing import List, Optional



class ClockSampler:
    """Samples ``nvidia-smi`` SM clocks and throttle reasons in a background thread during a timed
    region (B200_PROFILING.md "clocks DURING the timed region")."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_s: float = 0.2):
        self.gpu_index, self.period_s = gpu_index, period_s
        self.samples: List[List[str]] = []
        self._stop = threading.Event()
        self._th: Optional[threading.Thread] = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.gpu_index), "--query-gpu=" + self.FIELDS,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                row = [c.strip() for c in out.strip().split(",")]
                if len(row) >= 7:
                    self.samples.append(row)
            except Exception:
                pass
            self._stop.wait(self.period_s)

    def __enter__(self):
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._th is not None:
            self._th.join(timeout=10)

    def summary(self) -> dict:
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm = [float(r[0]) for r in self.samples if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.samples if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(r[3 + i].lower().startswith("active") for r in self.samples)]
        pw = [float(r[2]) for r in self.samples if r[2].replace(".", "").isdigit()]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(self.samples), "power_w_max": max(pw) if pw else None}
