"""Step tracer (the reference gets smdebug's profiler report attached by SageMaker, ``notebooks/
2_pytorch_dist_smddp_gpu.ipynb:166,1291-1298``; it has no tracing code of its own).

``Tracer`` records named ranges twice: on the host (``time.perf_counter``) and on the device (a pair of CUDA events on the
current stream), emits an NVTX range for external tools, and writes a Chrome-trace JSON (``chrome://tracing`` / Perfetto) with
one row per rank for host time and one for device time.  It is meant for eager steps and for whole-graph replays; ranges opened
inside a CUDA-graph capture are ignored (events cannot be recorded into a capture without becoming part of it).

    tr = Tracer(rank=dist.get_rank())
    with tr.range("step"):
        with tr.range("h2d"): ...
        with tr.range("replay"): graph.replay()
    tr.save("trace_rank0.json")
"""
from __future__ import annotations

import contextlib
import json
import time
from typing import List, Optional

import torch


class Tracer:
    def __init__(self, rank: int = 0, enabled: bool = True, nvtx: bool = True):
        self.rank = rank
        self.enabled = enabled
        self.cuda = torch.cuda.is_available()
        self.nvtx = nvtx and self.cuda
        self._t0 = time.perf_counter()
        self._host: List[tuple] = []          # (name, start_us, dur_us, depth)
        self._dev: List[tuple] = []           # (name, start_event, stop_event, depth)
        self._origin: Optional[torch.cuda.Event] = None
        self._depth = 0

    @contextlib.contextmanager
    def range(self, name: str):
        if not self.enabled:
            yield
            return
        capturing = self.cuda and torch.cuda.is_current_stream_capturing()
        t_start = time.perf_counter()
        ev0 = ev1 = None
        if self.cuda and not capturing:
            if self._origin is None:
                self._origin = torch.cuda.Event(enable_timing=True)
                self._origin.record()
            ev0 = torch.cuda.Event(enable_timing=True)
            ev0.record()
            if self.nvtx:
                torch.cuda.nvtx.range_push(name)
        depth = self._depth
        self._depth += 1
        try:
            yield
        finally:
            self._depth -= 1
            if ev0 is not None:
                if self.nvtx:
                    torch.cuda.nvtx.range_pop()
                ev1 = torch.cuda.Event(enable_timing=True)
                ev1.record()
                self._dev.append((name, ev0, ev1, depth))
            t_end = time.perf_counter()
            self._host.append((name, (t_start - self._t0) * 1e6, (t_end - t_start) * 1e6, depth))

    def events(self) -> List[dict]:
        """Chrome-trace events (``ph: X``); device rows are resolved after a synchronize."""
        out = []
        pid = self.rank
        for name, ts, dur, depth in self._host:
            out.append({"name": name, "ph": "X", "pid": pid, "tid": 0, "ts": ts, "dur": dur, "args": {"depth": depth}})
        if self._dev:
            torch.cuda.synchronize()
            for name, e0, e1, depth in self._dev:
                out.append({"name": name, "ph": "X", "pid": pid, "tid": 1, "ts": self._origin.elapsed_time(e0) * 1e3,
                            "dur": e0.elapsed_time(e1) * 1e3, "args": {"depth": depth}})
        out.append({"name": "process_name", "ph": "M", "pid": pid, "args": {"name": "rank %d" % self.rank}})
        out.append({"name": "thread_name", "ph": "M", "pid": pid, "tid": 0, "args": {"name": "host"}})
        out.append({"name": "thread_name", "ph": "M", "pid": pid, "tid": 1, "args": {"name": "device (current stream)"}})
        return out

    def summary_ms(self) -> dict:
        """Total device milliseconds per range name (host milliseconds without a GPU)."""
        tot: dict = {}
        if self._dev:
            torch.cuda.synchronize()
            for name, e0, e1, _ in self._dev:
                tot[name] = tot.get(name, 0.0) + e0.elapsed_time(e1)
        else:
            for name, _, dur, _ in self._host:
                tot[name] = tot.get(name, 0.0) + dur / 1e3
        return tot

    def save(self, path: str) -> str:
        with open(path, "w") as fh:
            json.dump({"traceEvents": self.events(), "displayTimeUnit": "ms"}, fh)
        return path
