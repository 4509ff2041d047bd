"""Device-side timing, clock sampling and JSON-lines metrics (SURVEY.md §5 rows
"Tracing / profiling" and "Metrics / logging").

The reference times with ``time.time()`` around one call and prints the mean of one
sample (notebooks/cv/onnx_experiments.py:92-95, :104, :133-140).  Here every number is
taken with CUDA events on the launching stream, bracketed by a synchronise, and the
slowest rank is reported (BASELINE.json:5 measurement rules).
"""
from __future__ import annotations

import json
import os
import shutil
import statistics
import subprocess
import threading
import time
from typing import Any, Callable, Dict, List, Optional

import torch

from ..parallel import process_group as pg


class DeviceTimer:
    """``with DeviceTimer() as t: ...`` -> ``t.ms`` (CUDA events; wall clock on CPU)."""

    def __init__(self, device: Optional[torch.device] = None, sync: bool = True):
        self.cuda = torch.cuda.is_available() and (device is None or device.type == "cuda")
        self.sync = sync
        self.ms = float("nan")

    def __enter__(self):
        if self.cuda:
            if self.sync:
                torch.cuda.synchronize()
            self._s, self._e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self._s.record()
        else:
            self._t0 = time.perf_counter()
        return self

    def __exit__(self, *exc):
        if self.cuda:
            self._e.record()
            self._e.synchronize()
            self.ms = self._s.elapsed_time(self._e)
        else:
            self.ms = (time.perf_counter() - self._t0) * 1e3
        return False


def time_kernel(fn: Callable[[], Any], warmup: int = 3, iters: int = 20, flush_l2: bool = True) -> Dict[str, float]:
    """Per-call device time of ``fn`` (median/min over ``iters``), L2 flushed in between by
    writing a 256 MiB buffer (> the 126 MB L2), as the profiling recipe requires."""
    assert torch.cuda.is_available()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if flush_l2 else None
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts: List[float] = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e))
    return {"median_ms": statistics.median(ts), "min_ms": min(ts), "iters": iters}


_SMI_FIELDS = (
    "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
    "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
    "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
)


class ClockSampler:
    """Samples ``nvidia-smi`` clocks/throttle reasons in the background during a timed
    region (the recipe's clocks line) and summarises them for the bench JSON."""

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self._proc: Optional[subprocess.Popen] = None
        self._lines: List[str] = []
        self._thread: Optional[threading.Thread] = None

    def start(self) -> "ClockSampler":
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return self
        self._proc = subprocess.Popen(
            [exe, f"--query-gpu={_SMI_FIELDS}", "--format=csv,noheader,nounits", "-i", str(self.gpu_index),
             "-lms", str(self.period_ms)],
            stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
        )

        def pump():
            assert self._proc is not None and self._proc.stdout is not None
            for line in self._proc.stdout:
                self._lines.append(line.strip())

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()
        return self

    def stop(self) -> Dict[str, Any]:
        if self._proc is not None:
            self._proc.terminate()  # the exact process we started
            try:
                self._proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self._proc.kill()
            if self._thread is not None:
                self._thread.join(timeout=2)
        sm: List[float] = []
        smax: List[float] = []
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self._lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                smax.append(float(parts[2]))
            except ValueError:
                continue
            for n, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {
            "sm_mhz": statistics.median(sm) if sm else None,
            "sm_max_mhz": max(smax) if smax else None,
            "reasons": sorted(reasons),
            "samples": len(sm),
        }


class JsonlLogger:
    """One JSON object per line; rank 0 writes (SURVEY.md §5 "Metrics")."""

    def __init__(self, path: Optional[str] = None, also_print: bool = False):
        self.path, self.also_print = path, also_print
        if path and pg.rank() == 0:
            os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)

    def log(self, **record: Any) -> None:
        if pg.rank() != 0:
            return
        record.setdefault("ts", time.time())
        line = json.dumps(record, default=str)
        if self.path:
            with open(self.path, "a") as f:
                f.write(line + "\n")
        if self.also_print:
            print(line, flush=True)


def load_measured_peaks(repo_root: Optional[str] = None) -> Dict[str, float]:
    """Roofline denominators: MEASURED_PEAKS.json if present, else the recipe's fallback."""
    root = repo_root or os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    p = os.path.join(root, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            d = json.load(f)
        d["source"] = "measured"
        return d
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}
