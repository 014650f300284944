"""Helpers shared by ``bench.py`` and the bench scripts: clock sampling during the
timed region, device-side timing with max-over-ranks, pinned host batches.

The reference's only measurement harness is ``torch.profiler`` around 8 steps (``main.py:187-220``; SURVEY.md §5.1,
§5.5); these helpers implement the protocol BASELINE.md §3 states instead."""
from __future__ import annotations

import os
import shutil
import statistics
import subprocess
import tempfile
import time
from typing import Any, Callable, Dict, List, Optional

import torch
import torch.distributed as dist

__all__ = ["ClockSampler", "timed_steps", "max_over_ranks", "dist_ready"]

_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
          "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
          "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")


class ClockSampler:
    """Runs ``nvidia-smi -lms 200`` next to the timed region and summarises SM
    clocks and throttle reasons (B200_PROFILING.md "clocks line")."""

    def __init__(self, gpu_index: int = 0) -> None:
        self.gpu_index = gpu_index
        self._proc: Optional[subprocess.Popen] = None
        self._path: Optional[str] = None

    def start(self) -> None:
        exe = shutil.which("nvidia-smi")
        if exe is None:
            return
        fd, self._path = tempfile.mkstemp(prefix="clocks_", suffix=".csv")
        os.close(fd)
        out = open(self._path, "w")
        try:
            self._proc = subprocess.Popen(
                [exe, f"--query-gpu={_QUERY}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu_index)], stdout=out, stderr=subprocess.DEVNULL)
        except Exception:
            self._proc = None

    def stop(self) -> Dict[str, Any]:
        if self._proc is None or self._path is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self._proc.terminate()
        try:
            self._proc.wait(timeout=5)
        except Exception:
            self._proc.kill()
        sm: List[float] = []
        sm_max = None
        reasons = set()
        power: List[float] = []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        with open(self._path) as f:
            for line in f:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    sm_max = float(parts[2])
                    power.append(float(parts[3]))
                except ValueError:
                    continue
                for name, val in zip(names, parts[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
        try:
            os.unlink(self._path)
        except OSError:
            pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": sm_max,
                "power_w_max": max(power) if power else None, "samples": len(sm),
                "reasons": sorted(reasons)}


def dist_ready() -> bool:
    return dist.is_available() and dist.is_initialized()


def max_over_ranks(value: float, device: torch.device) -> float:
    if not dist_ready():
        return value
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed_steps(step: Callable[[int], Any], steps: int, device: torch.device) -> float:
    """Time exactly ``steps`` calls of ``step`` with CUDA events on the current
    stream, bracketed by barrier + synchronize on both sides; returns the
    max-over-ranks elapsed milliseconds."""
    if dist_ready():
        dist.barrier()
    torch.cuda.synchronize(device)
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    t0.record(torch.cuda.current_stream(device))
    for k in range(steps):
        step(k)
    t1.record(torch.cuda.current_stream(device))
    torch.cuda.synchronize(device)
    if dist_ready():
        dist.barrier()
    torch.cuda.synchronize(device)
    return max_over_ranks(t0.elapsed_time(t1), device)
