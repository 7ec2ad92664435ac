"""Device timing helpers: CUDA events on the launching stream, max over ranks, L2 flush,
and an nvidia-smi clock sampler (profiling recipe, /opt/skills/guides/B200_PROFILING.md).
The reference only has ``time.time()`` stamps."""
from __future__ import annotations

import statistics
import subprocess
import threading
import time
from contextlib import contextmanager
from typing import Callable, Dict, List, Optional

import torch
import torch.distributed as dist

__all__ = ["DeviceTimer", "time_steps", "max_over_ranks", "L2Flusher", "ClockSampler", "nvtx_range"]


class DeviceTimer:
    """CUDA-event stopwatch (falls back to perf_counter on CPU)."""

    def __init__(self, device: Optional[torch.device] = None):
        self.cuda = torch.cuda.is_available() and (device is None or torch.device(device).type == "cuda")
        self._t0 = None

    def start(self):
        if self.cuda:
            self._e0, self._e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self._e0.record()
        else:
            self._t0 = time.perf_counter()

    def stop(self) -> float:
        """Milliseconds since start()."""
        if self.cuda:
            self._e1.record()
            self._e1.synchronize()
            return float(self._e0.elapsed_time(self._e1))
        return (time.perf_counter() - self._t0) * 1e3


def max_over_ranks(value: float, device=None) -> float:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
        t = torch.tensor([value], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return float(value)


def _barrier():
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def time_steps(step_fn: Callable[[int], None], steps: int, warmup: int = 3, between: Optional[Callable] = None) -> Dict:
    """Warm up, then time exactly ``steps`` calls bracketed by barrier+synchronize on both sides.
    Returns total/mean ms as the max over ranks."""
    for i in range(warmup):
        step_fn(i)
    cuda = torch.cuda.is_available()
    if cuda:
        torch.cuda.synchronize()
    _barrier()
    timer = DeviceTimer()
    timer.start()
    for i in range(steps):
        if between is not None:
            between()
        step_fn(warmup + i)
    total_ms = timer.stop()
    if cuda:
        torch.cuda.synchronize()
    _barrier()
    total_ms = max_over_ranks(total_ms)
    return {"total_ms": total_ms, "ms_per_step": total_ms / max(1, steps)}


class L2Flusher:
    """Writes a buffer larger than the 126 MB L2 between timed iterations."""

    def __init__(self, nbytes: int = 256 << 20, device="cuda"):
        self.buf = torch.empty(nbytes // 4, dtype=torch.int32, device=device)

    def __call__(self):
        self.buf.add_(1)


class ClockSampler:
    """Samples SM clocks / throttle reasons with nvidia-smi while a timed region runs."""
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_ms: int = 200):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.proc = None
        self.lines: List[str] = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-i", str(self.gpu_index), "-lms", str(self.period_ms)],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self._thr = threading.Thread(target=self._pump, daemon=True)
            self._thr.start()
        except Exception:
            self.proc = None
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> Dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1])); mx.append(float(parts[2])); power.append(float(parts[3]))
            except ValueError:
                continue
            for n, v in zip(names, parts[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s, p in zip(sm, power) if p > 0.5 * max(power)] or sm
        return {"sm_mhz": statistics.median(busy), "sm_max_mhz": max(mx), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


@contextmanager
def nvtx_range(name: str):
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield
