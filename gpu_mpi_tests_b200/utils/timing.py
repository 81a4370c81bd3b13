"""Timing: CUDA events on the launching stream with max-over-ranks, plus the host-clock
phase timers the reference prints.

The reference only has host clocks around fully synchronised phases (MPI_Wtime,
mpi_daxpy_nvtx.cc:168-327; clock_gettime, mpi_stencil2d_gt.cc:512-522).  Device time is the
judged quantity here (every multi-GPU number is the max over ranks of an event-timed
region), the host-clock lines are kept for output parity.
"""
from __future__ import annotations

import subprocess
import threading
import time
from contextlib import contextmanager

import torch
import torch.distributed as dist


class HostTimer:
    """Accumulating wall-clock timer (MPI_Wtime / clock_gettime stand-in)."""

    def __init__(self):
        self.total = 0.0
        self._t0 = None

    def start(self) -> None:
        self._t0 = time.perf_counter()

    def stop(self) -> float:
        dt = time.perf_counter() - self._t0
        self.total += dt
        return dt

    @contextmanager
    def section(self):
        self.start()
        try:
            yield self
        finally:
            self.stop()


def max_over_ranks(value: float, device: torch.device) -> float:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([value], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return float(value)


def _barrier(device: torch.device) -> None:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        if device.type == "cuda":
            dist.barrier(device_ids=[device.index])
        else:
            dist.barrier()


def time_steps(step_fn, steps: int, warmup: int, device: torch.device, tail=None) -> dict:
    """`warmup` untimed calls, then exactly `steps` calls timed on the device, bracketed by a
    barrier + synchronize on both sides; returns total/mean ms as the max over ranks.  `tail`,
    when given, runs inside the timed region after the last call (e.g. the flush of a pipelined
    loop, so that every result has been read back before the clock stops)."""
    for _ in range(warmup):
        step_fn()
    if tail is not None:
        tail()
    if device.type != "cuda":
        _barrier(device)
        t0 = time.perf_counter()
        for _ in range(steps):
            step_fn()
        if tail is not None:
            tail()
        _barrier(device)
        ms = (time.perf_counter() - t0) * 1e3
        ms = max_over_ranks(ms, device)
        return {"total_ms": ms, "ms_per_step": ms / max(steps, 1)}
    torch.cuda.synchronize(device)
    _barrier(device)
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step_fn()
    if tail is not None:
        tail()
    e1.record()
    torch.cuda.synchronize(device)
    ms = e0.elapsed_time(e1)
    _barrier(device)
    torch.cuda.synchronize(device)
    ms = max_over_ranks(ms, device)
    return {"total_ms": ms, "ms_per_step": ms / max(steps, 1)}


def time_graph(step_fn, iters: int, warmup: int, device: torch.device) -> tuple[float, str]:
    """Per-call ms of a launch-bound op: `iters` calls captured in ONE CUDA graph and replayed, so the
    Python/driver launch cost is out of the measurement for the peer-memory kernels and the library
    calls alike (device-timed, max over ranks).  Falls back to eager event timing when the op cannot
    be captured; the second value says which ("graph" / "eager")."""
    try:
        for _ in range(warmup):
            step_fn()
        torch.cuda.synchronize(device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(iters):
                step_fn()
        g.replay()
        torch.cuda.synchronize(device)
        _barrier(device)
        torch.cuda.synchronize(device)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize(device)
        ms = e0.elapsed_time(e1) / iters
        del g
        return max_over_ranks(ms, device), "graph"
    except Exception:  # noqa: BLE001
        torch.cuda.synchronize(device)
        return time_steps(step_fn, iters, warmup, device)["ms_per_step"], "eager"


def time_each(step_fn, iters: int, warmup: int, device: torch.device) -> list:
    """Per-call device times in ms (no max over ranks): for distributions / minima."""
    for _ in range(warmup):
        step_fn()
    torch.cuda.synchronize(device)
    out = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        step_fn()
        b.record()
        b.synchronize()
        out.append(a.elapsed_time(b))
    return out


class ClockSampler:
    """Samples `nvidia-smi` SM clocks and throttle reasons in a background thread while a timed
    region runs (B200_PROFILING.md: clocks line).  Degrades to an empty record when
    nvidia-smi is missing."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0, period_ms: int = 100):
        self.gpu_index, self.period_ms = gpu_index, period_ms
        self.samples: list = []
        self._proc = None
        self._thread = None

    def start(self) -> None:
        try:
            self._proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits",
                 "-lms", str(self.period_ms)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except (FileNotFoundError, OSError):
            self._proc = None
            return

        def pump():
            for line in self._proc.stdout:
                parts = [p.strip() for p in line.split(",")]
                if len(parts) >= 9:
                    self.samples.append(parts)

        self._thread = threading.Thread(target=pump, daemon=True)
        self._thread.start()

    def stop(self) -> dict:
        if self._proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        time.sleep(self.period_ms / 1000.0 * 1.5)
        self._proc.terminate()
        try:
            self._proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self._proc.kill()
        if self._thread is not None:
            self._thread.join(timeout=2)
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            try:
                sm.append(float(s[1]))
                smax.append(float(s[2]))
                power.append(float(s[3]))
            except ValueError:
                continue
            for name, val in zip(names, s[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(smax), "power_w_max": max(power) if power else None,
                "reasons": sorted(reasons), "samples": len(sm)}
