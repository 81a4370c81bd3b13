"""NVTX ranges and the profiler capture window.

The reference instruments its daxpy programs by hand: nvtxRangePushA / nvtxRangePop around
each phase and cudaProfilerStart / cudaProfilerStop around the measured region, so that
`nsys profile -c cudaProfilerApi` (or `ncu --profile-from-start off`) captures only that
window (daxpy_nvtx.cu:65-105, mpi_daxpy_nvtx.cc:167-328, summit/run.sh:15-21).  Same range
names here, through the native library's NVTX entry points when a GPU is present and as a
no-op (with an in-memory trace for tests) otherwise.
"""
from __future__ import annotations

from contextlib import contextmanager

import torch

from ..ops import _lib

_trace: list = []
_depth = 0


def _native() -> bool:
    return torch.cuda.is_available() and _lib.have_native()


def range_push(name: str) -> None:
    global _depth
    _trace.append(("push", _depth, name))
    _depth += 1
    if _native():
        _lib.lib.gmt_nvtx_push(name.encode())


def range_pop() -> None:
    global _depth
    _depth = max(0, _depth - 1)
    _trace.append(("pop", _depth, None))
    if _native():
        _lib.lib.gmt_nvtx_pop()


@contextmanager
def nvtx_range(name: str):
    """`with nvtx_range("cublasDaxpy"):` -- nested ranges are allowed, as in the reference
    (mpiAllGather > x / y, mpi_daxpy_nvtx.cc:283-290)."""
    range_push(name)
    try:
        yield
    finally:
        range_pop()


@contextmanager
def profiler_window():
    """cudaProfilerStart ... cudaProfilerStop (mpi_daxpy_nvtx.cc:167,328)."""
    if _native():
        _lib.lib.gmt_profiler_start()
    _trace.append(("profiler_start", 0, None))
    try:
        yield
    finally:
        if _native():
            _lib.lib.gmt_profiler_stop()
        _trace.append(("profiler_stop", 0, None))


def trace() -> list:
    """Events recorded so far (tests assert on range names and nesting)."""
    return list(_trace)


def reset_trace() -> None:
    global _depth
    _trace.clear()
    _depth = 0
