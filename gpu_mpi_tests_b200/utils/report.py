"""Every line the reference programs print, as format functions (golden strings).

The reference has no logging layer: each program printf()s per rank, unsynchronised
(SURVEY section 5.5).  Keeping the formats in one module lets the CLI programs, the golden
tests and avg.py agree on them.  C printf conversions map 1:1 onto Python %-formatting.

Next to the human-readable lines the framework writes one machine-readable JSON record per
run (JsonlSink) -- the reference has nothing of the kind.
"""
from __future__ import annotations

import json
import os
import sys
import time
from typing import Any, TextIO

# ---------------------------------------------------------------- daxpy family


def daxpy_value(v: float) -> str:
    """daxpy.cu:84 -- one line per element."""
    return "%f" % v


def daxpy_sum(total: float) -> str:
    """daxpy.cu:87."""
    return "SUM = %f" % total


def rank_device(rank: int, n_ranks: int, device: int, n_devices: int, mem: int) -> str:
    """mpi_daxpy.cc:58 -- NB 1-based rank and device numbers."""
    return "RANK[%d/%d] => DEVICE[%d/%d] mem=%d" % (rank + 1, n_ranks, device + 1, n_devices, mem)


def rank_device_error(n_ranks: int, n_devices: int) -> str:
    """mpi_daxpy.cc:45."""
    return "ERROR: Number of ranks (%d) not a multiple of number of GPUs (%d)" % (n_ranks, n_devices)


def memory_per_core(value: str | None) -> str:
    """mpi_daxpy.cc:104,106 (rank 0 only)."""
    return "MEMORY_PER_CORE is not set" if value is None else "MEMORY_PER_CORE=%s" % value


def rank_sum(rank: int, size: int, total: float) -> str:
    """mpi_daxpy.cc:157, mpi_daxpy_nvtx.cc:268."""
    return "%d/%d SUM = %f" % (rank, size, total)


def rank_allsum(rank: int, size: int, total: float) -> str:
    """mpi_daxpy_nvtx.cc:310."""
    return "%d/%d ALLSUM = %f" % (rank, size, total)


def rank_sum_gt(rank: int, size: int, device_id: int, vendor_id: int, total: float) -> str:
    """mpi_daxpy_gt.cc:92."""
    return "%d/%d [%d:0x%08x] SUM = %f" % (rank, size, device_id, vendor_id & 0xFFFFFFFF, total)


def daxpy_nvtx_banner(nodes: int, ranks: int, n: int, total: int) -> str:
    """mpi_daxpy_nvtx.cc:135 (rank 0)."""
    return "%d nodes, %d ranks, %d elements each, total %d" % (nodes, ranks, n, total)


def gpu_memory(free_b: int, total_b: int) -> str:
    """mpi_daxpy_nvtx.cc:203 (rank 0): free / total (used) in MB."""
    mb = 1024.0 * 1024.0
    return "GPU memory %0.3f / %0.3f (%0.3f) MB" % (free_b / mb, total_b / mb, (total_b - free_b) / mb)


def time_line(rank: int, size: int, phase: str, seconds: float) -> str:
    """mpi_daxpy_nvtx.cc:333-340: phase in {'total ', 'kernel', 'barrier', 'gather'} (padded)."""
    label = {"total": "total  ", "kernel": "kernel ", "barrier": "barrier", "gather": "gather "}.get(phase, phase)
    return "%d/%d TIME %s: %0.3f" % (rank, size, label, seconds)


_MEM_KIND = {-3: "NOT CUDA", -2: "UNMANAGED", -4: "INVALID"}


def meminfo(name: str, loc: int) -> str:
    """cuda_error.h:108-134 -- preferred location of a (managed) allocation."""
    if loc == -1:
        where = "CPU (%d)" % -1
    elif loc >= 0:
        where = "DEVICE (%d)" % loc
    elif loc == -4:
        where = "INVALID (%d)" % -2
    else:
        where = _MEM_KIND[loc]
    return "CUDA PreferredLocation of '%s' is %s" % (name, where)


_PTR_KIND = {-1: "NULL", 0: "Unregistered", 1: "Host", 2: "Device", 3: "Managed"}


def ptrinfo(name: str, ptr: int, kind: int) -> str:
    """cuda_error.h:71-96: 'CUDA pointer <label> (<hex>): <type>'."""
    return "CUDA pointer %s (%x): %s" % (name, ptr, _PTR_KIND.get(kind, "Unregistered"))


def cuda_error(file: str, line: int, what: str, code: int, text: str) -> str:
    """cuda_error.h:32-34 (stderr)."""
    return "%s(%i): CUDA Error (%s) %i: %s" % (file, line, what, code, text)


# ---------------------------------------------------------------- 1-D stencil


def stencil1d_banner(world: int, n_global: int, n_local: int) -> list[str]:
    """mpi_stencil_gt.cc:155-157 (rank 0)."""
    return ["n procs  = %d" % world, "n_global = %d" % n_global, "n_local  = %d" % n_local]


def divisor_error(rank: int, world: int, n_global: int) -> str:
    """mpi_stencil_gt.cc:142."""
    return "%d nmpi (%d) must be divisor of domain size (%d), exiting" % (rank, world, n_global)


def exchange_time_s(rank: int, world: int, seconds: float) -> str:
    """mpi_stencil_gt.cc:204 -- seconds."""
    return "%d/%d exchange time %0.8f" % (rank, world, seconds)


def exchange_time_ms(rank: int, world: int, ms: float) -> str:
    """mpi_stencil2d_gt.cc:537 (DEBUG), mpi_stencil2d_sycl.cc:530 -- milliseconds."""
    return "%d/%d exchange time %0.8f ms" % (rank, world, ms)


def err_norm_line(rank: int, world: int, device_id: int, vendor_id: int, err: float) -> str:
    """mpi_stencil_gt.cc:224, mpi_stencil2d_sycl.cc:550."""
    return "%d/%d [%d:0x%08x] err_norm = %.8f" % (rank, world, device_id, vendor_id & 0xFFFFFFFF, err)


# ---------------------------------------------------------------- 2-D gtensor test


def stencil2d_banner(world: int, n_global_deriv: int, n_global_other: int, n_iter: int, n_warmup: int) -> list[str]:
    """mpi_stencil2d_gt.cc:683-687 (rank 0)."""
    return [
        "n procs        = %d" % world,
        "n_global_deriv = %d" % n_global_deriv,
        "n_global_other = %d" % n_global_other,
        "n_iter         = %d" % n_iter,
        "n_warmup       = %d" % n_warmup,
    ]


def test_name(dim: int, space: str, use_buffers: bool) -> str:
    """mpi_stencil2d_gt.cc:379,381 -- note the padded 'device '."""
    label = "device " if space == "device" else "managed"
    return "TEST dim:%d, %s, buf:%d" % (dim, label, int(bool(use_buffers)))


def test_deriv_line(dim: int, space: str, use_buffers: bool, time_sum: float, err_sum: float) -> str:
    """mpi_stencil2d_gt.cc:568-571."""
    return test_name(dim, space, use_buffers) + "; %0.8f, err=%0.8f" % (time_sum, err_sum)


def test_sum_line(dim: int, space: str, time_sum: float) -> str:
    """mpi_stencil2d_gt.cc:645-648 -- always labelled buf:0."""
    return test_name(dim, space, False) + "; allreduce=%0.8f" % time_sum


def allreduce_time_ms(rank: int, world: int, ms: float) -> str:
    """mpi_stencil2d_gt.cc:636 (DEBUG)."""
    return "%d/%d allreduce time %0.8f ms" % (rank, world, ms)


def mpi_error(file: str, line: int, code: int) -> str:
    """mpi_stencil2d_gt.cc:35 (exit 2)."""
    return "%s:%d error %d" % (file, line, code)


# ---------------------------------------------------------------- SYCL variants


def sycl_banner(world: int, rank: int, ny: int, nx_global: int, nx_local: int, n_iter: int, n_warmup: int,
                stage_host: bool) -> list[str]:
    """mpi_stencil2d_sycl.cc:426-435."""
    return [
        "n procs    = %d" % world,
        "rank       = %d" % rank,
        "ny         = %d" % ny,
        "nx_global  = %d" % nx_global,
        "nx_local   = %d" % nx_local,
        "n_iter     = %d" % n_iter,
        "n_warmup   = %d" % n_warmup,
        "stage_host = %d" % int(bool(stage_host)),
    ]


def sycl_device_line(rank: int, bdf: str, uuid_prefix: str) -> str:
    """mpi_stencil2d_sycl.cc:422-424: '<rank> <BDF>(<uuid-hex>)'."""
    return "%d %s(%s)" % (rank, bdf, uuid_prefix)


def dev_bytes(nbytes: float) -> str:
    """mpi_stencil2d_sycl.cc:454-465."""
    gb = nbytes / (1024.0 ** 3)
    if gb >= 1.0:
        return "dev bytes  = %0.3f GB" % gb
    return "dev bytes  = %0.3f MB" % (nbytes / (1024.0 ** 2))


def sycl_oo_banner(world: int, rank: int, n_global: int, n_local: int, n_iter: int, n_warmup: int,
                   stage_host: bool) -> list[str]:
    """mpi_stencil2d_sycl_oo.cc:575-581."""
    return [
        "n procs    = %d" % world,
        "rank       = %d" % rank,
        "n_global   = %d" % n_global,
        "n_local    = %d" % n_local,
        "n_iter     = %d" % n_iter,
        "n_warmup   = %d" % n_warmup,
        "stage_host = %d" % int(bool(stage_host)),
    ]


def oo_exchange_time(rank: int, ms: float) -> str:
    """mpi_stencil2d_sycl_oo.cc:679."""
    return "%d: exchange time %0.8f ms" % (rank, ms)


def oo_err_norm(rank: int, vendor_id: int, err: float) -> str:
    """mpi_stencil2d_sycl_oo.cc:700."""
    return "%d: [0x%08x] err_norm = %.8f" % (rank, vendor_id & 0xFFFFFFFF, err)


# ---------------------------------------------------------------- Fortran programs


def mpienv_line(rank: int, value: int) -> str:
    """mpienv.f90:32, list-directed: 'rank ', rank, ' MEMORY_PER_CORE=', int."""
    return " rank %11d  MEMORY_PER_CORE=%11d" % (rank, value)


def gather_line(rank: int, nmpi: int, lsum: float, asum: float) -> str:
    """mpigatherinplace.f90:48, list-directed: rank, '/', nmpi, ' ', lsum, ' ', asum.
    Sums are computed in fp64 here (the reference accumulates in fp32 and saturates)."""
    return " %11d / %11d   %.8E   %.8E" % (rank, nmpi, lsum, asum)


# ---------------------------------------------------------------- sinks


class Printer:
    """stdout writer that can also capture lines (tests) and flushes per line."""

    def __init__(self, stream: TextIO | None = None, capture: bool = False):
        self.stream = stream if stream is not None else sys.stdout
        self.lines: list[str] | None = [] if capture else None

    def __call__(self, line: str | list[str]) -> None:
        for ln in ([line] if isinstance(line, str) else line):
            if self.lines is not None:
                self.lines.append(ln)
            # one write per line: ranks share the terminal and must not interleave mid-line
            self.stream.write(ln + "\n")
            self.stream.flush()


class JsonlSink:
    """One JSON record per run (GLUP/s, us per exchange, GB/s, max-over-ranks device time,
    fraction of roofline) appended to $GMT_JSONL or the given path; no-op when neither is set."""

    def __init__(self, path: str | None = None):
        self.path = path or os.environ.get("GMT_JSONL")

    def write(self, record: dict[str, Any]) -> None:
        if not self.path:
            return
        rec = {"ts": time.time(), **record}
        with open(self.path, "a") as f:
            f.write(json.dumps(rec) + "\n")
