"""Shared plumbing of the CLI programs: runtime start-up, fail-fast exits with the reference's
messages and exit codes, flag parsing that leaves the positional arguments untouched."""
from __future__ import annotations

import argparse
import os
import sys

from ..parallel import decomp, runtime
from ..utils import report


def base_parser(description: str) -> argparse.ArgumentParser:
    ap = argparse.ArgumentParser(description=description)
    ap.add_argument("--impl", choices=["fused", "baseline"], default=None,
                    help="peer-memory kernels or the library-only restatement (default: fused on GPUs)")
    ap.add_argument("--cpu", action="store_true", help="force the CPU/gloo flavour")
    ap.add_argument("--jsonl", default=None, metavar="PATH",
                    help="append one machine-readable JSON record per rank to PATH (the printed lines stay as they are)")
    return ap


# machine-readable twin of the printed lines: programs call record(key=value) next to their
# printf-style output; finish() appends {"program", "rank", "world", "impl", ...} to --jsonl PATH
_state = {"path": None, "program": None, "impl": None, "fields": {}}


def record(**fields) -> None:
    _state["fields"].update(fields)


def record_append(key: str, item) -> None:
    """For programs that print several result lines (one list entry per line)."""
    _state["fields"].setdefault(key, []).append(item)


def start(args) -> tuple:
    """MPI_Init + set_rank_device: returns (rt, out, impl).  A bad rank/device ratio prints the
    reference's message and exits 1 (mpi_daxpy.cc:43-47)."""
    if getattr(args, "cpu", False):
        os.environ["GMT_FORCE_CPU"] = "1"
    out = report.Printer()
    try:
        rt = runtime.init()
    except runtime.RankDeviceError as e:
        out(str(e))
        sys.exit(1)
    impl = args.impl or ("fused" if rt.is_cuda else "baseline")
    if impl == "fused" and not rt.is_cuda:
        out("fused kernels need a CUDA device; using the baseline (library-only) path")
        impl = "baseline"
    _state.update(path=getattr(args, "jsonl", None) or os.environ.get("GMT_JSONL"), program=os.path.basename(sys.argv[0] or "").replace(".py", ""),
                  impl=impl, fields={})
    if _state["program"] in ("", "__main__"):
        _state["program"] = (sys.modules["__main__"].__spec__.name.rsplit(".", 1)[-1]
                             if getattr(sys.modules.get("__main__"), "__spec__", None) else "program")
    return rt, out, impl


def finish(rt) -> None:
    sys.stdout.flush()
    if _state["path"]:
        import json
        rec = {"program": _state["program"], "rank": rt.rank, "world": rt.world_size, "impl": _state["impl"],
               "device": rt.device_id if rt.is_cuda else "cpu", **_state["fields"]}
        line = json.dumps(rec) + "\n"
        fd = os.open(_state["path"], os.O_WRONLY | os.O_CREAT | os.O_APPEND, 0o644)
        try:
            os.write(fd, line.encode())    # one write per record: lines of different ranks never interleave
        finally:
            os.close(fd)
    runtime.finalize()


def slab_or_exit(n_global: int, rt, out) -> decomp.Slab1D:
    """Divisibility check with the reference's message and exit(1) (mpi_stencil_gt.cc:141-145)."""
    try:
        return decomp.Slab1D(n_global, rt.world_size, rt.rank)
    except decomp.DecompositionError as e:
        out(str(e))
        sys.exit(1)


def getenv_memory_per_core(rt, out) -> None:
    """Rank 0 echoes MEMORY_PER_CORE (mpi_daxpy.cc:101-108)."""
    if rt.rank == 0:
        out(report.memory_per_core(os.environ.get("MEMORY_PER_CORE")))


def run_main(body) -> int:
    """Wrap a program body: a failed communication call prints `file:line error code` and exits
    with status 2, like the reference's check() (mpi_stencil2d_gt.cc:32-40)."""
    from ..baseline.exchange import CommError
    try:
        return body()
    except CommError as e:
        where, _, line = e.where.partition(":")
        print(report.mpi_error(where, int(line or 0), e.code), flush=True)
        sys.exit(2)
