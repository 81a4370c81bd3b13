"""mpi_daxpy_nvtx_{managed,unmanaged} (mpi_daxpy_nvtx.cc:85-343): the daxpy benchmark.

Weak-scaled by node (48 Mi doubles per node), phases wrapped in the reference's NVTX ranges
inside a profiler capture window, then `SUM`, `ALLSUM` and four `TIME` lines per rank.
`--managed` is the -DMANAGED build, `--barrier` the -DBARRIER one.  With --impl fused the
daxpy and its local sum are one kernel and the two allgathers are peer-store kernels; with
--impl baseline everything is library calls (torch axpy, all_gather_into_tensor, host sums)."""
from __future__ import annotations

import sys
import time

from ..models.daxpy import DaxpyWorkload, N_PER_NODE, elements_per_rank
from ..parallel import array as ga
from ..utils import nvtx, report
from ..utils.nvtx import nvtx_range
from . import _common


def main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("--managed", action="store_true", help="cudaMallocManaged arrays (-DMANAGED)")
    ap.add_argument("--barrier", action="store_true", help="time an explicit barrier before the gathers (-DBARRIER)")
    ap.add_argument("--n-per-node", type=int, default=N_PER_NODE)
    ap.add_argument("--multicast", action="store_true", help="allgather through NVLS multimem.st when available")
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    W, r = rt.world_size, rt.rank
    nodes = rt.node_count
    nall = nodes * args.n_per_node
    n = elements_per_rank(nodes, W, args.n_per_node)
    if r == 0:
        out(report.daxpy_nvtx_banner(nodes, W, n, nall))
    _common.getenv_memory_per_core(rt, out)
    out(rt.rank_device_line())

    space_name = "managed" if args.managed else "device"
    k_time = b_time = g_time = 0.0
    with nvtx.profiler_window():
        start_time = time.perf_counter()
        wl = DaxpyWorkload(rt, n, impl=impl, mem_space=space_name)
        with nvtx_range("allocateArrays"):
            wl.allocate()
            wl.sync()
        if r == 0:
            out(report.gpu_memory(*rt.mem_info()))
        with nvtx_range("initializeArrays"):
            with nvtx_range("copyInput"):
                wl.initialize()
                wl.sync()
        out(report.meminfo("d_x", ga.preferred_location(wl.x[r], 8)))
        out(report.meminfo("d_y", ga.preferred_location(wl.y[r], 8)))

        t0 = time.perf_counter()
        with nvtx_range("cublasDaxpy"):
            wl.daxpy()
            wl.sync()
        k_time = time.perf_counter() - t0

        with nvtx_range("localSum"):
            with nvtx_range("copyOutput"):
                s = wl.local_sum(r)
        out(report.rank_sum(r, W, s))

        with nvtx_range("copyPrepAllxInplace"):
            wl.prep_allx()
            wl.sync()
        if args.barrier:
            t0 = time.perf_counter()
            with nvtx_range("mpiBarrier"):
                wl.barrier()
                wl.sync()
            b_time = time.perf_counter() - t0

        t0 = time.perf_counter()
        with nvtx_range("mpiAllGather"):
            with nvtx_range("x"):
                wl.gather_x(multicast=args.multicast)
            with nvtx_range("y"):
                wl.gather_y(multicast=args.multicast)
            wl.sync()
        g_time = time.perf_counter() - t0

        with nvtx_range("allSum"):
            with nvtx_range("copyAlly"):
                total = wl.all_sum(r)
        out(report.rank_allsum(r, W, total))
        wl.check()
        with nvtx_range("free"):
            wl.free()
        total_time = time.perf_counter() - start_time

    out(report.time_line(r, W, "total", total_time))
    out(report.time_line(r, W, "kernel", k_time))
    out(report.time_line(r, W, "barrier", b_time))
    out(report.time_line(r, W, "gather", g_time))
    _common.record(n=n, managed=args.managed, sum=s, allsum=total, total_s=total_time, kernel_s=k_time, barrier_s=b_time,
                   gather_s=g_time, gather_GBs_in=2 * (W - 1) * n * 8 / g_time / 1e9 if g_time > 0 else None)
    _common.finish(rt)
    return 0


if __name__ == "__main__":
    sys.exit(main())
