"""mpi_stencil2d_gt (mpi_stencil2d_gt.cc:651-734): the 2-D flagship test.

argv[1] = n_local_deriv (default 1024), argv[2] = n_iter (default 1000); the other dimension is
fixed at 512*1024; warm-up 5 (the banner prints 10, like the reference).  Runs the reference's
matrix: TEST dim:{0,1} x {device, managed} x buf:{1,0} (8 lines: sum over ranks of the seconds
spent exchanging over n_iter iterations, and of err_norm), then 4 all-reduce lines.

--impl baseline   the reference's algorithm on library calls: pack / library send-recv /
                  unpack with the same staging meaning of `buf` (dim 0: host staging on top of
                  device staging; dim 1: device staging vs zero-copy), host-clock timing of the
                  exchange only, stencil untimed.
--impl fused      the printed seconds are n_iter calls of the peer-memory exchange alone
                  (gmt_halo_exchange2d_f64: export, flags, landing; device-timed), the quantity
                  the reference times.  The product's step fuses that exchange into the stencil
                  kernel; its seconds, and what the exchange still costs when hidden under the
                  sweep (fused minus plain sweep), go to the --jsonl record (fused_s, exposed_s).
                  `buf` has no meaning (there are no staging buffers) but all 8 lines are printed.
The all-reduce lines time the in-place SUM of n_local_deriv doubles (8 KiB by default): the
fused path runs axis-sum + all-reduce as one kernel (device-timed), the baseline path times the
library all_reduce alone with the host clock, as the reference does."""
from __future__ import annotations

import sys
import time

import torch
import torch.distributed as dist

from ..models.stencil2d import Stencil2DWorkload
from ..ops import local
from ..parallel import array as ga
from ..parallel.collectives import PeerCollectives
from ..parallel.symm import make_space
from ..utils import report
from . import _common

PI = 3.14159265358979  # the reference's literal differs in the last digits; value is never checked


def test_deriv(rt, out, impl, space, dim, mem, use_buffers, n_local_deriv, n_other, n_iter, n_warmup, debug):
    nx, ny = (n_local_deriv, n_other) if dim == 0 else (n_other, n_local_deriv)
    mode = "dim0" if dim == 0 else "dim1"
    if impl == "baseline":
        stage = ("host" if use_buffers else "device") if dim == 0 else ("device" if use_buffers else "none")
        wl = Stencil2DWorkload(rt, nx, ny, mode=mode, impl="baseline", stage=stage, mem_space=mem)
    else:
        wl = Stencil2DWorkload(rt, nx, ny, mode=mode, impl="fused", with_err=False, space=space, mem_space=mem)
    wl.setup()
    r = rt.rank
    total = 0.0
    extra = {}
    if impl == "baseline":
        for i in range(n_warmup + n_iter):
            t0 = time.perf_counter()
            wl.plan.exchange()
            dt = time.perf_counter() - t0
            if i >= n_warmup:
                total += dt
            wl.plan.stencil()
            if rt.is_cuda:
                torch.cuda.synchronize()
    else:
        m = wl.measure_fused(n_iter, n_warmup)
        total = m["exchange_s"]
        extra = {"fused_s": m["fused_s"], "plain_s": m["plain_s"], "exposed_s": m["exposed_s"]}
    err = wl.err_norm(r)
    wl.check()
    if debug:
        out(report.exchange_time_ms(r, rt.world_size, total / max(n_iter, 1) * 1000))
        out(report.err_norm_line(r, rt.world_size, rt.device_id, rt.vendor_id, err))
    time_sum = rt.reduce_scalar(total)
    err_sum = rt.reduce_scalar(err)
    if r == 0:
        out(report.test_deriv_line(dim, mem, use_buffers, time_sum, err_sum))
    _common.record_append("tests", {"test": "deriv", "dim": dim, "space": mem, "buf": int(use_buffers), "n_iter": n_iter,
                                    "exchange_s": total, "err_norm": err, **extra,
                                    "exchange_s_sum": time_sum if r == 0 else None, "err_sum": err_sum if r == 0 else None})
    del wl
    if rt.is_cuda:
        torch.cuda.empty_cache()


def test_sum(rt, out, impl, space, dim, mem, n_local_deriv, n_other, n_iter, n_warmup, debug):
    nx, ny = (n_local_deriv, n_other) if dim == 0 else (n_other, n_local_deriv)
    r, W = rt.rank, rt.world_size
    z = ga.full((ny, nx), PI / W, mem, rt.device)
    n_out = n_local_deriv  # the reference sizes d_sum by the reduced axis' own extent
    d_sum = ga.empty(n_out, mem, rt.device)
    total = 0.0
    if impl == "baseline":
        for i in range(n_warmup + n_iter):
            local.axis_sum(z, dim, out=d_sum, n_out=n_out)
            if rt.is_cuda:
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            if W > 1:
                dist.all_reduce(d_sum)
            if rt.is_cuda:
                torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            if i >= n_warmup:
                total += dt
    else:
        pc = PeerCollectives(space, stage_cap=max(n_out, 8))
        for _ in range(n_warmup):
            pc.axis_sum_allreduce({r: z}, dim, {r: d_sum}, n_out)
        torch.cuda.synchronize()
        rt.barrier()
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        for _ in range(n_iter):
            pc.axis_sum_allreduce({r: z}, dim, {r: d_sum}, n_out)
        e1.record()
        for _ in range(n_iter):
            local.axis_sum(z, dim, out=d_sum, n_out=n_out)
        e2.record()
        torch.cuda.synchronize()
        pc.check_status()
        total = max(0.0, (e0.elapsed_time(e1) - e1.elapsed_time(e2)) * 1e-3)
    if debug:
        out(report.allreduce_time_ms(r, W, total / max(n_iter, 1) * 1000))
    time_sum = rt.reduce_scalar(total)
    if r == 0:
        out(report.test_sum_line(dim, mem, time_sum))
    _common.record_append("tests", {"test": "sum", "dim": dim, "space": mem, "n_iter": n_iter, "allreduce_s": total,
                                    "allreduce_s_sum": time_sum if r == 0 else None})


def _main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("n_local_deriv", nargs="?", type=int, default=1024)
    ap.add_argument("n_iter", nargs="?", type=int, default=1000)
    ap.add_argument("--n-other", type=int, default=512 * 1024, help="extent of the other dimension")
    ap.add_argument("--no-managed", action="store_true", help="skip the managed-memory cases (TEST_MANAGED=OFF)")
    ap.add_argument("--debug", action="store_true", help="per-rank lines of the DEBUG build")
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    n_warmup_printed, n_warmup = 10, 5
    W = rt.world_size
    if rt.rank == 0:
        out(report.stencil2d_banner(W, args.n_local_deriv * W, args.n_other, args.n_iter, n_warmup_printed))
    space = make_space(rt) if impl == "fused" else None
    spaces = ["device"] if args.no_managed else ["device", "managed"]
    for dim in (0, 1):
        for mem in spaces:
            for buf in (True, False):
                test_deriv(rt, out, impl, space, dim, mem, buf, args.n_local_deriv, args.n_other, args.n_iter,
                           n_warmup, args.debug)
    for dim in (0, 1):
        for mem in spaces:
            test_sum(rt, out, impl, space, dim, mem, args.n_local_deriv, args.n_other, args.n_iter, n_warmup,
                     args.debug)
    _common.finish(rt)
    return 0


def main(argv=None) -> int:
    return _common.run_main(lambda: _main(argv))


if __name__ == "__main__":
    sys.exit(main())
