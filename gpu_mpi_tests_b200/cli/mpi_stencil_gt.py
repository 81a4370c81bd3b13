"""mpi_stencil_gt (mpi_stencil_gt.cc:124-230): 1-D distributed stencil.

argv[1] = n_global in Mi points (default 32), which must divide by the number of ranks.  Prints
the rank-0 banner, per rank `exchange time` (seconds, single un-warmed shot like the reference)
and `[dev:0xid] err_norm`.  With --impl fused the printed time is one call of the peer-memory
exchange alone (gmt_halo_exchange1d_f64: two 16-byte peer stores, flags, landing; device-timed,
un-warmed like the reference's); the derivative then runs as the product does it, with the
exchange fused into the stencil kernel.  With --impl baseline it is the host-clock time of the
library exchange."""
from __future__ import annotations

import sys
import time

import torch

from ..models.stencil1d import Stencil1DWorkload
from ..utils import report
from . import _common


def _main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("n_global_mi", nargs="?", type=int, default=32)
    ap.add_argument("--points", type=int, default=None, help="n_global in points (overrides the Mi argument)")
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    n_global = args.points if args.points is not None else args.n_global_mi * 1024 * 1024
    slab = _common.slab_or_exit(n_global, rt, out)
    if rt.rank == 0:
        out(report.stencil1d_banner(rt.world_size, n_global, slab.n_local))
    r = rt.rank
    wl = Stencil1DWorkload(rt, slab.n_local, impl=impl, with_err=True, n_global=n_global)
    wl.setup()
    if rt.is_cuda:
        torch.cuda.synchronize()
    rt.barrier()
    if impl == "baseline":
        t0 = time.perf_counter()
        wl.exchange_only()
        seconds = time.perf_counter() - t0
        wl.stencil_only()
    else:
        e0, e1 = (torch.cuda.Event(enable_timing=True) for _ in range(2))
        e0.record()
        wl.exchange_only()        # boundary_exchange alone: what the reference times (:200-204)
        e1.record()
        wl.step()                 # exchange + derivative, one kernel (the product's step)
        torch.cuda.synchronize()
        seconds = e0.elapsed_time(e1) * 1e-3
    out(report.exchange_time_s(r, rt.world_size, seconds))
    err = wl.err_norm(r)
    wl.check()
    out(report.err_norm_line(r, rt.world_size, rt.device_id, rt.vendor_id, err))
    _common.record(n_global=n_global, n_local=slab.n_local, exchange_s=seconds, err_norm=err)
    _common.finish(rt)
    return 0


def main(argv=None) -> int:
    return _common.run_main(lambda: _main(argv))


if __name__ == "__main__":
    sys.exit(main())
