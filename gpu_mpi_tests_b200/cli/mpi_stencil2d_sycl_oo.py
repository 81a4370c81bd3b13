"""mpi_stencil2d_sycl_oo (mpi_stencil2d_sycl_oo.cc:517-705): the dim-0 case on a SQUARE global
domain with small array classes.  argv: [n_global_Ki=8] [stage_host] [n_iter=100]; n_global =
argv*1024 must divide by the rank count; slab along dim 0; warm-up 5.  --debug shrinks to
n_global/1024 points, 1 iteration, no warm-up and dumps the boundary rows and the received
ghost rows rank by rank on stderr (the DEBUG build, mpi_stencil2d_sycl_oo.cc:545-549,636-659)."""
from __future__ import annotations

import sys

from ..utils import report
from . import _common
from .mpi_stencil2d_sycl import run_dim0


def _main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("n_global_ki", nargs="?", type=int, default=8)
    ap.add_argument("stage_host", nargs="?", default="0")
    ap.add_argument("n_iter", nargs="?", type=int, default=100)
    ap.add_argument("--debug", action="store_true")
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    stage_host = args.stage_host[:1] == "1"
    n_global = args.n_global_ki * 1024
    n_iter, n_warmup = args.n_iter, 5
    if args.debug:
        n_global, n_iter, n_warmup = max(args.n_global_ki, rt.world_size * 4), 1, 0
    slab = _common.slab_or_exit(n_global, rt, out)
    if rt.rank == 0:
        out(report.sycl_oo_banner(rt.world_size, rt.rank, n_global, slab.n_local, n_iter, n_warmup, stage_host))
    if impl == "fused" and slab.n_local % 4:
        out("n_local is not a multiple of 4: using the baseline exchange")
        impl = "baseline"
    run_dim0(rt, out, impl, slab.n_local, n_global, n_iter, n_warmup, stage_host, "oo", debug=args.debug)
    _common.finish(rt)
    return 0


def main(argv=None) -> int:
    return _common.run_main(lambda: _main(argv))


if __name__ == "__main__":
    sys.exit(main())
