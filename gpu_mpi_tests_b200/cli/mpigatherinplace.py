"""mpigatherinplace (mpigatherinplace.f90:1-58): in-place allgather of N doubles per rank
(default 128 Mi = 1 GiB per rank), slot values floor(rank*i/N); prints rank, nmpi, the sum of
the own slot and the sum of the gathered array.  The reference does this in HOST memory with no
timer; here the array lives in peer-mapped device memory (--impl fused) or in a plain tensor
moved by the communication library (--impl baseline / CPU).  Sums are fp64 (the reference's
fp32 accumulators saturate at 2^24).  `--gather ROOT` runs MPI_Gather(MPI_IN_PLACE) instead."""
from __future__ import annotations

import sys

from ..models.gather import GatherWorkload, N_DEFAULT
from ..utils import report
from . import _common


def main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("n", nargs="?", type=int, default=N_DEFAULT, help="doubles per rank")
    ap.add_argument("--gather", type=int, default=None, metavar="ROOT", help="gather to ROOT instead of allgather")
    ap.add_argument("--multicast", action="store_true")
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    wl = GatherWorkload(rt, args.n, impl=impl)
    wl.fill()
    if args.gather is None:
        wl.allgather(multicast=args.multicast)
    else:
        wl.gather_to_root(args.gather)
    lsum, asum = wl.sums(rt.rank)
    wl.check()
    out(report.gather_line(rt.rank, rt.world_size, lsum, asum))
    _common.record(n=args.n, op="allgather" if args.gather is None else f"gather:{args.gather}", lsum=lsum, asum=asum)
    _common.finish(rt)
    return 0


if __name__ == "__main__":
    sys.exit(main())
