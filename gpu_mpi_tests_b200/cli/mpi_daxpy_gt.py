"""mpi_daxpy_gt (mpi_daxpy_gt.cc:48-97): the per-rank smoke test on the portable array API:
host init, copy to device, axpy, sync, copy back, "r/N [dev:0xid] SUM = ..."."""
from __future__ import annotations

import sys

import torch

from ..ops import local
from ..parallel import array as ga
from ..utils import oracle, report
from . import _common


def main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("--n", type=int, default=1024)
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    n, a = args.n, 2.0
    h_x, h_y = oracle.daxpy_init_small(n)
    d_x, d_y = ga.empty(n, "device", rt.device), ga.empty(n, "device", rt.device)
    d_x.copy_(h_x); d_y.copy_(h_y)
    local.daxpy(a, d_x, d_y)
    if rt.is_cuda:
        torch.cuda.synchronize()
    h_y = d_y.cpu()
    out(report.rank_sum_gt(rt.rank, rt.world_size, rt.device_id, rt.vendor_id, float(h_y.sum())))
    _common.finish(rt)
    return 0


if __name__ == "__main__":
    sys.exit(main())
