"""mpi_daxpy (mpi_daxpy.cc:65-169): the smoke test per rank -- rank/device report, MEMORY_PER_CORE
echo, device AND managed copies of x and y, preferred-location report for all six pointers,
daxpy on the managed pointers, sum read by the host through unified memory.  No communication."""
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
    _common.getenv_memory_per_core(rt, out)
    out(rt.rank_device_line())
    x, y = oracle.daxpy_init_small(n)
    d_x, d_y = ga.empty(n, "device", rt.device), ga.empty(n, "device", rt.device)
    m_x, m_y = ga.empty(n, "managed", rt.device), ga.empty(n, "managed", rt.device)
    d_x.copy_(x); d_y.copy_(y); m_x.copy_(x); m_y.copy_(y)
    for name, t in (("x", x), ("y", y), ("d_x", d_x), ("d_y", d_y), ("m_x", m_x), ("m_y", m_y)):
        out(report.meminfo(name, ga.preferred_location(t)))
    local.daxpy(a, m_x, m_y)  # on the managed pointers (mpi_daxpy.cc:140-141)
    if rt.is_cuda:
        torch.cuda.synchronize()
    # host reads the managed result (mpi_daxpy.cc:152-156)
    total = float(m_y.cpu().sum()) if m_y.is_cuda else float(m_y.sum())
    out(report.rank_sum(rt.rank, rt.world_size, total))
    _common.finish(rt)
    return 0


if __name__ == "__main__":
    sys.exit(main())
