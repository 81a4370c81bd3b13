"""daxpy / daxpy_nvtx (daxpy.cu:35-94, daxpy_nvtx.cu:38-108): single-GPU smoke test.
y = 2x + y with x[i] = i+1, y[i] = -i-1, n = 1024; prints every y[i] and SUM = 524800.000000."""
from __future__ import annotations

import sys

import torch

from ..ops import local
from ..parallel import array as ga
from ..utils import nvtx, oracle, report
from . import _common


def run(argv=None, with_nvtx: bool = False) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("--n", type=int, default=1024)
    ap.add_argument("--quiet", action="store_true", help="do not print the n element lines")
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    n, a = args.n, 2.0
    win = nvtx.profiler_window() if with_nvtx else None
    if win:
        win.__enter__()
    # host arrays + H2D like the reference (daxpy.cu:55-70); the kernel itself runs on the device
    hx, hy = oracle.daxpy_init_small(n)
    if with_nvtx:
        nvtx.range_push("copyInput")
    dx = ga.empty(n, "device", rt.device)
    dy = ga.empty(n, "device", rt.device)
    dx.copy_(hx)
    dy.copy_(hy)
    if with_nvtx:
        nvtx.range_pop()
        nvtx.range_push("cublasDaxpy")
    if impl == "fused" or not rt.is_cuda:
        local.daxpy(a, dx, dy)
    else:
        dy.add_(dx, alpha=a)
    if rt.is_cuda:
        torch.cuda.synchronize()
    if with_nvtx:
        nvtx.range_pop()
        nvtx.range_push("copyOutput")
    hy = dy.cpu()
    if with_nvtx:
        nvtx.range_pop()
    total = 0.0
    for v in hy.tolist():
        if not args.quiet:
            out(report.daxpy_value(v))
        total += v
    out(report.daxpy_sum(total))
    if win:
        win.__exit__(None, None, None)
    _common.finish(rt)
    return 0


def main(argv=None) -> int:
    return run(argv, with_nvtx=False)


if __name__ == "__main__":
    sys.exit(main())
