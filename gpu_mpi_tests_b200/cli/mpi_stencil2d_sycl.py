"""mpi_stencil2d_sycl (mpi_stencil2d_sycl.cc:377-556): the dim-0 case of the 2-D test without
the array library.  argv: [nx_local=1024] [stage_host: first char '1'] [n_iter=100]; warm-up 5,
ny = 512*1024.  Prints a device line per rank, the parameter block, the device-bytes estimate,
per rank `exchange time <ms per iteration>` and `err_norm`."""
from __future__ import annotations

import sys
import time

import torch

from ..models.stencil2d import Stencil2DWorkload
from ..ops import local
from ..parallel.symm import make_space
from ..utils import report
from . import _common


def device_line(rt) -> str:
    if rt.props is not None:
        p = rt.props
        bdf = "%04x:%02x:%02x.0" % (p.pci_domain, p.pci_bus, p.pci_device)
        uuid = "".join("%02x" % b for b in bytes(p.uuid)[:8])
    else:
        bdf, uuid = "cpu", "0" * 16
    return report.sycl_device_line(rt.rank, bdf, uuid)


def _rank_by_rank(rt, fn) -> None:
    """Run fn() on one rank at a time, separated by barriers (the DEBUG build's dump loop,
    mpi_stencil2d_sycl_oo.cc:636-659)."""
    for turn in range(rt.world_size):
        if turn == rt.rank:
            fn()
            sys.stderr.flush()
        rt.barrier()


def debug_dump_boundary_rows(rt, z, n_bnd: int = 2) -> None:
    """DEBUG build: the two boundary rows a rank is about to send on each side, first 20 columns
    (mpi_stencil2d_sycl_oo.cc:641-654).  z is the column-major array as torch [j, i]."""
    nrows, ncols = z.shape[1], z.shape[0]

    def dump():
        for i in list(range(n_bnd, 2 * n_bnd)) + list(range(nrows - 2 * n_bnd, nrows - n_bnd)):
            vals = " ".join("%f" % v for v in z[:min(20, ncols), i].tolist())
            sys.stderr.write("%d: [%d, :] %s\n" % (rt.rank, i, vals))
    _rank_by_rank(rt, dump)


def debug_dump_ghosts(rt, z, has_left: bool, has_right: bool, n_bnd: int = 2, limit: int = 20) -> None:
    """DEBUG build: what arrived in the ghost rows (the reference prints its receive buffers,
    mpi_stencil2d_sycl_oo.cc:481-505); bounded to `limit` columns."""
    nrows, ncols = z.shape[1], z.shape[0]

    def dump():
        for name, has, row0 in (("rbuf_l", has_left, 0), ("rbuf_r", has_right, nrows - n_bnd)):
            if not has:
                continue
            for i in range(min(limit, ncols)):
                for j in range(n_bnd):
                    sys.stderr.write("%d: %s[%d, %d] = %f\n" % (rt.rank, name, j, i, float(z[i, row0 + j])))
    _rank_by_rank(rt, dump)


def run_dim0(rt, out, impl, nx_local, ny, n_iter, n_warmup, stage_host, line_style: str, debug: bool = False):
    space = make_space(rt) if impl == "fused" else None
    r, W = rt.rank, rt.world_size
    if impl == "baseline":
        wl = Stencil2DWorkload(rt, nx_local, ny, mode="dim0", impl="baseline",
                               stage="host" if stage_host else "device")
    else:
        wl = Stencil2DWorkload(rt, nx_local, ny, mode="dim0", impl="fused", with_err=False, space=space)
    wl.setup()
    if debug:
        debug_dump_boundary_rows(rt, wl.z[r])
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
        # the printed time is the peer-memory exchange alone (device-timed), as in the reference
        # (mpi_stencil2d_sycl.cc:515-520); fused-step and exposed-cost seconds go to --jsonl
        m = wl.measure_fused(n_iter, n_warmup)
        total = m["exchange_s"]
        extra = {"fused_s": m["fused_s"], "plain_s": m["plain_s"], "exposed_s": m["exposed_s"]}
    if debug:
        if rt.is_cuda:
            torch.cuda.synchronize()
        debug_dump_ghosts(rt, wl.z[r], r > 0, r < W - 1)
    ms = total / max(n_iter, 1) * 1000
    err = wl.err_norm(r)
    wl.check()
    from . import _common as _c
    _c.record(nx_local=nx_local, ny=ny, n_iter=n_iter, stage_host=bool(stage_host), exchange_ms=ms, err_norm=err, **extra)
    if line_style == "sycl":
        out(report.exchange_time_ms(r, W, ms))
        out(report.err_norm_line(r, W, rt.device_id, rt.vendor_id, err))
    else:
        out(report.oo_exchange_time(r, ms))
        out(report.oo_err_norm(r, rt.vendor_id, err))


def _main(argv=None) -> int:
    ap = _common.base_parser(__doc__)
    ap.add_argument("nx_local", nargs="?", type=int, default=1024)
    ap.add_argument("stage_host", nargs="?", default="0")
    ap.add_argument("n_iter", nargs="?", type=int, default=100)
    ap.add_argument("--ny", type=int, default=512 * 1024)
    args = ap.parse_args(argv)
    rt, out, impl = _common.start(args)
    stage_host = args.stage_host[:1] == "1"
    n_warmup = 5
    W = rt.world_size
    out(device_line(rt))
    if rt.rank == 0:
        out(report.sycl_banner(W, rt.rank, args.ny, args.nx_local * W, args.nx_local, args.n_iter, n_warmup, stage_host))
        nbytes = 8.0 * ((args.nx_local + 4) * args.ny + args.nx_local * args.ny + 4 * 2 * args.ny)
        out(report.dev_bytes(nbytes))
    run_dim0(rt, out, impl, args.nx_local, args.ny, args.n_iter, n_warmup, stage_host, "sycl")
    _common.finish(rt)
    return 0


def main(argv=None) -> int:
    return _common.run_main(lambda: _main(argv))


if __name__ == "__main__":
    sys.exit(main())
