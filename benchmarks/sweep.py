#!/usr/bin/env python
"""Sweep of the five BASELINE.json configs at the current world size, fused (peer-memory
kernels) against the library-only baseline (NCCL + un-fused kernels), every number timed on
the device with CUDA events and reduced with max over ranks.

    python benchmarks/sweep.py                      # one GPU
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 benchmarks/sweep.py [--quick]

Writes one JSON object per measurement to stdout (rank 0) and, with --out, to a file.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

import torch
import torch.distributed as dist

from gpu_mpi_tests_b200.models.daxpy import DaxpyWorkload
from gpu_mpi_tests_b200.models.stencil1d import Stencil1DWorkload
from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.parallel import runtime
from gpu_mpi_tests_b200.parallel.collectives import PeerCollectives
from gpu_mpi_tests_b200.parallel.symm import make_space
from gpu_mpi_tests_b200.utils import timing

F64 = torch.float64
RESULTS = []


def emit(rt, rec):
    RESULTS.append(rec)
    if rt.rank == 0:
        print(json.dumps(rec), flush=True)


def t_ms(fn, steps, warmup, dev):
    return timing.time_steps(fn, steps, warmup, dev)["ms_per_step"]


def graph_ms(fn, iters, warmup, dev):
    """Latency of a launch-bound op: see utils/timing.time_graph."""
    return timing.time_graph(fn, iters, warmup, dev)


def free():
    torch.cuda.empty_cache()


def stencils(rt, space, args):
    W, dev, r = rt.world_size, rt.device, rt.rank
    steps, warm = args.steps, args.warmup
    # ---- config 2: 1-D, 2^27 points per GPU
    n1 = 1 << (24 if args.quick else 27)
    for impl in ("fused", "baseline"):
        wl = Stencil1DWorkload(rt, n1, impl=impl, with_err=False, space=space if impl == "fused" else None)
        wl.setup()
        ms = t_ms(wl.step, steps, warm, dev)
        rec = {"config": "stencil1d", "impl": impl, "n_gpus": W, "n_per_gpu": n1, "ms_per_step": ms,
               "glups": n1 * W / ms / 1e6, "hbm_gbs_per_gpu": 16.0 * n1 / ms / 1e6}
        if impl == "baseline":
            rec["exchange_us"] = t_ms(wl.exchange_only, steps, warm, dev) * 1e3
            rec["stencil_us"] = t_ms(wl.stencil_only, steps, warm, dev) * 1e3
        else:
            plain = t_ms(wl.stencil_only, steps, warm, dev)
            rec["stencil_only_ms"] = plain
            rec["exposed_exchange_us"] = (ms - plain) * 1e3
        wl.check()
        emit(rt, rec)
        del wl
        free()
    # ---- config 3 (16384^2 grid) and the reference-default slab tiles
    shapes = [("grid", 16384, 16384), ("dim0", 1024, 524288), ("dim1", 524288, 1024)]
    if args.quick:
        shapes = [("grid", 4096, 4096), ("dim0", 1024, 16384), ("dim1", 16384, 1024)]
    for mode, nx, ny in shapes:
        for impl in ("fused", "baseline"):
            wl = Stencil2DWorkload(rt, nx, ny, mode=mode, impl=impl, with_err=False,
                                   space=space if impl == "fused" else None, stage="device")
            wl.setup()
            ms = t_ms(wl.step, steps, warm, dev)
            rec = {"config": f"stencil2d_{mode}", "impl": impl, "n_gpus": W, "tile": [nx, ny], "ms_per_step": ms,
                   "glups": nx * ny * W / ms / 1e6, "hbm_gbs_per_gpu": 16.0 * nx * ny / ms / 1e6,
                   "grid": [wl.px, wl.py]}
            if impl == "baseline":
                rec["exchange_us"] = t_ms(wl.plan.exchange, steps, warm, dev) * 1e3
                rec["stencil_us"] = t_ms(wl.plan.stencil, steps, warm, dev) * 1e3
            else:
                z, o = wl.z[r], wl.out[r]
                if mode == "grid":
                    fn = lambda: local.stencil2d_xy(z, wl.sx, wl.sy, out=o)
                elif mode == "dim0":
                    fn = lambda: local.stencil2d(z, 0, wl.sx, out=o)
                else:
                    fn = lambda: local.stencil2d(z, 1, wl.sy, out=o)
                plain = t_ms(fn, steps, warm, dev)
                rec["stencil_only_ms"] = plain
                rec["exposed_exchange_us"] = (ms - plain) * 1e3
                rec["halo_bytes_out_per_gpu"] = wl.plan.halo_bytes_out(r)
            wl.check()
            emit(rt, rec)
            del wl
            free()


def exchanges(rt, space, args):
    """The halo exchange ALONE -- the reference's timed quantity (mpi_stencil_gt.cc:200-204,
    mpi_stencil2d_gt.cc:512-522) -- peer-memory kernel vs the library restatement (pack kernels, NCCL
    send/recv, unpack kernels, device-wide syncs), for the message sizes the reference programs use:
    16 B (1-D), 128 KiB strided (mpi_stencil2d_sycl_oo default), 1 MiB strided, 8 MiB strided (dim 0),
    8 MiB contiguous (dim 1), and the flagship's four 256 KiB faces."""
    W, dev, r = rt.world_size, rt.device, rt.rank
    cases = [("1d_16B", "1d", 1 << 22, 0), ("dim0_128KiB_strided", "dim0", 1024, 8192),
             ("dim0_1MiB_strided", "dim0", 1024, 65536), ("dim0_8MiB_strided", "dim0", 1024, 524288),
             ("dim1_8MiB_contiguous", "dim1", 524288, 1024), ("grid_4x256KiB", "grid", 16384, 16384)]
    if args.quick:
        cases = cases[:3]
    for label, mode, nx, ny in cases:
        for impl in ("fused", "baseline"):
            if mode == "1d":
                wl = Stencil1DWorkload(rt, nx, impl=impl, with_err=False, space=space if impl == "fused" else None)
                fn = wl.exchange_only
            else:
                wl = Stencil2DWorkload(rt, nx, ny, mode=mode, impl=impl, with_err=False,
                                       space=space if impl == "fused" else None, stage="device")
                fn = wl.exchange
            wl.setup()
            if impl == "fused":
                ms, how = graph_ms(fn, 50, 5, dev)
                eager = t_ms(fn, 50, 5, dev)
                rec = {"us": ms * 1e3, "timing": how, "us_eager_python_call": eager * 1e3}
                if mode != "1d":
                    rec["halo_bytes_out_per_gpu"] = wl.plan.halo_bytes_out(r)
            else:
                rec = {"us": t_ms(fn, 30, 5, dev) * 1e3, "timing": "eager (device-wide syncs inside every call)"}
            wl.check()
            emit(rt, {"config": "halo_exchange_alone", "case": label, "impl": impl, "n_gpus": W, **rec})
            del wl
            free()


def daxpy(rt, space, args):
    W, dev, r = rt.world_size, rt.device, rt.rank
    sizes = [1 << 20, (48 << 20) // W]
    for n in sizes:
        iters = 200 if n <= (1 << 20) else args.steps
        for impl in ("fused", "baseline"):
            wl = DaxpyWorkload(rt, n, impl=impl, space=space if impl == "fused" else None, gather=False)
            wl.allocate()
            wl.initialize()
            wl.a = 1e-9  # keep y bounded over many repetitions
            if n <= (1 << 20):
                ms, how = graph_ms(wl.global_sum_step, 50, 10, dev)
            else:
                ms, how = t_ms(wl.global_sum_step, iters, 10, dev), "eager"
            wl.check()
            emit(rt, {"config": "daxpy_global_sum", "impl": impl, "n_gpus": W, "n_per_gpu": n, "us_per_step": ms * 1e3,
                      "timing": how,
                      "gbs_per_gpu": 24.0 * n / ms / 1e6, "gbs_total": 24.0 * n * W / ms / 1e6})
            del wl
            free()


def collectives(rt, space, args):
    W, dev, r = rt.world_size, rt.device, rt.rank
    if W == 1:
        return
    # ---- config 5: in-place allgather / gather sweep, slice sizes 1 KB .. 1 GB (capped by --max-slice)
    sizes = [1 << k for k in range(10, 31, 2)]
    sizes = [s for s in sizes if s <= args.max_slice]
    nmax = max(sizes) // 8
    pc = PeerCollectives(space, buf_doubles=nmax * W, stage_cap=65536)   # LL cells for slices up to 512 KiB
    ref = torch.empty(nmax * W, dtype=F64, device=dev)
    for sbytes in sizes:
        n = sbytes // 8
        iters = 200 if sbytes <= (1 << 16) else (50 if sbytes <= (1 << 22) else 10)
        variants = [("fused_p2p", lambda: pc.allgather_inplace(n, mode="p2p"))]
        if n <= pc.ll_gather_max:
            variants.append(("fused_ll", lambda: pc.allgather_inplace(n, mode="ll")))
        if pc.has_multicast:
            variants.append(("fused_multicast", lambda: pc.allgather_inplace(n, mode="multimem")))
            variants.append(("fused_multicast_weak", lambda: pc.allgather_inplace(n, mode="multimem_weak")))
            variants.append(("fused_mc_plain_store", lambda: pc.allgather_inplace(n, mode="mc_store")))
        variants.append(("fused_gather_root0", lambda: pc.gather_inplace(n, 0)))
        variants.append(("nccl_allgather", lambda: dist.all_gather_into_tensor(ref[:n * W], ref[r * n:(r + 1) * n])))
        for name, fn in variants:
            if sbytes <= (1 << 20):
                ms, how = graph_ms(fn, 50, 5, dev)
            else:
                ms, how = t_ms(fn, iters, 5, dev), "eager"
            emit(rt, {"config": "allgather_inplace", "impl": name, "n_gpus": W, "slice_bytes": sbytes, "timing": how,
                      "us": ms * 1e3, "ingress_gbs_per_gpu": (W - 1) * sbytes / ms / 1e6})
    pc.check_status()
    del pc, ref
    free()
    # ---- in-place all-reduce, 8 KiB (test_sum) and a few larger vectors
    for n in (1024, 1 << 14, 1 << 18):
        pc = PeerCollectives(space, stage_cap=n)
        v = torch.ones(n, dtype=F64, device=dev) * 1e-12
        algos = ["ll", "one_shot", "two_shot"] + (["multimem"] if pc.has_multicast else [])
        for algo in algos:
            ms, how = graph_ms(lambda: pc.allreduce({r: v}, algo=algo), 50, 10, dev)
            emit(rt, {"config": "allreduce_inplace", "impl": f"fused_{algo}", "n_gpus": W, "bytes": n * 8, "us": ms * 1e3,
                      "timing": how})
        ms, how = graph_ms(lambda: dist.all_reduce(v), 50, 10, dev)
        emit(rt, {"config": "allreduce_inplace", "impl": "nccl", "n_gpus": W, "bytes": n * 8, "us": ms * 1e3, "timing": how})
        pc.check_status()
        del pc
    pc = PeerCollectives(space)
    ms, how = graph_ms(pc.barrier, 50, 10, dev)
    emit(rt, {"config": "barrier", "impl": "fused", "n_gpus": W, "us": ms * 1e3, "timing": how})
    tb = torch.zeros(1, device=dev)
    ms, how = graph_ms(lambda: dist.all_reduce(tb), 50, 10, dev)
    emit(rt, {"config": "barrier", "impl": "nccl_allreduce_4B", "n_gpus": W, "us": ms * 1e3, "timing": how})
    free()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--max-slice", type=int, default=1 << 30, help="largest allgather slice in bytes (BASELINE config 5: 1 GiB)")
    ap.add_argument("--only", default="stencils,exchanges,daxpy,collectives")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rt = runtime.init()
    if not rt.is_cuda:
        raise SystemExit("needs CUDA")
    space = make_space(rt)
    emit(rt, {"config": "env", "n_gpus": rt.world_size, "peer_space": space.kind, "multicast": space.has_multicast,
              "gpu": torch.cuda.get_device_name(rt.device)})
    only = args.only.split(",")
    if "stencils" in only:
        stencils(rt, space, args)
    if "exchanges" in only:
        exchanges(rt, space, args)
    if "daxpy" in only:
        daxpy(rt, space, args)
    if "collectives" in only:
        collectives(rt, space, args)
    if args.out and rt.rank == 0:
        Path(args.out).write_text("\n".join(json.dumps(x) for x in RESULTS) + "\n")
    runtime.finalize()


if __name__ == "__main__":
    main()
