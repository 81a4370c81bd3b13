#!/usr/bin/env python
"""Short single-GPU program for ncu: a few launches of each hot kernel at the BASELINE shapes.
    python tools/profile_target.py [grid|grid4|dim0|dim0x2|dim1|1d|daxpy|exch|coll|all]

grid4 / dim0x2 run 4 (2x2) / 2 virtual ranks in ONE cooperative launch on this GPU, so the
capture shows the fused kernel with its real exchange work (pushes into neighbours' inboxes,
flag waits, ghost landing) -- a world of one has no neighbours."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.parallel.halo import FusedStencil1D, FusedStencil2D
from gpu_mpi_tests_b200.parallel.symm import VirtualSpace

what = sys.argv[1] if len(sys.argv) > 1 else "all"
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
space = VirtualSpace(1, dev)
import os

N = int(os.environ.get("PROFILE_N", "3"))
if what in ("grid", "all"):
    p = FusedStencil2D(space, 1, 1, 16384, 16384, 2, 2, with_err=True)
    p.init_analytic()
    for _ in range(N):
        p.step()
    torch.cuda.synchronize()
    print("grid err", p.err_norm(0))
    del p
if what in ("grid4", "all"):
    p = FusedStencil2D(VirtualSpace(4, dev), 2, 2, 8192, 8192, 2, 2, with_err=True)
    p.init_analytic(poison=1e30)
    for _ in range(N):
        p.step()
    torch.cuda.synchronize()
    p.check_status()
    print("grid4 err", [p.err_norm(r) for r in range(4)])
    del p
if what in ("dim0x2", "all"):
    p = FusedStencil2D(VirtualSpace(2, dev), 2, 1, 1024, 262144, 2, 0, with_err=False)
    p.init_analytic(poison=1e30)
    for _ in range(N):
        p.step()
    torch.cuda.synchronize()
    p.check_status()
    del p
if what in ("dim0", "all"):
    p = FusedStencil2D(space, 1, 1, 1024, 524288, 2, 0, with_err=False)
    p.init_analytic()
    for _ in range(N):
        p.step()
    torch.cuda.synchronize()
    del p
if what in ("dim1", "all"):
    p = FusedStencil2D(space, 1, 1, 524288, 1024, 0, 2, with_err=False)
    p.init_analytic()
    for _ in range(N):
        p.step()
    for v in ("vector", "tma"):      # plain one-shot kernel and the TMA cluster/multicast kernel
        for _ in range(N):
            local.stencil2d(p.z[0], 1, 1.0, out=p.out[0], variant=v)
    torch.cuda.synchronize()
    del p
if what in ("1d", "all"):
    p = FusedStencil1D(space, 1 << 27, with_err=False)
    p.init_analytic()
    for _ in range(N):
        p.step()
    y = p.z[0]
    o = p.out[0]
    for v in ("vector", "tma"):
        for _ in range(N):
            local.stencil1d(y, 1.0, out=o, variant=v)
    torch.cuda.synchronize()
    del p
if what in ("daxpy", "all"):
    n = 48 << 20
    x = torch.rand(n, dtype=torch.float64, device=dev)
    y = torch.rand(n, dtype=torch.float64, device=dev)
    for _ in range(N):
        local.daxpy(1e-9, x, y)
        local.daxpy_sum(1e-9, x, y)
    torch.cuda.synchronize()
if what in ("exch", "round2"):
    # round-2 kernels: the exchange alone (2 and 4 virtual ranks, real pushes / flags / landing), the dim-0
    # step with ghosts from the inbox, the fused 1-D step on the TMA data path
    p = FusedStencil2D(VirtualSpace(2, dev), 2, 1, 1024, 262144, 2, 0, with_err=False)
    p.init_analytic(poison=1e30)
    for _ in range(N):
        p.exchange()
    torch.cuda.synchronize()
    p.check_status()
    del p
    p = FusedStencil2D(VirtualSpace(4, dev), 2, 2, 8192, 8192, 2, 2, with_err=False)
    p.init_analytic(poison=1e30)
    for _ in range(N):
        p.exchange()
    torch.cuda.synchronize()
    p.check_status()
    del p
    p = FusedStencil2D(VirtualSpace(2, dev), 2, 1, 1024, 262144, 2, 0, with_err=False, land_ghosts=False)
    p.init_analytic(poison=1e30)
    for _ in range(N):
        p.step()
    torch.cuda.synchronize()
    p.check_status()
    del p
    p = FusedStencil1D(VirtualSpace(2, dev), 1 << 26, with_err=False, variant="tma")
    p.init_analytic()
    for _ in range(N):
        p.step()
        p.exchange()
    torch.cuda.synchronize()
    p.check_status()
    del p
if what in ("coll", "round2"):
    from gpu_mpi_tests_b200.parallel.collectives import PeerCollectives
    F64 = torch.float64
    for world, n in ((1, 1 << 20), (1, 48 << 20), (4, 6 << 20)):
        pc = PeerCollectives(VirtualSpace(world, dev))
        xs = {r: torch.rand(n, dtype=F64, device=dev) for r in range(world)}
        ys = {r: torch.rand(n, dtype=F64, device=dev) for r in range(world)}
        outs = {r: torch.zeros(1, dtype=F64, device=dev) for r in range(world)}
        for _ in range(N):
            pc.daxpy_allreduce(1e-9, xs, ys, outs)
        torch.cuda.synchronize()
        pc.check_status()
        del pc, xs, ys
    world = 4
    n_slice = (4 << 20) // 8
    pc = PeerCollectives(VirtualSpace(world, dev), buf_doubles=n_slice * world, stage_cap=65536)
    v = {r: torch.ones(1 << 16, dtype=F64, device=dev) * 1e-9 for r in range(world)}
    z = {r: torch.rand(1024, 4096, dtype=F64, device=dev) for r in range(world)}
    o = {r: torch.zeros(1024, dtype=F64, device=dev) for r in range(world)}
    for _ in range(N):
        pc.allgather_inplace(n_slice, mode="p2p")
        pc.allgather_inplace(8192, mode="ll")
        pc.allreduce({r: v[r][:1024] for r in range(world)}, algo="ll")
        pc.allreduce(v, algo="two_shot")
        pc.axis_sum_allreduce(z, 0, o, 1024)
        pc.barrier()
    torch.cuda.synchronize()
    pc.check_status()
print("profile target done")
