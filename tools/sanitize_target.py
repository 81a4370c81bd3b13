#!/usr/bin/env python
"""Small-shape tour of every kernel for compute-sanitizer (memcheck / racecheck, one tool per
run): plain kernels, fused halo kernels and collectives on virtual ranks."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.parallel.collectives import PeerCollectives
from gpu_mpi_tests_b200.parallel.halo import FusedStencil1D, FusedStencil2D
from gpu_mpi_tests_b200.parallel.symm import VirtualSpace

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
F64 = torch.float64
for n in (7, 4096, 9000):
    x = torch.rand(n, dtype=F64, device=dev)
    y = torch.rand(n, dtype=F64, device=dev)
    local.daxpy(2.0, x, y)
    local.daxpy_sum(2.0, x, y)
    local.total(x)
    local.sumsq_diff(x, y)
    yy = torch.rand(n + 4, dtype=F64, device=dev)
    for v in ("generic", "vector", "tma"):
        local.stencil1d(yy, 1.5, variant=v)
for nx, ny in ((8, 5), (1028, 33), (2048, 130)):
    z0 = torch.rand(ny, nx + 4, dtype=F64, device=dev)
    z1 = torch.rand(ny + 4, nx, dtype=F64, device=dev)
    zz = torch.rand(ny + 4, nx + 4, dtype=F64, device=dev)
    local.stencil2d(z0, 0, 1.0)
    local.stencil2d(z1, 1, 1.0)
    local.stencil2d_xy(zz, 1.0, 2.0)
    local.axis_sum(z0, 0)
    local.axis_sum(z0, 1)
    local.copy2d(z0[:, 2:4], torch.empty(ny, 2, dtype=F64, device=dev))
p = FusedStencil1D(VirtualSpace(3, dev), 8192, with_err=True, spin_limit=1 << 24)
p.init_analytic()
for _ in range(3):
    p.step()
for gx, gy, px, py, nx, ny in ((2, 2, 2, 2, 1024, 140), (2, 0, 3, 1, 64, 1100), (0, 2, 1, 3, 2048, 130)):
    q = FusedStencil2D(VirtualSpace(px * py, dev), px, py, nx, ny, gx, gy, with_err=True, spin_limit=1 << 24)
    q.init_analytic()
    for _ in range(3):
        q.step()
    torch.cuda.synchronize()
    q.check_status()
pc = PeerCollectives(VirtualSpace(4, dev), buf_doubles=4 * 1000, stage_cap=4096, spin_limit=1 << 24)
for _ in range(2):
    pc.allgather_inplace(1000)
    pc.gather_inplace(1000, 1)
    v = {r: torch.rand(3000, dtype=F64, device=dev) for r in range(4)}
    pc.allreduce(v, algo="one_shot")
    pc.allreduce(v, algo="two_shot")
    xs = {r: torch.rand(5000, dtype=F64, device=dev) for r in range(4)}
    ys = {r: torch.rand(5000, dtype=F64, device=dev) for r in range(4)}
    outs = {r: torch.zeros(1, dtype=F64, device=dev) for r in range(4)}
    pc.daxpy_allreduce(2.0, xs, ys, outs)
    z = {r: torch.rand(40, 64, dtype=F64, device=dev) for r in range(4)}
    o = {r: torch.zeros(32, dtype=F64, device=dev) for r in range(4)}
    pc.axis_sum_allreduce(z, 0, o, 32)
    pc.axis_sum_allreduce(z, 1, o, 32)
    pc.barrier()
torch.cuda.synchronize()
pc.check_status()
p.check_status()
print("sanitize target done")
