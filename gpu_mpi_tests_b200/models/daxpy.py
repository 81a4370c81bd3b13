"""Distributed daxpy workload (mpi_daxpy_nvtx.cc:85-343 and its smaller siblings).

Phases of the reference, kept as methods so that the CLI can wrap each in the same NVTX range
and host timer: allocate, initialise (x[i]=(i+1)/n, y=-x), daxpy (a=2), local sum,
copy-prep of the own slot, optional barrier, the two allgathers (x in place, y from a separate
send buffer), sum over everything gathered.

impl="fused" (CUDA):  daxpy and the local sum are one kernel; the allgathers are peer-store
                      kernels (parallel/collectives.py); the sums run on the device.
                      `global_sum_step()` is the single-kernel daxpy + all-reduce the north-star
                      metric is quoted on.
impl="baseline":      library calls only -- torch's axpy (cuBLAS-class library kernel),
                      torch.distributed all_gather_into_tensor (aliased = MPI_IN_PLACE), D2H
                      copies and host sums like the reference.  Also the CPU/gloo flavour.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import local
from ..parallel import array as ga
from ..parallel import runtime as _rt
from ..parallel.collectives import PeerCollectives
from ..parallel.symm import PeerSpace, make_space

F64 = torch.float64
N_PER_NODE = 48 * 1024 * 1024  # mpi_daxpy_nvtx.cc:86


def elements_per_rank(nodes: int, world: int, n_per_node: int = N_PER_NODE) -> int:
    """Weak scaling by node: n = nodes * n_per_node / world (mpi_daxpy_nvtx.cc:131-132)."""
    return nodes * n_per_node // world


class DaxpyWorkload:
    def __init__(self, rt: "_rt.Runtime | None" = None, n: int = 1 << 20, impl: str = "fused",
                 mem_space: str = "device", space: PeerSpace | None = None, a: float = 2.0,
                 gather: bool = True, spin_limit: int = 0):
        self.rt = rt or _rt.current()
        self.n, self.impl, self.mem_space, self.a = int(n), impl, mem_space, float(a)
        self.device = self.rt.device if space is None else space.device
        if impl == "fused":
            if self.device.type != "cuda":
                raise RuntimeError("impl='fused' needs CUDA; use impl='baseline' on the CPU/gloo path")
            self.space = space if space is not None else make_space(self.rt)
            self.world, self.ranks = self.space.world, list(self.space.ranks_here)
        elif impl == "baseline":
            self.space = None
            self.world, self.ranks = self.rt.world_size, [self.rt.rank]
        else:
            raise ValueError("impl must be fused or baseline")
        self.gather = gather
        self.x, self.y, self.allx, self.ally, self.sum_out = {}, {}, {}, {}, {}
        self.coll_x = self.coll_y = self.coll_sum = None
        self.spin_limit = spin_limit

    # ------------------------------------------------------------------ phases
    def allocate(self) -> None:
        """allocateArrays (mpi_daxpy_nvtx.cc:177-199)."""
        n, W = self.n, self.world
        for r in self.ranks:
            self.x[r] = ga.empty(n, self.mem_space, self.device)
            self.y[r] = ga.empty(n, self.mem_space, self.device)
            self.sum_out[r] = torch.zeros(1, dtype=F64, device=self.device)
        if self.impl == "fused":
            self.coll_sum = PeerCollectives(self.space, spin_limit=self.spin_limit)
            if self.gather:
                self.coll_x = PeerCollectives(self.space, buf_doubles=n * W, spin_limit=self.spin_limit)
                self.coll_y = PeerCollectives(self.space, buf_doubles=n * W, spin_limit=self.spin_limit)
                for r in self.ranks:
                    self.allx[r], self.ally[r] = self.coll_x.buffer(r), self.coll_y.buffer(r)
        elif self.gather:
            r = self.rt.rank
            self.allx[r] = ga.empty(n * W, self.mem_space, self.device)
            self.ally[r] = ga.empty(n * W, self.mem_space, self.device)

    def initialize(self) -> None:
        """initializeArrays: x[i] = (i+1)/n, y = -x (mpi_daxpy_nvtx.cc:207-225), on the device."""
        n = self.n
        for r in self.ranks:
            local.iota(self.x[r], 1.0, 1.0, 1.0 / n)
            local.iota(self.y[r], 1.0, 1.0, -1.0 / n)

    def daxpy(self) -> None:
        """y <- a x + y.  Fused impl: the local sum comes out of the same pass."""
        for r in self.ranks:
            if self.impl == "fused":
                local.daxpy_sum(self.a, self.x[r], self.y[r], out=self.sum_out[r])
            elif self.device.type == "cuda":
                self.y[r].add_(self.x[r], alpha=self.a)   # library axpy
            else:
                local.daxpy(self.a, self.x[r], self.y[r])

    def local_sum(self, rank: int) -> float:
        """localSum (mpi_daxpy_nvtx.cc:251-268): expect (n+1)/2."""
        if self.impl == "fused":
            return float(self.sum_out[rank].item())
        return float(self.y[rank].cpu().sum()) if self.device.type == "cuda" else float(self.y[rank].sum())

    def prep_allx(self) -> None:
        """copyPrepAllxInplace: D2D copy of x into slot `rank` of allx (mpi_daxpy_nvtx.cc:270-272)."""
        n = self.n
        for r in self.ranks:
            self.allx[r][r * n:(r + 1) * n].copy_(self.x[r], non_blocking=True)

    def barrier(self) -> None:
        if self.impl == "fused":
            self.coll_sum.barrier()
        else:
            self.rt.barrier()

    def gather_x(self, multicast: bool = False) -> None:
        """MPI_Allgather(MPI_IN_PLACE) on allx (mpi_daxpy_nvtx.cc:285)."""
        if self.impl == "fused":
            self.coll_x.allgather_inplace(self.n, multicast=multicast)
        elif self.world > 1:
            r, n = self.rt.rank, self.n
            dist.all_gather_into_tensor(self.allx[r], self.allx[r][r * n:(r + 1) * n])

    def gather_y(self, multicast: bool = False) -> None:
        """MPI_Allgather(d_y -> d_ally) (mpi_daxpy_nvtx.cc:288)."""
        if self.impl == "fused":
            self.coll_y.allgather(self.n, self.y, multicast=multicast)
        else:
            r = self.rt.rank
            if self.world > 1:
                dist.all_gather_into_tensor(self.ally[r], self.y[r])
            else:
                self.ally[r].copy_(self.y[r])

    def all_sum(self, rank: int) -> float:
        """allSum over the gathered y (mpi_daxpy_nvtx.cc:293-310): expect world*(n+1)/2."""
        if self.impl == "fused":
            return float(local.total(self.ally[rank]).item())
        t = self.ally[rank]
        return float(t.cpu().sum()) if t.is_cuda else float(t.sum())

    # ------------------------------------------------------------------ the fused headline op
    def global_sum_step(self) -> None:
        """y <- a x + y and sum over all ranks of sum(y) into sum_out, ONE kernel per rank."""
        if self.impl == "fused":
            self.coll_sum.daxpy_allreduce(self.a, self.x, self.y, self.sum_out)
        else:
            r = self.rt.rank
            if self.device.type == "cuda":
                self.y[r].add_(self.x[r], alpha=self.a)
            else:
                local.daxpy(self.a, self.x[r], self.y[r])
            self.sum_out[r].copy_(self.y[r].sum().reshape(1))
            if self.world > 1:
                dist.all_reduce(self.sum_out[r])

    def sync(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def check(self) -> None:
        for c in (self.coll_x, self.coll_y, self.coll_sum):
            if c is not None:
                c.check_status()

    def free(self) -> None:
        """free (mpi_daxpy_nvtx.cc:313-325)."""
        self.x.clear(); self.y.clear(); self.allx.clear(); self.ally.clear()
        self.coll_x = self.coll_y = None
