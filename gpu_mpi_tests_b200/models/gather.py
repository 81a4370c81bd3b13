"""In-place allgather / gather workload (mpigatherinplace.f90 and BASELINE config 5).

Each rank fills its slot of an array of N*world doubles with floor(rank*i/N), i = 1..N
(mpigatherinplace.f90:34-37), all ranks allgather in place (:39-40), and every rank prints the
sum of its own slot and of the whole array (:46-48).  The reference works on host memory and
has no timer; here the array lives in peer-mapped device memory and the collective is a
peer-store kernel (impl="fused") or the communication library's all_gather_into_tensor aliased
in place (impl="baseline", also the CPU/gloo flavour).  `gather_to_root` is the
MPI_Gather(MPI_IN_PLACE) variant BASELINE asks for.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import local
from ..parallel import runtime as _rt
from ..parallel.collectives import PeerCollectives
from ..parallel.symm import PeerSpace, make_space
from ..utils import oracle

F64 = torch.float64
N_DEFAULT = 128 * 1024 * 1024  # mpigatherinplace.f90:11 (1 GiB per rank)


class GatherWorkload:
    def __init__(self, rt: "_rt.Runtime | None" = None, n: int = 1 << 20, impl: str = "fused",
                 space: PeerSpace | None = None, spin_limit: int = 0):
        self.rt = rt or _rt.current()
        self.n, self.impl = int(n), impl
        self.device = self.rt.device if space is None else space.device
        if impl == "fused":
            if self.device.type != "cuda":
                raise RuntimeError("impl='fused' needs CUDA")
            self.space = space if space is not None else make_space(self.rt)
            self.world, self.ranks = self.space.world, list(self.space.ranks_here)
            # slices up to 512 KiB travel as LL cells: give the staging area that many cells per source rank
            self.coll = PeerCollectives(self.space, buf_doubles=self.n * self.world, spin_limit=spin_limit,
                                        stage_cap=min(max(self.n, 8192), 65536))
            self.allx = {r: self.coll.buffer(r) for r in self.ranks}
        else:
            self.space, self.coll = None, None
            self.world, self.ranks = self.rt.world_size, [self.rt.rank]
            self.allx = {self.rt.rank: torch.empty(self.n * self.world, dtype=F64, device=self.device)}

    def slot(self, rank: int, owner: int) -> torch.Tensor:
        return self.allx[rank][owner * self.n:(owner + 1) * self.n]

    def fill(self) -> None:
        for r in self.ranks:
            self.allx[r].fill_(-1.0)
            self.slot(r, r).copy_(oracle.gather_slot_values(r, self.n, self.device))

    def allgather(self, multicast: bool = False) -> None:
        if self.impl == "fused":
            self.coll.allgather_inplace(self.n, multicast=multicast)
        elif self.world > 1:
            r = self.rt.rank
            dist.all_gather_into_tensor(self.allx[r], self.slot(r, r))

    def gather_to_root(self, root: int = 0) -> None:
        if self.impl == "fused":
            self.coll.gather_inplace(self.n, root)
        elif self.world > 1:
            r = self.rt.rank
            if self.rt.backend == "nccl" or self.device.type == "cpu":
                lst = [self.slot(r, p) for p in range(self.world)] if r == root else None
                dist.gather(self.slot(r, r).clone(), lst, dst=root)

    def sums(self, rank: int) -> tuple[float, float]:
        """(lsum, asum): own-slot sum and whole-array sum, in fp64."""
        return float(local.total(self.slot(rank, rank)).item()), float(local.total(self.allx[rank]).item())

    def expected(self, rank: int) -> tuple[float, float]:
        return oracle.gather_slot_sum(rank, self.n), oracle.gather_all_sum(self.world, self.n)

    def bytes_in(self) -> int:
        """Bytes each rank receives per allgather."""
        return (self.world - 1) * self.n * 8

    def check(self) -> None:
        if self.coll is not None:
            self.coll.check_status()
