"""1-D stencil workload (mpi_stencil_gt.cc:124-230): y = x^3 on [0, 8), slab-decomposed, one
ghost exchange of 2 points per side, 5-point derivative, err_norm against 3x^2.

impl="fused": one kernel does the exchange and the derivative (parallel/halo.py);
impl="baseline": the exchange goes through the communication library directly on the array's
ghost cells, then the un-fused kernel runs (baseline/exchange.py); also the CPU/gloo flavour.
"""
from __future__ import annotations

import torch

from ..ops import local
from ..parallel import runtime as _rt
from ..parallel.halo import FusedStencil1D
from ..parallel.symm import PeerSpace, make_space

F64 = torch.float64


class Stencil1DWorkload:
    def __init__(self, rt: "_rt.Runtime | None" = None, n_local: int = 1 << 22, impl: str = "fused",
                 with_err: bool = True, space: PeerSpace | None = None, n_global: int | None = None,
                 spin_limit: int = 0, variant: str = "vector"):
        self.rt = rt or _rt.current()
        self.n, self.impl, self.with_err = int(n_local), impl, with_err
        self.device = self.rt.device if space is None else space.device
        if impl == "fused":
            if self.device.type != "cuda":
                raise RuntimeError("impl='fused' needs CUDA; use impl='baseline' on the CPU/gloo path")
            self.space = space if space is not None else make_space(self.rt)
            self.plan = FusedStencil1D(self.space, self.n, with_err=with_err, spin_limit=spin_limit, n_global=n_global)
            self.ranks = self.plan.ranks
        elif impl == "baseline":
            from ..baseline.exchange import BaselineStencil1D
            self.plan = BaselineStencil1D(self.rt, self.n, n_global=n_global, variant=variant)
            self.ranks = [self.rt.rank]
        else:
            raise ValueError("impl must be fused or baseline")
        self.z, self.out = self.plan.z, self.plan.out
        self.dx, self.scale, self.n_global = self.plan.dx, self.plan.scale, self.plan.n_global

    def setup(self) -> None:
        self.plan.init_analytic()

    def step(self) -> None:
        self.plan.step()

    def exchange_only(self) -> None:
        """The separately timed boundary_exchange of the reference (mpi_stencil_gt.cc:83-122,200-204):
        library send/recv on the baseline path, one small peer-memory kernel on the fused path."""
        self.plan.exchange()

    def stencil_only(self) -> None:
        if self.impl == "baseline":
            self.plan.stencil()
        else:
            for r in self.ranks:
                local.stencil1d(self.z[r], self.scale, out=self.out[r])

    def analytic(self, rank: int) -> torch.Tensor:
        xs = (rank * self.n + torch.arange(self.n, dtype=F64, device=self.device)) * self.dx
        return 3.0 * xs ** 2

    def err_norm(self, rank: int) -> float:
        if self.impl == "fused" and self.with_err:
            return self.plan.err_norm(rank)
        return local.diff_norm(self.out[rank], self.analytic(rank))

    def check(self) -> None:
        if hasattr(self.plan, "check_status"):
            self.plan.check_status()
