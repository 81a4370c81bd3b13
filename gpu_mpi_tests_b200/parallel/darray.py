"""Distributed arrays: the user-facing object that ties decomposition, memory space, ghost
cells and the two exchange implementations together.

The reference has no such object -- every program re-derives n_local, the ghost offsets and
the neighbour ranks by hand (mpi_stencil_gt.cc:131-148, mpi_stencil2d_gt.cc:392-418) and calls
its own copy of boundary_exchange.  `DistArray1D` / `DistArray2D` give the same capabilities as
methods:

    a = DistArray2D(rt, nx, ny, dims=(0,))          # slab along dim 0 like the reference's test
    a.fill_analytic()                                # z = x^3 + y^2, physical ghosts analytic
    a.exchange()                                     # ghost cells <- neighbours (library path)
    d = a.derivative()                               # exchange fused into the stencil kernel

`exchange()` fills the ghost cells and nothing else: on GPUs it is the peer-memory exchange kernel
(gmt_halo_exchange1d/2d_f64: boundary values stored straight into the neighbours' inboxes, flags,
landing -- no pack buffers, no library call), so ANY compute step of the user's can follow it; on
CPUs it is the library (gloo) exchange of baseline/exchange.py.  `derivative()` uses the fused
kernel on GPUs (the exchange happens inside the sweep and also leaves the ghost cells filled) and
exchange()+stencil on CPUs.  `apply(fn)` runs a user function on the exchanged local array.
Arrays follow the reference's column-major convention: `local` has torch shape
(n_dim1 [+4], n_dim0 [+4]).
"""
from __future__ import annotations

import torch

from ..ops import local as ops
from . import decomp
from . import runtime as _rt

F64 = torch.float64
N_BND = 2


class DistArray1D:
    """1-D array of n_global points on [0, 8), equal slabs, 2 ghost points per side."""

    def __init__(self, rt: "_rt.Runtime | None" = None, n_global: int = 1 << 20, impl: str | None = None):
        self.rt = rt or _rt.current()
        self.slab = decomp.Slab1D(n_global, self.rt.world_size, self.rt.rank)
        self.impl = impl or ("fused" if self.rt.is_cuda else "baseline")
        from ..models.stencil1d import Stencil1DWorkload
        self._wl = Stencil1DWorkload(self.rt, self.slab.n_local, impl=self.impl, with_err=False, n_global=n_global)
        self.rank = self.rt.rank

    @property
    def local(self) -> torch.Tensor:
        """The rank's ghosted storage (n_local + 4)."""
        return self._wl.z[self.rank]

    @property
    def interior(self) -> torch.Tensor:
        return self.local[N_BND:-N_BND]

    @property
    def dx(self) -> float:
        return self._wl.dx

    def fill_analytic(self) -> None:
        self._wl.setup()

    def exchange(self) -> None:
        """Fill the ghost cells from the neighbours; the interior and the result are untouched."""
        self._wl.exchange_only()

    def apply(self, fn):
        """exchange(), then fn(local ghosted array) -- the hook for a user's own compute step."""
        self.exchange()
        return fn(self.local)

    def derivative(self) -> torch.Tensor:
        """dy/dx with the 5-point stencil, ghosts refreshed on the way."""
        self._wl.step()
        return self._wl.out[self.rank]

    def err_norm(self) -> float:
        return self._wl.err_norm(self.rank)


class DistArray2D:
    """2-D array with per-rank interior (nx, ny) and ghosts in the decomposed dims.

    dims=(0,)  slab along dim 0 (world x 1 ranks), dims=(1,) slab along dim 1, dims=(0, 1) a
    px x py process grid with four face neighbours."""

    def __init__(self, rt: "_rt.Runtime | None" = None, nx: int = 1024, ny: int = 1024, dims: tuple = (0, 1),
                 impl: str | None = None, mem_space: str = "device", stage: str = "device"):
        self.rt = rt or _rt.current()
        self.impl = impl or ("fused" if self.rt.is_cuda else "baseline")
        mode = {(0,): "dim0", (1,): "dim1", (0, 1): "grid"}[tuple(sorted(dims))]
        from ..models.stencil2d import Stencil2DWorkload
        self._wl = Stencil2DWorkload(self.rt, nx, ny, mode=mode, impl=self.impl, with_err=False, mem_space=mem_space,
                                     stage=stage)
        self.rank = self.rt.rank
        self.nx, self.ny, self.gx, self.gy = nx, ny, self._wl.gx, self._wl.gy

    @property
    def local(self) -> torch.Tensor:
        return self._wl.z[self.rank]

    @property
    def interior(self) -> torch.Tensor:
        z = self.local
        return z[self.gy:self.gy + self.ny, self.gx:self.gx + self.nx]

    def ghost(self, side: int) -> torch.Tensor:
        """Ghost cells of a side (0 low x, 1 high x, 2 low y, 3 high y) as a strided view."""
        z, gx, gy, nx, ny = self.local, self.gx, self.gy, self.nx, self.ny
        return [z[gy:gy + ny, 0:gx], z[gy:gy + ny, gx + nx:], z[0:gy, gx:gx + nx], z[gy + ny:, gx:gx + nx]][side]

    def fill_analytic(self, poison: float = 0.0) -> None:
        self._wl.setup(poison=poison)

    def exchange(self) -> None:
        """Fill the ghost cells of every decomposed side from the neighbours; the interior and the
        result array are untouched."""
        self._wl.exchange()

    def apply(self, fn):
        """exchange(), then fn(local ghosted array) -- the hook for a user's own compute step."""
        self.exchange()
        return fn(self.local)

    def derivative(self) -> torch.Tensor:
        """d/dx (dims has 0) + d/dy (dims has 1), ghosts refreshed on the way."""
        self._wl.step()
        return self._wl.out[self.rank]

    def analytic_derivative(self) -> torch.Tensor:
        return self._wl.analytic(self.rank)

    def err_norm(self) -> float:
        return ops.diff_norm(self._wl.out[self.rank], self._wl.analytic(self.rank))
