"""Domain decomposition arithmetic.

The reference decomposes one axis into equal slabs, one per rank, with `rank-1` / `rank+1`
as the only neighbours and no wrap-around (mpi_stencil_gt.cc:131-148,
mpi_stencil2d_gt.cc:392-415).  BASELINE config 3 additionally asks for a 2-D process grid with
four face neighbours, which the reference does not have; `Grid2D` provides it (faces only:
the stencils are axis-aligned, so corners are never exchanged).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

from ..utils import report

N_BND = 2


class DecompositionError(ValueError):
    """The domain does not divide evenly (mpi_stencil_gt.cc:141-145)."""


@dataclass(frozen=True)
class Slab1D:
    """Equal slabs of a 1-D axis of `n_global` points over `world` ranks."""
    n_global: int
    world: int
    rank: int

    def __post_init__(self):
        if self.n_global % self.world != 0:
            raise DecompositionError(report.divisor_error(self.rank, self.world, self.n_global))

    @property
    def n_local(self) -> int:
        return self.n_global // self.world

    @property
    def n_local_with_ghost(self) -> int:
        return self.n_local + 2 * N_BND

    @property
    def start(self) -> int:
        return self.rank * self.n_local

    @property
    def lo(self) -> int:
        """Rank of the low neighbour or -1 at the physical boundary."""
        return self.rank - 1 if self.rank > 0 else -1

    @property
    def hi(self) -> int:
        return self.rank + 1 if self.rank < self.world - 1 else -1


def weak_slab(n_local: int, world: int, rank: int) -> Slab1D:
    """Weak scaling: n_global = n_local * world (mpi_stencil2d_gt.cc:395)."""
    return Slab1D(n_local * world, world, rank)


def best_grid(world: int) -> tuple[int, int]:
    """Most square px x py factorisation with px >= py (px splits dim 0, py splits dim 1)."""
    py = int(math.isqrt(world))
    while world % py:
        py -= 1
    return world // py, py


@dataclass(frozen=True)
class Grid2D:
    """px x py process grid; rank = cx + cy*px.  Side order matches the kernels:
    0 = low x, 1 = high x, 2 = low y, 3 = high y.

    periodic=True wraps both directions (an extension: the reference's exchanges never wrap,
    mpi_stencil_gt.cc:94-107), so that EVERY rank has four live faces whatever the grid shape --
    with one rank along a direction the rank is its own neighbour there."""
    px: int
    py: int
    rank: int
    periodic: bool = False

    def __post_init__(self):
        if not (0 <= self.rank < self.px * self.py):
            raise DecompositionError(f"rank {self.rank} outside {self.px}x{self.py} grid")

    @property
    def world(self) -> int:
        return self.px * self.py

    @property
    def cx(self) -> int:
        return self.rank % self.px

    @property
    def cy(self) -> int:
        return self.rank // self.px

    def neighbour(self, side: int) -> int:
        cx, cy = self.cx, self.cy
        if side == 0:
            cx -= 1
        elif side == 1:
            cx += 1
        elif side == 2:
            cy -= 1
        elif side == 3:
            cy += 1
        else:
            raise ValueError("side must be 0..3")
        if self.periodic:
            return (cx % self.px) + (cy % self.py) * self.px
        if 0 <= cx < self.px and 0 <= cy < self.py:
            return cx + cy * self.px
        return -1

    def neighbours(self) -> list[int]:
        return [self.neighbour(s) for s in range(4)]

    def physical_mask(self) -> int:
        """Bit s set when side s is a physical boundary (ghosts filled analytically)."""
        return sum(1 << s for s in range(4) if self.neighbour(s) < 0)


def slab_as_grid(dim: int, world: int, rank: int) -> Grid2D:
    """The reference's 1-D slab decomposition along `dim` expressed as a degenerate grid."""
    return Grid2D(world, 1, rank) if dim == 0 else Grid2D(1, world, rank)
