"""CPU oracle: plain-torch/NumPy fp64 restatement of every piece of arithmetic in the framework.

Used (a) as the numerics reference for every CUDA kernel test, (b) as the compute path of the
CPU/gloo flavour of the programs.  Nothing here is performance code.

Reference arithmetic:
  daxpy              y = a*x + y                               daxpy.cu:73, mpi_daxpy_nvtx.cc:245
  init               x[i]=(i+1)/n, y=-x                        mpi_daxpy_nvtx.cc:209-212
                     x[i]=i+1, y[i]=-i-1 (n=1024)              daxpy.cu:55-58
  stencil5           {1/12,-2/3,0,2/3,-1/12} * scale           mpi_stencil_gt.cc:46-59
  1-D test function  y=x^3, dy/dx=3x^2 on [0,8)                mpi_stencil_gt.cc:167-196
  2-D test function  z=x^3+y^2, dz/dx=3x^2, dz/dy=2y           mpi_stencil2d_gt.cc:427-456
  gather slot fill   allx(r*N+i) = floor(r*i/N), i=1..N        mpigatherinplace.f90:34-37
"""
from __future__ import annotations

import math

import torch

STENCIL5 = (1.0 / 12.0, -2.0 / 3.0, 0.0, 2.0 / 3.0, -1.0 / 12.0)
N_BND = 2  # (5 - 1) / 2 ghost points per side
DOMAIN_LENGTH = 8.0  # lx / ln in the reference programs


# ------------------------------------------------------------------ daxpy
def daxpy(a: float, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """In-place y <- a*x + y; returns y."""
    return y.add_(x, alpha=a)


def daxpy_init_small(n: int = 1024, device="cpu"):
    """x[i]=i+1, y[i]=-i-1 (daxpy.cu:55-58)."""
    i = torch.arange(n, dtype=torch.float64, device=device)
    return i + 1.0, -i - 1.0


def daxpy_small_expected_sum(n: int = 1024) -> float:
    """After y = 2x + y with the small init: y[i] = i+1, SUM = n(n+1)/2 (524800 for n=1024)."""
    return n * (n + 1) / 2.0


def daxpy_init_scaled(n: int, device="cpu"):
    """x[i]=(i+1)/n, y=-x (mpi_daxpy_nvtx.cc:209-212)."""
    x = (torch.arange(n, dtype=torch.float64, device=device) + 1.0) / float(n)
    return x, -x.clone()


def daxpy_scaled_expected_sum(n: int) -> float:
    """SUM = (n+1)/2 (mpi_daxpy_nvtx.cc:268)."""
    return (n + 1) / 2.0


def daxpy_scaled_expected_allsum(n: int, world_size: int) -> float:
    """ALLSUM = world_size*(n+1)/2 (mpi_daxpy_nvtx.cc:310)."""
    return world_size * (n + 1) / 2.0


# ------------------------------------------------------------------ stencils
def stencil1d_5(y_ghosted: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    """out[i] = scale * sum_k c_k y[i+k]; input has n+4 points, output n."""
    c = STENCIL5
    y = y_ghosted
    return (c[0] * y[:-4] + c[1] * y[1:-3] + c[2] * y[2:-2] + c[3] * y[3:-1] + c[4] * y[4:]) * scale


def stencil2d_1d_5(z: torch.Tensor, dim: int, scale: float = 1.0) -> torch.Tensor:
    """5-point derivative along `dim` of a COLUMN-MAJOR 2-D array.

    Arrays are stored as torch tensors of shape (n_dim1, n_dim0) so that dim 0 is contiguous:
    element (i, j) of the reference's column-major array is ``t[j, i]``.
    dim 0: z is (nx+4, ny) -> out (nx, ny);  dim 1: z is (nx, ny+4) -> out (nx, ny).
    """
    c = STENCIL5
    if dim == 0:
        return (c[0] * z[:, :-4] + c[1] * z[:, 1:-3] + c[3] * z[:, 3:-1] + c[4] * z[:, 4:]) * scale
    if dim == 1:
        return (c[0] * z[:-4] + c[1] * z[1:-3] + c[3] * z[3:-1] + c[4] * z[4:]) * scale
    raise ValueError("dim must be 0 or 1")


def stencil2d_xy(z: torch.Tensor, sx: float, sy: float) -> torch.Tensor:
    """sx*d/dx + sy*d/dy of a (nx+4, ny+4) array ghosted in both dims (BASELINE config 3)."""
    return stencil2d_1d_5(z[2:-2], 0, sx) + stencil2d_1d_5(z[:, 2:-2], 1, sy)


def fn_x_cubed(x):
    return x * x * x


def fn_x_cubed_deriv(x):
    return 3.0 * x * x


def init_stencil1d(n_local: int, rank: int, world: int, n_global: int | None = None, device="cpu"):
    """(y_ghosted, dydx_actual, scale) for one rank of the 1-D test (mpi_stencil_gt.cc:167-196).

    Ghost cells at a physical boundary get the analytic continuation; ghost cells facing a
    neighbour are left at 0 and must be filled by the exchange.
    """
    n_global = n_global if n_global is not None else n_local * world
    lx = DOMAIN_LENGTH
    dx = lx / n_global
    scale = n_global / lx
    x_start = rank * (lx / world)
    i = torch.arange(n_local, dtype=torch.float64, device=device)
    x = x_start + i * dx
    y = torch.zeros(n_local + 2 * N_BND, dtype=torch.float64, device=device)
    y[N_BND:N_BND + n_local] = fn_x_cubed(x)
    if rank == 0:
        g = (torch.arange(N_BND, dtype=torch.float64, device=device) - N_BND) * dx
        y[:N_BND] = fn_x_cubed(g)
    if rank == world - 1:
        g = lx + torch.arange(N_BND, dtype=torch.float64, device=device) * dx
        y[N_BND + n_local:] = fn_x_cubed(g)
    return y, fn_x_cubed_deriv(x), scale


def init_stencil2d(n_local_deriv: int, n_other: int, dim: int, rank: int, world: int, device="cpu"):
    """(z_ghosted, dz_actual, scale) for one rank of the 2-D slab test
    (mpi_stencil2d_gt.cc:427-497).  Torch layout (n_dim1, n_dim0), see stencil2d_1d_5."""
    n_global_deriv = n_local_deriv * world
    ln = DOMAIN_LENGTH
    delta = ln / n_global_deriv
    scale = n_global_deriv / ln
    start = rank * (ln / world)
    f64 = torch.float64
    if dim == 0:
        nx, ny = n_local_deriv, n_other
        x = start + torch.arange(-N_BND, nx + N_BND, dtype=f64, device=device) * delta
        y = torch.arange(ny, dtype=f64, device=device) * delta
        z = x[None, :] ** 3 + y[:, None] ** 2  # (ny, nx+4)
        if rank != 0:
            z[:, :N_BND] = 0.0
        if rank != world - 1:
            z[:, N_BND + nx:] = 0.0
        actual = (3.0 * x[N_BND:N_BND + nx] ** 2)[None, :].expand(ny, nx).contiguous()
    elif dim == 1:
        nx, ny = n_other, n_local_deriv
        x = torch.arange(nx, dtype=f64, device=device) * delta
        y = start + torch.arange(-N_BND, ny + N_BND, dtype=f64, device=device) * delta
        z = x[None, :] ** 3 + y[:, None] ** 2  # (ny+4, nx)
        if rank != 0:
            z[:N_BND] = 0.0
        if rank != world - 1:
            z[N_BND + ny:] = 0.0
        actual = (2.0 * y[N_BND:N_BND + ny])[:, None].expand(ny, nx).contiguous()
    else:
        raise ValueError("dim must be 0 or 1")
    return z, actual, scale


def err_norm(numeric: torch.Tensor, actual: torch.Tensor) -> float:
    """sqrt(sum((numeric-actual)^2)) (mpi_stencil_gt.cc:222)."""
    d = (numeric - actual).double()
    return math.sqrt(float((d * d).sum()))


def err_norm_pass_bound(points: int) -> float:
    """Round-off-only bound from SURVEY section 2.7: a stale ghost cell gives >~1e5."""
    return 1e-2 * math.sqrt(max(points, 1) / float(1 << 25)) + 1e-6


# ------------------------------------------------------------------ gather
def gather_slot_values(rank: int, n: int, device="cpu") -> torch.Tensor:
    """Slot fill of mpigatherinplace.f90:34-37: value(i) = floor(rank*i/N), i = 1..N."""
    i = torch.arange(1, n + 1, dtype=torch.int64, device=device)
    return torch.div(rank * i, n, rounding_mode="floor").to(torch.float64)


def gather_slot_sum(rank: int, n: int) -> float:
    """Closed form of sum_{i=1..N} floor(r*i/N) = ((r-1)(N-1) + gcd(r,N) - 1)/2 + r."""
    if rank == 0:
        return 0.0
    return ((rank - 1) * (n - 1) + math.gcd(rank, n) - 1) / 2.0 + rank


def gather_all_sum(world: int, n: int) -> float:
    return float(sum(gather_slot_sum(r, n) for r in range(world)))
