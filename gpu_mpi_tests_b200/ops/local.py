"""Single-rank fp64 ops on torch tensors.

CUDA tensors go to the sm_100a kernels in libgmt_b200 (csrc/daxpy.cu, reduce.cu, stencil.cu,
stencil_tma.cu) on the current torch stream; CPU tensors go to the oracle arithmetic
(utils/oracle.py), which is what the gloo flavour of the programs runs on.  There is no
third path: a CUDA tensor without the native library is an error, not a fallback.

2-D arrays follow the reference's column-major convention (mpi_stencil2d_sycl.cc:40-43): the
torch tensor has shape (n_dim1, ld) and element (i, j) is ``t[j, i]``.
"""
from __future__ import annotations

import torch

from ..utils import oracle
from ._lib import lib

F64 = torch.float64
_scratch: dict = {}


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def reduce_scratch(device: torch.device) -> torch.Tensor:
    """Per-device, per-stream zero-initialised scratch for the reduction kernels (the kernels
    leave it zeroed again)."""
    key = (device.index, _stream())
    t = _scratch.get(key)
    if t is None:
        t = torch.zeros(int(lib.gmt_reduce_scratch_bytes()), dtype=torch.uint8, device=device)
        _scratch[key] = t
    return t


def _chk(t: torch.Tensor, name: str) -> None:
    if t.dtype != F64:
        raise TypeError(f"{name}: expected float64, got {t.dtype}")
    if t.dim() == 1 and t.stride(0) != 1:
        raise ValueError(f"{name}: must be contiguous")


def _ld(t: torch.Tensor) -> int:
    """Leading dimension of a column-major 2-D view stored as a (ncols, ld) torch tensor."""
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError("2-D arrays must have unit stride along the last torch dim (dim 0 of the array)")
    return t.stride(0) if t.shape[0] > 1 else max(t.shape[1], t.stride(0))


# ------------------------------------------------------------------ daxpy
def daxpy(a: float, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
    """y <- a*x + y (cublasDaxpy call sites, mpi_daxpy_nvtx.cc:245)."""
    _chk(x, "x"); _chk(y, "y")
    if y.is_cuda:
        lib.gmt_daxpy_f64(y.numel(), float(a), x.data_ptr(), y.data_ptr(), _stream())
        return y
    return oracle.daxpy(a, x, y)


def daxpy_sum(a: float, x: torch.Tensor, y: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """y <- a*x + y and a 1-element tensor holding sum(y_new), in one pass over memory."""
    _chk(x, "x"); _chk(y, "y")
    if y.is_cuda:
        out = out if out is not None else torch.empty(1, dtype=F64, device=y.device)
        lib.gmt_daxpy_sum_f64(y.numel(), float(a), x.data_ptr(), y.data_ptr(), out.data_ptr(),
                              reduce_scratch(y.device).data_ptr(), _stream())
        return out
    oracle.daxpy(a, x, y)
    s = y.sum().reshape(1)
    if out is not None:
        out.copy_(s)
        return out
    return s


def total(x: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """sum(x) as a 1-element tensor (host sum loops mpi_daxpy_nvtx.cc:253-256,296-298)."""
    _chk(x, "x")
    if x.is_cuda:
        out = out if out is not None else torch.empty(1, dtype=F64, device=x.device)
        lib.gmt_sum_f64(x.numel(), x.data_ptr(), out.data_ptr(), reduce_scratch(x.device).data_ptr(), _stream())
        return out
    s = x.sum().reshape(1)
    if out is not None:
        out.copy_(s)
        return out
    return s


def sumsq_diff(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor | None = None) -> torch.Tensor:
    """sum((a-b)^2) as a 1-element tensor; sqrt of it is the reference's err_norm
    (mpi_stencil_gt.cc:222, diff_norm mpi_stencil2d_sycl.cc:165-181).  Contiguous 1-D or
    column-major 2-D views."""
    if a.is_cuda:
        out = out if out is not None else torch.empty(1, dtype=F64, device=a.device)
        scr = reduce_scratch(a.device).data_ptr()
        if a.dim() == 2:
            ny, nx = a.shape
            lib.gmt_sumsq_diff2d_f64(nx, ny, a.data_ptr(), _ld(a), b.data_ptr(), _ld(b), out.data_ptr(), scr,
                                     _stream())
        else:
            _chk(a, "a"); _chk(b, "b")
            lib.gmt_sumsq_diff_f64(a.numel(), a.data_ptr(), b.data_ptr(), out.data_ptr(), scr, _stream())
        return out
    d = (a - b).double()
    s = (d * d).sum().reshape(1)
    if out is not None:
        out.copy_(s)
        return out
    return s


def diff_norm(a: torch.Tensor, b: torch.Tensor) -> float:
    """sqrt(sum((a-b)^2)) as a host float (synchronises)."""
    return float(sumsq_diff(a, b).item()) ** 0.5


def axis_sum(z: torch.Tensor, axis: int, out: torch.Tensor | None = None, n_out: int | None = None) -> torch.Tensor:
    """Axis sum of a column-major (nx, ny) array (gt::sum_axis_to, mpi_stencil2d_gt.cc:611,620).
    axis 0 -> out[j] = sum_i z(i,j); axis 1 -> out[i] = sum_j z(i,j).  Only the first n_out
    entries are produced when n_out is given (the reference sizes the output by the reduced
    axis' own extent)."""
    ny, nx = z.shape
    surviving = ny if axis == 0 else nx
    n_out = surviving if n_out is None else n_out
    if out is None:
        out = torch.empty(n_out, dtype=F64, device=z.device)
    if z.is_cuda:
        lib.gmt_axis_sum_f64(nx, ny, _ld(z), z.data_ptr(), int(axis), n_out, out.data_ptr(), _stream())
        return out
    full = z.sum(dim=1) if axis == 0 else z.sum(dim=0)
    out[:n_out].copy_(full[:n_out])
    return out


def fill(x: torch.Tensor, value: float) -> torch.Tensor:
    """x[:] <- value (fill constructor mpi_stencil2d_gt.cc:598)."""
    if x.is_cuda and x.is_contiguous():
        lib.gmt_fill_f64(x.numel(), float(value), x.data_ptr(), _stream())
        return x
    return x.fill_(value)


def iota(x: torch.Tensor, start: float, step: float, mul: float = 1.0) -> torch.Tensor:
    """x[i] <- (start + i*step) * mul, on the device (host init loops mpi_daxpy_nvtx.cc:209-212)."""
    _chk(x, "x")
    if x.is_cuda:
        lib.gmt_iota_f64(x.numel(), float(start), float(step), float(mul), x.data_ptr(), _stream())
        return x
    i = torch.arange(x.numel(), dtype=F64)
    return x.copy_((start + i * step) * mul)


# ------------------------------------------------------------------ stencils
VARIANTS = {"generic": 0, "vector": 1, "tma": 2}


def stencil1d(y_ghosted: torch.Tensor, scale: float, out: torch.Tensor | None = None,
              variant: str = "vector") -> torch.Tensor:
    """out[i] = scale * sum_k c_k y[i+k] (stencil1d_5, mpi_stencil_gt.cc:54-59)."""
    _chk(y_ghosted, "y")
    n = y_ghosted.numel() - 4
    if out is None:
        out = torch.empty(n, dtype=F64, device=y_ghosted.device)
    if y_ghosted.is_cuda:
        lib.gmt_stencil1d_f64(n, y_ghosted.data_ptr(), out.data_ptr(), float(scale), VARIANTS[variant], _stream())
        return out
    return out.copy_(oracle.stencil1d_5(y_ghosted, scale))


def stencil2d(z: torch.Tensor, dim: int, scale: float, out: torch.Tensor | None = None,
              variant: str = "vector") -> torch.Tensor:
    """5-point derivative along `dim` of a column-major array ghosted in that dim only
    (stencil2d_1d_5_d0/_d1, mpi_stencil2d_gt.cc:84-110)."""
    rows, cols = z.shape  # torch (n_dim1, n_dim0)
    nx = cols - 4 if dim == 0 else cols
    ny = rows if dim == 0 else rows - 4
    if out is None:
        out = torch.empty(ny, nx, dtype=F64, device=z.device)
    if z.is_cuda:
        lib.gmt_stencil2d_f64(int(dim), nx, ny, z.data_ptr(), _ld(z), out.data_ptr(), _ld(out), float(scale),
                              VARIANTS[variant], _stream())
        return out
    return out.copy_(oracle.stencil2d_1d_5(z, dim, scale))


def stencil2d_xy(z: torch.Tensor, sx: float, sy: float, out: torch.Tensor | None = None) -> torch.Tensor:
    """sx*d/dx + sy*d/dy of an array ghosted in both dims (BASELINE config 3)."""
    rows, cols = z.shape
    nx, ny = cols - 4, rows - 4
    if out is None:
        out = torch.empty(ny, nx, dtype=F64, device=z.device)
    if z.is_cuda:
        lib.gmt_stencil2d_xy_f64(nx, ny, z.data_ptr(), _ld(z), out.data_ptr(), _ld(out), float(sx), float(sy),
                                 _stream())
        return out
    return out.copy_(oracle.stencil2d_xy(z, sx, sy))


def copy2d(src: torch.Tensor, dst: torch.Tensor) -> torch.Tensor:
    """dst(i,j) = src(i,j) for equally shaped column-major views: the pack / unpack kernel of
    the baseline exchange (mpi_stencil2d_gt.cc:166,172,239,251; buf_from_view / buf_to_view
    mpi_stencil2d_sycl.cc:82-116)."""
    if src.shape != dst.shape:
        raise ValueError(f"copy2d: shape mismatch {tuple(src.shape)} vs {tuple(dst.shape)}")
    if dst.is_cuda and src.is_cuda and src.dim() == 2 and src.stride(1) == 1 and dst.stride(1) == 1:
        cols, rows = src.shape
        lib.gmt_copy2d_f64(rows, cols, src.data_ptr(), _ld(src), dst.data_ptr(), _ld(dst), _stream())
        return dst
    return dst.copy_(src)


# ------------------------------------------------------------------ initialisers
def init_stencil1d(y: torch.Tensor, actual: torch.Tensor | None, n: int, x0: float, dx: float,
                   left_physical: bool, right_physical: bool) -> None:
    """y = x^3 on the ghosted array, actual = 3x^2 (mpi_stencil_gt.cc:177-196), device side."""
    if y.is_cuda:
        lib.gmt_init_stencil1d_f64(n, 2, float(x0), float(dx), y.data_ptr(),
                                   actual.data_ptr() if actual is not None else 0, int(left_physical),
                                   int(right_physical), _stream())
        return
    i = torch.arange(-2, n + 2, dtype=F64)
    x = x0 + i * dx
    y.copy_(x ** 3)
    if not left_physical:
        y[:2] = 0.0
    if not right_physical:
        y[n + 2:] = 0.0
    if actual is not None:
        actual.copy_(3.0 * x[2:n + 2] ** 2)


def init_stencil2d(z: torch.Tensor, actual: torch.Tensor | None, nx: int, ny: int, gx: int, gy: int,
                   x0: float, y0: float, dx: float, dy: float, deriv_mode: int = 0, sx: float = 1.0,
                   sy: float = 1.0, ghost_mask: int = 0, poison: float = 0.0) -> None:
    """z = x^3 + y^2 on the ghosted array, ghosts analytic only on the sides in ghost_mask
    (1 low x, 2 high x, 4 low y, 8 high y); actual per deriv_mode (1: 3x^2, 2: 2y,
    3: sx*3x^2 + sy*2y)  (mpi_stencil2d_gt.cc:445-497), device side."""
    if z.is_cuda:
        lib.gmt_init_stencil2d_f64(nx, ny, gx, gy, float(x0), float(y0), float(dx), float(dy), z.data_ptr(), _ld(z),
                                   actual.data_ptr() if actual is not None else 0,
                                   _ld(actual) if actual is not None else nx, int(deriv_mode), float(sx),
                                   float(sy), int(ghost_mask), float(poison), _stream())
        return
    xi = x0 + torch.arange(-gx, nx + gx, dtype=F64) * dx
    yj = y0 + torch.arange(-gy, ny + gy, dtype=F64) * dy
    full = xi[None, :] ** 3 + yj[:, None] ** 2
    z.copy_(full)
    if gx:
        if not ghost_mask & 1:
            z[:, :gx] = poison
        if not ghost_mask & 2:
            z[:, gx + nx:] = poison
    if gy:
        if not ghost_mask & 4:
            z[:gy] = poison
        if not ghost_mask & 8:
            z[gy + ny:] = poison
    if gx and gy:
        for rs in (slice(0, gy), slice(gy + ny, None)):
            for cs in (slice(0, gx), slice(gx + nx, None)):
                z[rs, cs] = poison
    if actual is not None and deriv_mode:
        xs, ys = xi[gx:gx + nx], yj[gy:gy + ny]
        ax = (3.0 * xs ** 2)[None, :].expand(ny, nx)
        ay = (2.0 * ys)[:, None].expand(ny, nx)
        actual.copy_({1: ax, 2: ay, 3: sx * ax + sy * ay}[deriv_mode])
