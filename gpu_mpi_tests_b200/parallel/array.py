"""Array layer: memory spaces and column-major 2-D arrays.

What the reference gets from gtensor containers (`gt::empty_device`, `gtensor2<T,N,S>` with
S in {device, managed}, mpi_stencil2d_gt.cc:43-73), raw cudaMalloc / cudaMallocManaged /
cudaMallocHost (mpi_daxpy_nvtx.cc:177-199) and the hand-written `span2d` / `array2d` classes
and slice copies of mpi_stencil2d_sycl_oo.cc:51-266.

Storage is always a torch tensor.  "managed" memory is not something torch can allocate, so
it comes from the native library (cudaMallocManaged) and is wrapped through
``__cuda_array_interface__``; it is usable by every kernel and directly from the host, which
is what the reference's managed flavours exercise.

2-D arrays are COLUMN-MAJOR like the reference (first index fastest): an (nrows, ncols)
array is stored as a torch tensor of shape (ncols, nrows); `Array2D` hides the transposition.
"""
from __future__ import annotations

import ctypes as C
import weakref

import numpy as np
import torch

from ..ops import _lib, local

SPACES = ("device", "managed", "host")
_SPACE_ID = {"device": 0, "managed": 1, "host": 2}


class _NativeBlock:
    """Owner of one gmt_alloc allocation, exposed to torch via __cuda_array_interface__."""

    def __init__(self, nbytes: int, space: str):
        self.space = space
        self.nbytes = nbytes
        p = C.c_void_p()
        _lib.lib.gmt_alloc(_SPACE_ID[space], nbytes, C.byref(p))
        self.ptr = int(p.value)
        self.__cuda_array_interface__ = {
            "shape": (max(nbytes // 8, 1),), "typestr": "<f8", "data": (self.ptr, False), "version": 3,
            "strides": None,
        }
        self._finalizer = weakref.finalize(self, _NativeBlock._release, _SPACE_ID[space], self.ptr)

    @staticmethod
    def _release(space_id: int, ptr: int) -> None:
        try:
            _lib.lib.gmt_free(space_id, C.c_void_p(ptr))
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass


_keepalive: "weakref.WeakValueDictionary[int, _NativeBlock]" = weakref.WeakValueDictionary()


def empty(shape, space: str = "device", device: torch.device | None = None) -> torch.Tensor:
    """Uninitialised fp64 array in a memory space.

    device  - torch caching allocator on `device` (CPU tensor when device is the CPU)
    managed - cudaMallocManaged (CPU tensor when there is no GPU: the closest equivalent)
    host    - pinned host memory (plain CPU tensor when there is no GPU)
    """
    if space not in SPACES:
        raise ValueError(f"unknown memory space {space!r}; expected one of {SPACES}")
    shape = (shape,) if isinstance(shape, int) else tuple(shape)
    device = device or (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available()
                        else torch.device("cpu"))
    if device.type != "cuda":
        return torch.empty(shape, dtype=torch.float64)
    if space == "device":
        return torch.empty(shape, dtype=torch.float64, device=device)
    if space == "host":
        return torch.empty(shape, dtype=torch.float64, pin_memory=True)
    n = int(np.prod(shape)) if len(shape) else 1
    blk = _NativeBlock(max(n, 1) * 8, "managed")
    t = torch.as_tensor(blk, device=device)[:n].view(shape)
    _keepalive[t.data_ptr()] = blk
    t._gmt_block = blk  # tensor views made from t keep the storage, the storage keeps blk alive via this ref
    return t


def zeros(shape, space: str = "device", device: torch.device | None = None) -> torch.Tensor:
    return empty(shape, space, device).zero_()


def full(shape, value: float, space: str = "device", device: torch.device | None = None) -> torch.Tensor:
    """gtensor fill constructor (mpi_stencil2d_gt.cc:598)."""
    return local.fill(empty(shape, space, device), value)


def pointer_info(t: torch.Tensor) -> tuple[int, int]:
    """(type, device) as in PTRINFO (cuda_error.h:66-96): 0 unregistered, 1 host, 2 device, 3 managed."""
    if not _lib.have_native() or not torch.cuda.is_available():
        return (0, -1)
    ty, dev = C.c_int(), C.c_int()
    _lib.lib.gmt_ptr_info(C.c_void_p(t.data_ptr()), C.byref(ty), C.byref(dev))
    return ty.value, dev.value


def preferred_location(t: torch.Tensor, nbytes: int | None = None) -> int:
    """MEMINFO (cuda_error.h:99-136): -3 not CUDA, -2 unmanaged, -1 CPU, >= 0 device, -4 invalid."""
    if not _lib.have_native() or not torch.cuda.is_available():
        return -3
    loc = C.c_int()
    nb = nbytes if nbytes is not None else t.numel() * t.element_size()
    _lib.lib.gmt_mem_preferred_location(C.c_void_p(t.data_ptr()), nb, C.byref(loc))
    return loc.value


# ---------------------------------------------------------------------- slices
def normalize_slice(start: int, end: int, extent: int) -> tuple[int, int]:
    """The reference's slice rule (mpi_stencil2d_sycl_oo.cc:174-181): negative indices count
    from the end, and `end == 0` with `start > end` means "to the end"."""
    if start < 0:
        start += extent
    if end < 0:
        end += extent
    elif end == 0 and start > end:
        end = extent
    if not (0 <= start < end <= extent):
        raise IndexError(f"bad slice [{start}:{end}) of extent {extent}")
    return start, end


class Array2D:
    """Column-major (nrows, ncols) fp64 array: `a(row, col)` element access, views, slice
    copies.  The owning/non-owning split of array2d/span2d (mpi_stencil2d_sycl_oo.cc:51-152)
    collapses into torch's storage sharing."""

    def __init__(self, nrows: int, ncols: int, space: str = "device", device: torch.device | None = None,
                 tensor: torch.Tensor | None = None):
        self.nrows, self.ncols, self.space = int(nrows), int(ncols), space
        if tensor is None:
            tensor = empty((self.ncols, self.nrows), space, device)
        if tuple(tensor.shape) != (self.ncols, self.nrows):
            raise ValueError("tensor shape must be (ncols, nrows)")
        self.t = tensor

    @property
    def size(self) -> int:
        return self.nrows * self.ncols

    @property
    def ld(self) -> int:
        return self.t.stride(0) if self.ncols > 1 else self.nrows

    def __call__(self, row: int, col: int) -> float:
        if not (0 <= row < self.nrows and 0 <= col < self.ncols):
            raise IndexError(f"({row}, {col}) outside ({self.nrows}, {self.ncols})")
        return float(self.t[col, row])

    def set(self, row: int, col: int, value: float) -> None:
        if not (0 <= row < self.nrows and 0 <= col < self.ncols):
            raise IndexError(f"({row}, {col}) outside ({self.nrows}, {self.ncols})")
        self.t[col, row] = value

    def view(self, dim: int, start: int, end: int) -> torch.Tensor:
        """Strided torch view of rows (dim 0) or columns (dim 1) [start, end) with the
        reference's index rule."""
        if dim == 0:
            s, e = normalize_slice(start, end, self.nrows)
            return self.t[:, s:e]
        if dim == 1:
            s, e = normalize_slice(start, end, self.ncols)
            return self.t[s:e, :]
        raise ValueError("dim must be 0 or 1")

    def copy_from(self, other: "Array2D") -> None:
        """copy() (mpi_stencil2d_sycl_oo.cc:154-162)."""
        if (other.nrows, other.ncols) != (self.nrows, self.ncols):
            raise ValueError("copy: shape mismatch")
        self.t.copy_(other.t, non_blocking=True)


def copy_src_slice(src: Array2D, dest: Array2D, dim: int, start: int, end: int) -> None:
    """dest <- src[start:end) along dim (pack; mpi_stencil2d_sycl_oo.cc:216-266)."""
    v = src.view(dim, start, end)
    if tuple(v.shape) != tuple(dest.t.shape):
        raise ValueError(f"copy_src_slice: slice {tuple(v.shape)} does not fit dest {tuple(dest.t.shape)}")
    local.copy2d(v, dest.t)


def copy_dest_slice(src: Array2D, dest: Array2D, dim: int, start: int, end: int) -> None:
    """dest[start:end) along dim <- src (unpack; mpi_stencil2d_sycl_oo.cc:164-214)."""
    v = dest.view(dim, start, end)
    if tuple(v.shape) != tuple(src.t.shape):
        raise ValueError(f"copy_dest_slice: src {tuple(src.t.shape)} does not fit slice {tuple(v.shape)}")
    local.copy2d(src.t, v)
