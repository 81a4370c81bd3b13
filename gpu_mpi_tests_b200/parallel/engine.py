"""Python handle on the native exchange engine (csrc/engine.cu).

The engine is the C++ orchestration layer: one process, every GPU of the node, a px x py grid
of ranks placed on the GPUs with the reference's block rule (set_rank_device,
mpi_daxpy.cc:36-62), peers wired through CUDA peer access, one fused halo+stencil launch per
GPU per step.  Native programs use it directly (native/gmt_stencil2d.cpp); this wrapper exists
so tests can drive the same object and look at its arrays as torch tensors.  The torch.distributed
path (one process per GPU, `halo.FusedStencil2D`) launches the same kernels.
"""
from __future__ import annotations

import ctypes as C

import torch

from ..ops._lib import lib, require_native


class _DevPtr:
    """Adapter so torch can view engine-owned device memory (CUDA array interface v2)."""

    def __init__(self, ptr: int, shape: tuple, owner):
        self.__cuda_array_interface__ = {"shape": shape, "typestr": "<f8", "data": (ptr, False), "version": 2,
                                         "strides": None}
        self._owner = owner


class NativeEngine2D:
    def __init__(self, px: int, py: int, nx: int, ny: int, gx: int = 2, gy: int = 2, with_err: bool = True,
                 devices: list[int] | None = None, spin_limit: int = 0):
        require_native()
        self.px, self.py, self.nx, self.ny, self.gx, self.gy = px, py, nx, ny, gx, gy
        self._h = C.c_void_p()
        dev_arr = (C.c_int * (px * py))(*devices) if devices is not None else None
        lib.gmt_engine2d_create(px, py, dev_arr, nx, ny, gx, gy, int(with_err), spin_limit, C.byref(self._h))
        self.world = int(lib.gmt_engine2d_world(self._h))
        self.n_device_groups = int(lib.gmt_engine2d_device_groups(self._h))

    @staticmethod
    def default_device(n_ranks: int, rank: int, n_devices: int) -> int:
        d = C.c_int()
        lib.gmt_engine_default_device(n_ranks, rank, n_devices, C.byref(d))
        return d.value

    def close(self) -> None:
        if self._h:
            lib.gmt_engine2d_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # noqa: D105
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass

    def rank_device(self, rank: int) -> int:
        d = C.c_int()
        lib.gmt_engine2d_rank_device(self._h, rank, C.byref(d))
        return d.value

    def array(self, rank: int, which: str = "out") -> torch.Tensor:
        """Engine-owned array as a torch view: shape (columns, ld), i.e. [j, i] like every other
        column-major array in this package.  Valid until close()."""
        p, cols, ld = C.c_void_p(), C.c_int64(), C.c_int64()
        lib.gmt_engine2d_array(self._h, rank, 0 if which == "z" else 1, C.byref(p), C.byref(cols), C.byref(ld))
        dev = torch.device("cuda", self.rank_device(rank))
        with torch.cuda.device(dev):
            return torch.as_tensor(_DevPtr(p.value, (cols.value, ld.value), self), device=dev)

    def init_analytic(self, poison: float = 0.0) -> None:
        lib.gmt_engine2d_init_analytic(self._h, float(poison))

    def step(self) -> None:
        lib.gmt_engine2d_step(self._h)

    def synchronize(self) -> None:
        lib.gmt_engine2d_synchronize(self._h)

    def check(self) -> None:
        lib.gmt_engine2d_status(self._h, None)

    def time_steps(self, steps: int, warmup: int = 5) -> float:
        ms = C.c_double()
        lib.gmt_engine2d_time_steps(self._h, steps, warmup, C.byref(ms))
        return ms.value

    def err_norm(self, rank: int) -> float:
        v = C.c_double()
        lib.gmt_engine2d_err_norm(self._h, rank, C.byref(v))
        return v.value
