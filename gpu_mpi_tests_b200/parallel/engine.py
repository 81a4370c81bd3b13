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

    def exchange(self) -> None:
        """The halo exchange alone (no sweep); can be mixed with step()."""
        lib.gmt_engine2d_exchange(self._h)

    def time_exchanges(self, steps: int, warmup: int = 5) -> float:
        """ms per exchange-only call, device time, max over the GPUs."""
        ms = C.c_double()
        lib.gmt_engine2d_time_exchanges(self._h, steps, warmup, C.byref(ms))
        return ms.value

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


class NativeCollectives:
    """Handle on the native collective engine (csrc/engine_coll.cu): `world` ranks placed on the
    GPUs of this process, symmetric buffers wired through CUDA peer access, one launch per GPU
    per collective on the engine's own per-GPU streams."""

    def __init__(self, world: int, buf_doubles: int = 0, stage_cap: int = 4096, devices: list[int] | None = None,
                 spin_limit: int = 0):
        require_native()
        self._h = C.c_void_p()
        dev_arr = (C.c_int * world)(*devices) if devices is not None else None
        lib.gmt_coll_engine_create(world, dev_arr, buf_doubles, stage_cap, spin_limit, C.byref(self._h))
        self.world = int(lib.gmt_coll_engine_world(self._h))
        self.n_device_groups = int(lib.gmt_coll_engine_device_groups(self._h))
        self._keep = {}

    def close(self) -> None:
        if self._h:
            lib.gmt_coll_engine_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):  # noqa: D105
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def rank_device(self, rank: int) -> int:
        d = C.c_int()
        lib.gmt_coll_engine_rank_device(self._h, rank, C.byref(d))
        return d.value

    def buffer(self, rank: int) -> torch.Tensor:
        p, n = C.c_void_p(), C.c_int64()
        lib.gmt_coll_engine_buffer(self._h, rank, C.byref(p), C.byref(n))
        dev = torch.device("cuda", self.rank_device(rank))
        with torch.cuda.device(dev):
            return torch.as_tensor(_DevPtr(p.value, (n.value,), self), device=dev)

    def stream(self, rank: int) -> "torch.cuda.ExternalStream":
        s = C.c_void_p()
        lib.gmt_coll_engine_stream(self._h, rank, C.byref(s))
        return torch.cuda.ExternalStream(s.value, device=torch.device("cuda", self.rank_device(rank)))

    def bind(self, rank: int, inp=None, out=None, y=None) -> None:
        self._keep[rank] = (inp, out, y)   # keep the operands alive while the engine points at them
        ptr = lambda t: None if t is None else C.c_void_p(t.data_ptr())  # noqa: E731
        lib.gmt_coll_engine_bind(self._h, rank, ptr(inp), ptr(out), ptr(y))

    def allgather(self, n: int, from_in: bool = False) -> None:
        lib.gmt_coll_engine_allgather(self._h, n, int(from_in))

    def gather(self, n: int, root: int = 0) -> None:
        lib.gmt_coll_engine_gather(self._h, n, root)

    def allreduce(self, n: int, algo: int = 0) -> None:
        lib.gmt_coll_engine_allreduce(self._h, n, algo)

    def daxpy_allreduce(self, n: int, a: float) -> None:
        lib.gmt_coll_engine_daxpy_allreduce(self._h, n, float(a))

    def axis_sum_allreduce(self, nx: int, ny: int, ld: int, axis: int, n_out: int) -> None:
        lib.gmt_coll_engine_axis_sum_allreduce(self._h, nx, ny, ld, axis, n_out)

    def barrier(self) -> None:
        lib.gmt_coll_engine_barrier(self._h)

    def synchronize(self) -> None:
        lib.gmt_coll_engine_synchronize(self._h)

    def check(self) -> None:
        lib.gmt_coll_engine_status(self._h, None)

    def timed(self, fn, iters: int = 1) -> float:
        """Device milliseconds per call of fn(), max over the GPUs."""
        lib.gmt_coll_engine_mark(self._h, 0)
        for _ in range(iters):
            fn()
        lib.gmt_coll_engine_mark(self._h, 1)
        ms = C.c_double()
        lib.gmt_coll_engine_elapsed_ms(self._h, C.byref(ms))
        return ms.value / iters
