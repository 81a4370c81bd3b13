"""ctypes binding of libgmt_b200.so (C ABI in include/gmt_b200.h).

The library is built in-tree by ``gpu_mpi_tests_b200._native.build`` and loaded from
``gpu_mpi_tests_b200/_native/``.  On a machine with a GPU a missing library is a hard error
(no silent PyTorch fallback): the CUDA kernels *are* the product.  On a CPU-only machine the
package still imports and the CPU/gloo path (``gpu_mpi_tests_b200.utils.oracle`` arithmetic on
torch CPU tensors) is used instead.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

_NATIVE_DIR = Path(__file__).resolve().parents[1] / "_native"
LIB_PATH = _NATIVE_DIR / "libgmt_b200.so"

c_i64 = C.c_int64
c_dbl = C.c_double
c_ptr = C.c_void_p


class GmtError(RuntimeError):
    """A libgmt_b200 entry point returned non-zero (cf. CHECK in cuda_error.h:16-41)."""

    def __init__(self, name: str, code: int, message: str):
        super().__init__(f"{name} failed with code {code}: {message}")
        self.code = code
        self.entry_point = name


class DeviceProps(C.Structure):
    _fields_ = [
        ("name", C.c_char * 256),
        ("total_global_mem", C.c_uint64),
        ("pci_domain", C.c_int), ("pci_bus", C.c_int), ("pci_device", C.c_int),
        ("sm_count", C.c_int), ("cc_major", C.c_int), ("cc_minor", C.c_int),
        ("clock_khz", C.c_int), ("mem_clock_khz", C.c_int), ("l2_bytes", C.c_int),
        ("managed_memory", C.c_int), ("concurrent_managed_access", C.c_int),
        ("max_smem_optin", C.c_int),
        ("uuid", C.c_ubyte * 16),
    ]


class Halo1dRank(C.Structure):
    _fields_ = [
        ("z", c_ptr), ("out", c_ptr), ("n", c_i64), ("scale", c_dbl),
        ("has", C.c_int32 * 2),
        ("my_inbox", (c_ptr * 2) * 2), ("peer_inbox", (c_ptr * 2) * 2),
        ("my_flag", c_ptr * 2), ("peer_flag", c_ptr * 2),
        ("epoch", c_ptr), ("done", c_ptr), ("status", c_ptr),
        ("err_scratch", c_ptr), ("err_sumsq", c_ptr),
        ("x0", c_dbl), ("dx", c_dbl),
        ("spin_limit", c_i64),
    ]


class Halo2dRank(C.Structure):
    _fields_ = [
        ("z", c_ptr), ("out", c_ptr),
        ("nx", c_i64), ("ny", c_i64), ("ldz", c_i64), ("ldo", c_i64),
        ("gx", C.c_int32), ("gy", C.c_int32),
        ("sx", c_dbl), ("sy", c_dbl),
        ("has", C.c_int32 * 4),
        ("my_inbox", (c_ptr * 2) * 4), ("peer_inbox", (c_ptr * 2) * 4),
        ("my_flag", c_ptr * 4), ("peer_flag", c_ptr * 4),
        ("epoch", c_ptr), ("done", c_ptr), ("status", c_ptr),
        ("err_scratch", c_ptr), ("err_sumsq", c_ptr),
        ("x0", c_dbl), ("y0", c_dbl), ("dx", c_dbl), ("dy", c_dbl),
        ("spin_limit", c_i64),
    ]


class CollRank(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32),
        ("peer_buf", c_ptr), ("mc_buf", c_ptr),
        ("peer_stage", c_ptr), ("mc_stage", c_ptr), ("stage_cap", c_i64),
        ("peer_flag", c_ptr),
        ("epoch", c_ptr), ("done", c_ptr), ("status", c_ptr), ("scratch", c_ptr),
        ("inp", c_ptr), ("out", c_ptr), ("y", c_ptr),   # `in` in the C struct
        ("spin_limit", c_i64),
    ]


_SIGS = {
    # name: (restype, argtypes)   -- restype None means "int status, raise on non-zero"
    "gmt_version": (C.c_int, []),
    "gmt_device_count": (None, [C.POINTER(C.c_int)]),
    "gmt_set_device": (None, [C.c_int]),
    "gmt_get_device": (None, [C.POINTER(C.c_int)]),
    "gmt_device_props_get": (None, [C.c_int, C.POINTER(DeviceProps)]),
    "gmt_mem_info": (None, [C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]),
    "gmt_ptr_info": (None, [c_ptr, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "gmt_mem_preferred_location": (None, [c_ptr, C.c_size_t, C.POINTER(C.c_int)]),
    "gmt_can_access_peer": (None, [C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "gmt_enable_peer_access": (None, [C.c_int]),
    "gmt_device_synchronize": (None, []),
    "gmt_stream_synchronize": (None, [c_ptr]),
    "gmt_alloc": (None, [C.c_int, C.c_size_t, C.POINTER(c_ptr)]),
    "gmt_free": (None, [C.c_int, c_ptr]),
    "gmt_memcpy_async": (None, [c_ptr, c_ptr, C.c_size_t, C.c_int, c_ptr]),
    "gmt_memset_async": (None, [c_ptr, C.c_int, C.c_size_t, c_ptr]),
    "gmt_ipc_get_handle": (None, [c_ptr, C.POINTER(C.c_ubyte * 64)]),
    "gmt_ipc_open_handle": (None, [C.POINTER(C.c_ubyte * 64), C.POINTER(c_ptr)]),
    "gmt_ipc_close_handle": (None, [c_ptr]),
    "gmt_nvtx_push": (None, [C.c_char_p]),
    "gmt_nvtx_pop": (None, []),
    "gmt_profiler_start": (None, []),
    "gmt_profiler_stop": (None, []),
    "gmt_reduce_scratch_bytes": (C.c_size_t, []),
    "gmt_daxpy_f64": (None, [c_i64, c_dbl, c_ptr, c_ptr, c_ptr]),
    "gmt_daxpy_sum_f64": (None, [c_i64, c_dbl, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gmt_daxpy_sum_noalloc_f64": (None, [c_i64, c_dbl, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gmt_noalloc_max_n": (c_i64, []),
    "gmt_noalloc_scratch_bytes": (C.c_size_t, []),
    "gmt_sum_f64": (None, [c_i64, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gmt_sumsq_diff_f64": (None, [c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "gmt_sumsq_diff2d_f64": (None, [c_i64, c_i64, c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr, c_ptr]),
    "gmt_axis_sum_f64": (None, [c_i64, c_i64, c_i64, c_ptr, C.c_int, c_i64, c_ptr, c_ptr]),
    "gmt_fill_f64": (None, [c_i64, c_dbl, c_ptr, c_ptr]),
    "gmt_fill_u64": (None, [c_i64, C.c_uint64, c_ptr, c_ptr]),
    "gmt_iota_f64": (None, [c_i64, c_dbl, c_dbl, c_dbl, c_ptr, c_ptr]),
    "gmt_slot_pattern_f64": (None, [c_i64, c_i64, c_i64, c_ptr, c_ptr]),
    "gmt_stencil1d_f64": (None, [c_i64, c_ptr, c_ptr, c_dbl, C.c_int, c_ptr]),
    "gmt_stencil2d_f64": (None, [C.c_int, c_i64, c_i64, c_ptr, c_i64, c_ptr, c_i64, c_dbl, C.c_int, c_ptr]),
    "gmt_stencil2d_d1_cluster_f64": (None, [c_i64, c_i64, c_ptr, c_i64, c_ptr, c_i64, c_dbl, c_ptr]),
    "gmt_stencil2d_xy_f64": (None, [c_i64, c_i64, c_ptr, c_i64, c_ptr, c_i64, c_dbl, c_dbl, c_ptr]),
    "gmt_copy2d_f64": (None, [c_i64, c_i64, c_ptr, c_i64, c_ptr, c_i64, c_ptr]),
    "gmt_init_stencil1d_f64": (None, [c_i64, C.c_int, c_dbl, c_dbl, c_ptr, c_ptr, C.c_int, C.c_int, c_ptr]),
    "gmt_init_stencil2d_f64": (None, [c_i64, c_i64, C.c_int, C.c_int, c_dbl, c_dbl, c_dbl, c_dbl, c_ptr, c_i64,
                                      c_ptr, c_i64, C.c_int, c_dbl, c_dbl, C.c_int, c_dbl, c_ptr]),
    "gmt_fused_stencil1d_f64": (None, [c_ptr, C.c_int, c_i64, C.c_int, c_ptr]),
    "gmt_fused_stencil2d_f64": (None, [c_ptr, C.c_int, c_i64, c_i64, C.c_int, C.c_int, C.c_int, c_ptr]),
    "gmt_fused_stencil1d_tma_f64": (None, [c_ptr, C.c_int, c_i64, c_ptr]),
    "gmt_halo_exchange1d_f64": (None, [c_ptr, C.c_int, c_ptr]),
    "gmt_halo_exchange2d_f64": (None, [c_ptr, C.c_int, c_i64, c_i64, C.c_int, C.c_int, c_ptr]),
    "gmt_halo2d_flag_count": (c_i64, [C.c_int, c_i64, c_i64]),
    "gmt_debug_set_push_delay": (None, [C.c_uint]),
    "gmt_debug_occupy": (None, [C.c_int, C.c_ulonglong, c_ptr]),
    "gmt_debug_set_grid_factor": (None, [C.c_int]),
    "gmt_publish_epoch_flags": (None, [c_ptr, c_i64, c_ptr, c_ptr]),
    "gmt_allgather_f64": (None, [c_ptr, C.c_int, c_i64, C.c_int, C.c_int, c_ptr]),
    "gmt_gather_f64": (None, [c_ptr, C.c_int, c_i64, C.c_int, c_ptr]),
    "gmt_allreduce_f64": (None, [c_ptr, C.c_int, c_i64, C.c_int, c_ptr]),
    "gmt_daxpy_allreduce_f64": (None, [c_ptr, C.c_int, c_i64, c_dbl, c_ptr]),
    "gmt_axis_sum_allreduce_f64": (None, [c_ptr, C.c_int, c_i64, c_i64, c_i64, C.c_int, c_i64, c_ptr]),
    "gmt_barrier": (None, [c_ptr, C.c_int, c_ptr]),
    "gmt_coll_stage_bytes": (C.c_size_t, [C.c_int, c_i64]),
    "gmt_coll_min_stage_cap": (c_i64, []),
    "gmt_coll_ll_gather_max": (c_i64, []),
    # native exchange engine (engine.cu)
    "gmt_engine_default_device": (None, [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]),
    "gmt_engine2d_create": (None, [C.c_int, C.c_int, C.POINTER(C.c_int), c_i64, c_i64, C.c_int, C.c_int, C.c_int,
                                   c_i64, C.POINTER(c_ptr)]),
    "gmt_engine2d_destroy": (None, [c_ptr]),
    "gmt_engine2d_world": (C.c_int, [c_ptr]),
    "gmt_engine2d_device_groups": (C.c_int, [c_ptr]),
    "gmt_engine2d_rank_device": (None, [c_ptr, C.c_int, C.POINTER(C.c_int)]),
    "gmt_engine2d_array": (None, [c_ptr, C.c_int, C.c_int, C.POINTER(c_ptr), C.POINTER(c_i64), C.POINTER(c_i64)]),
    "gmt_engine2d_init_analytic": (None, [c_ptr, C.c_double]),
    "gmt_engine2d_step": (None, [c_ptr]),
    "gmt_engine2d_exchange": (None, [c_ptr]),
    "gmt_engine2d_time_exchanges": (None, [c_ptr, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "gmt_engine2d_synchronize": (None, [c_ptr]),
    "gmt_engine2d_status": (None, [c_ptr, C.POINTER(C.c_int)]),
    "gmt_engine2d_time_steps": (None, [c_ptr, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "gmt_engine2d_err_norm": (None, [c_ptr, C.c_int, C.POINTER(C.c_double)]),
    # native collective engine (engine_coll.cu)
    "gmt_coll_engine_create": (None, [C.c_int, C.POINTER(C.c_int), c_i64, c_i64, c_i64, C.POINTER(c_ptr)]),
    "gmt_coll_engine_destroy": (None, [c_ptr]),
    "gmt_coll_engine_world": (C.c_int, [c_ptr]),
    "gmt_coll_engine_device_groups": (C.c_int, [c_ptr]),
    "gmt_coll_engine_rank_device": (None, [c_ptr, C.c_int, C.POINTER(C.c_int)]),
    "gmt_coll_engine_buffer": (None, [c_ptr, C.c_int, C.POINTER(c_ptr), C.POINTER(c_i64)]),
    "gmt_coll_engine_stream": (None, [c_ptr, C.c_int, C.POINTER(c_ptr)]),
    "gmt_coll_engine_bind": (None, [c_ptr, C.c_int, c_ptr, c_ptr, c_ptr]),
    "gmt_coll_engine_allgather": (None, [c_ptr, c_i64, C.c_int]),
    "gmt_coll_engine_gather": (None, [c_ptr, c_i64, C.c_int]),
    "gmt_coll_engine_allreduce": (None, [c_ptr, c_i64, C.c_int]),
    "gmt_coll_engine_daxpy_allreduce": (None, [c_ptr, c_i64, c_dbl]),
    "gmt_coll_engine_axis_sum_allreduce": (None, [c_ptr, c_i64, c_i64, c_i64, C.c_int, c_i64]),
    "gmt_coll_engine_barrier": (None, [c_ptr]),
    "gmt_coll_engine_synchronize": (None, [c_ptr]),
    "gmt_coll_engine_status": (None, [c_ptr, C.POINTER(C.c_int)]),
    "gmt_coll_engine_mark": (None, [c_ptr, C.c_int]),
    "gmt_coll_engine_elapsed_ms": (None, [c_ptr, C.POINTER(C.c_double)]),
}


class _Lib:
    """Lazy loader: attribute access returns a checked wrapper around the C entry point."""

    def __init__(self):
        self._dll = None
        self._cache = {}

    def available(self) -> bool:
        return LIB_PATH.exists()

    def load(self):
        if self._dll is None:
            # build outputs are git-ignored: a fresh checkout compiles them on first use, an edited tree
            # recompiles what is stale (one process builds, the other ranks wait on a file lock);
            # no library and no compiler -> a loud error, never a fallback
            try:
                from .._native.build import ensure_built
                ensure_built()
            except Exception as exc:  # noqa: BLE001
                if not LIB_PATH.exists():
                    raise GmtError("load", -1, f"{LIB_PATH} not found and building it failed ({exc}); run "
                                   "`python -m gpu_mpi_tests_b200._native.build` (or __graft_entry__.build())") from exc
            self._dll = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)
            self._dll.gmt_last_error.restype = C.c_char_p
        return self._dll

    def raw(self, name: str):
        """The bare ctypes function (for tests that exercise the C ABI directly)."""
        dll = self.load()
        fn = getattr(dll, name)
        if name in _SIGS:
            restype, argtypes = _SIGS[name]
            fn.argtypes = argtypes
            fn.restype = C.c_int if restype is None else restype
        return fn

    def last_error(self) -> str:
        return (self.load().gmt_last_error() or b"").decode(errors="replace")

    def __getattr__(self, name: str):
        if name.startswith("_"):
            raise AttributeError(name)
        if name in self._cache:
            return self._cache[name]
        if name not in _SIGS:
            raise AttributeError(f"unknown libgmt_b200 entry point {name}")
        fn = self.raw(name)
        restype, _ = _SIGS[name]
        if restype is None:
            def checked(*args, _fn=fn, _name=name):
                rc = _fn(*args)
                if rc != 0:
                    raise GmtError(_name, rc, self.last_error())
                return rc
            wrapper = checked
        else:
            wrapper = fn
        self._cache[name] = wrapper
        return wrapper


lib = _Lib()


def have_native() -> bool:
    return lib.available()


def require_native() -> None:
    """Fail loudly when CUDA is present but the extension is not (no silent fallback)."""
    if not lib.available():
        raise GmtError("load", -1, f"{LIB_PATH} is missing; the CUDA path cannot run without it")
    lib.load()


def env_flag(name: str, default: bool = False) -> bool:
    v = os.environ.get(name)
    if v is None:
        return default
    return v.strip().lower() not in ("", "0", "false", "no", "off")
