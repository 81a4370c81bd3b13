"""Process-group bootstrap and rank -> GPU binding.

Replaces MPI_Init/Comm_size/Comm_rank/Finalize, the five copies of set_rank_device
(mpi_daxpy.cc:36-62, mpi_daxpy_nvtx.cc:43-69, mpi_daxpy_gt.cc:26-45, mpi_stencil_gt.cc:61-81,
mpi_stencil2d_gt.cc:112-133), the SYCL twin get_rank_queue (mpi_stencil2d_sycl.cc:183-209),
get_node_count (mpi_daxpy_nvtx.cc:72-82) and the device/vendor id report
(mpi_stencil_gt.cc:151-152).

One process per GPU; ``torch.distributed`` does the rendezvous (NCCL when CUDA is present,
gloo otherwise) and serves as control plane (scalar reporting reduces, barriers) and as the
*baseline* data plane.  The product data plane is peer memory (parallel/symm.py).

Single-process use needs no launcher: with WORLD_SIZE unset the runtime is a world of one
and every collective helper degenerates to the identity.
"""
from __future__ import annotations

import datetime
import os
from dataclasses import dataclass, field

import torch
import torch.distributed as dist

from ..ops import _lib
from ..utils import report


class RankDeviceError(RuntimeError):
    """n_ranks is larger than, and not a multiple of, n_devices (mpi_daxpy.cc:43-47)."""


def map_rank_to_device(n_ranks: int, rank: int, n_devices: int) -> tuple[int, int]:
    """The reference's block mapping with oversubscription (mpi_daxpy.cc:41-55).

    Returns (device, ranks_per_device).  With more ranks than devices, consecutive ranks share
    a device (the "42 procs on 6 GPUs under MPS" use case, daxpy.cu:6-7).
    """
    if n_devices <= 0:
        raise RankDeviceError("no devices")
    if n_ranks > n_devices:
        if n_ranks % n_devices != 0:
            raise RankDeviceError(report.rank_device_error(n_ranks, n_devices))
        ranks_per_device = n_ranks // n_devices
        return rank // ranks_per_device, ranks_per_device
    return rank, 1


def pack_vendor_id(pci_domain: int, pci_bus: int, pci_device: int) -> int:
    """32-bit id printed as [%d:0x%08x]: PCI domain/bus/device packed like gtensor's CUDA
    backend does for device_get_vendor_id (mpi_stencil_gt.cc:152)."""
    return ((pci_domain & 0xFFFF) << 16) | ((pci_bus & 0xFF) << 8) | (pci_device & 0xFF)


@dataclass
class Runtime:
    world_size: int = 1
    rank: int = 0
    local_rank: int = 0
    local_world_size: int = 1
    backend: str = "none"
    device: torch.device = field(default_factory=lambda: torch.device("cpu"))
    n_devices: int = 0
    device_id: int = 0
    vendor_id: int = 0
    ranks_per_device: int = 1
    owns_pg: bool = False
    props: object | None = None

    # -------------------------------------------------------------- facts
    @property
    def is_cuda(self) -> bool:
        return self.device.type == "cuda"

    @property
    def node_count(self) -> int:
        """world_size / ranks-on-this-node (get_node_count, mpi_daxpy_nvtx.cc:72-82)."""
        return max(1, self.world_size // max(1, self.local_world_size))

    @property
    def distributed(self) -> bool:
        return self.world_size > 1

    def total_device_memory(self) -> int:
        if self.is_cuda and self.props is not None:
            return int(self.props.total_global_mem)
        try:
            return os.sysconf("SC_PAGE_SIZE") * os.sysconf("SC_PHYS_PAGES")
        except (ValueError, OSError):
            return 0

    def mem_info(self) -> tuple[int, int]:
        """(free, total) bytes (cudaMemGetInfo, mpi_daxpy_nvtx.cc:201-205)."""
        if self.is_cuda:
            import ctypes as C
            f, t = C.c_uint64(), C.c_uint64()
            _lib.lib.gmt_mem_info(C.byref(f), C.byref(t))
            return int(f.value), int(t.value)
        total = self.total_device_memory()
        return total, total

    def rank_device_line(self) -> str:
        """RANK[r/N] => DEVICE[d/D] mem=<total / ranks per device> (mpi_daxpy.cc:56-61)."""
        return report.rank_device(self.rank, self.world_size, self.device_id, max(1, self.n_devices),
                                  self.total_device_memory() // max(1, self.ranks_per_device))

    # -------------------------------------------------------------- control plane
    def barrier(self) -> None:
        if self.distributed:
            if self.backend == "nccl":
                dist.barrier(device_ids=[self.device.index])
            else:
                dist.barrier()

    def reduce_scalar(self, value: float, op: str = "sum") -> float:
        """All-reduce of one host double (stands in for MPI_Reduce-to-root of a scalar,
        mpi_stencil2d_gt.cc:563,566,643; every rank gets the result)."""
        if not self.distributed:
            return float(value)
        t = torch.tensor([value], dtype=torch.float64, device=self.device)
        dist.all_reduce(t, op={"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op])
        return float(t.item())

    def gather_objects(self, obj) -> list:
        if not self.distributed:
            return [obj]
        out = [None] * self.world_size
        dist.all_gather_object(out, obj)
        return out

    def finalize(self) -> None:
        """MPI_Finalize."""
        if self.owns_pg and dist.is_initialized():
            try:
                if self.is_cuda:
                    torch.cuda.synchronize()
                dist.barrier(device_ids=[self.device.index]) if self.backend == "nccl" else dist.barrier()
            except Exception:  # noqa: BLE001 - best effort on the way out
                pass
            dist.destroy_process_group()
            self.owns_pg = False


_current: Runtime | None = None


def current() -> Runtime:
    global _current
    if _current is None:
        _current = init()
    return _current


def use_cuda_default() -> bool:
    if _lib.env_flag("GMT_FORCE_CPU"):
        return False
    return torch.cuda.is_available()


def init(backend: str | None = None, use_cuda: bool | None = None, allow_oversubscribe: bool = True,
         timeout_s: float = 600.0) -> Runtime:
    """MPI_Init + set_rank_device.  Reads RANK / WORLD_SIZE / LOCAL_RANK / LOCAL_WORLD_SIZE /
    MASTER_ADDR / MASTER_PORT (the torchrun contract)."""
    global _current
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", str(world)))
    cuda = use_cuda_default() if use_cuda is None else use_cuda
    rt = Runtime(world_size=world, rank=rank, local_rank=local_rank, local_world_size=local_world)

    if cuda:
        _lib.require_native()
        n_dev = torch.cuda.device_count()
        # the reference maps by global rank; LOCAL_RANK is the multi-node-correct choice
        dev, rpd = map_rank_to_device(local_world, local_rank, n_dev)
        if rpd > 1 and not allow_oversubscribe:
            raise RankDeviceError(f"{local_world} ranks on {n_dev} devices needs oversubscription")
        torch.cuda.set_device(dev)
        _lib.lib.gmt_set_device(dev)
        props = _lib.DeviceProps()
        import ctypes as C
        _lib.lib.gmt_device_props_get(dev, C.byref(props))
        rt.device = torch.device("cuda", dev)
        rt.n_devices, rt.device_id, rt.ranks_per_device, rt.props = n_dev, dev, rpd, props
        rt.vendor_id = pack_vendor_id(props.pci_domain, props.pci_bus, props.pci_device)
        default_backend = "nccl" if rpd == 1 else "gloo"  # NCCL refuses two ranks on one device
    else:
        n_dev = int(os.environ.get("GMT_FAKE_DEVICES", str(max(1, local_world))))
        dev, rpd = map_rank_to_device(local_world, local_rank, n_dev)
        rt.device = torch.device("cpu")
        rt.n_devices, rt.device_id, rt.ranks_per_device = n_dev, dev, rpd
        rt.vendor_id = 0
        default_backend = "gloo"

    if world > 1:
        rt.backend = backend or default_backend
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            kwargs = {}
            if rt.backend == "nccl":
                kwargs["device_id"] = rt.device
            dist.init_process_group(backend=rt.backend, rank=rank, world_size=world,
                                    timeout=datetime.timedelta(seconds=timeout_s), **kwargs)
            rt.owns_pg = True
    _current = rt
    return rt


def finalize() -> None:
    global _current
    if _current is not None:
        _current.finalize()
        _current = None
