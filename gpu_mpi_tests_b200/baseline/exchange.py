"""Library-only restatement of the reference's halo exchange + stencil (the baseline).

Structure follows boundary_exchange (mpi_stencil_gt.cc:83-122) and boundary_exchange_x/_y
(mpi_stencil2d_gt.cc:136-373): pack boundary strips into staging buffers with a kernel, sync,
post non-blocking receives and sends through the communication LIBRARY (NCCL on GPUs, gloo
on CPUs -- standing in for GPU-aware MPI, which is not installable here), wait, unpack with a
kernel, sync; then run the un-fused stencil kernel.  The staging knobs are the reference's:

  stage='device'  device staging buffers (dim 0 always; dim 1 when buf:1)
  stage='none'    dim 1 only: send/recv directly on the array's contiguous ghost slabs (buf:0)
  stage='host'    additionally bounce through pinned host buffers and a CPU transport
                  (dim 0 with buf:1 = the reference's stage_host, mpi_stencil2d_gt.cc:147-169)
  mem_space       'device' or 'managed' arrays (TEST_MANAGED flavours)

Nothing in this module may be called from the fused product path.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import local
from ..parallel import array as ga
from ..parallel import decomp
from ..parallel import runtime as _rt

F64 = torch.float64
DOMAIN = 8.0
_gloo_group = None


def _host_group():
    """A gloo group next to the NCCL one, for host-staged transfers."""
    global _gloo_group
    if _gloo_group is None:
        _gloo_group = dist.new_group(backend="gloo")
    return _gloo_group


def _sync(device: torch.device) -> None:
    if device.type == "cuda":
        torch.cuda.synchronize(device)


class CommError(RuntimeError):
    """A communication-library call failed: the analogue of a non-MPI_SUCCESS return code caught
    by the reference's check() (mpi_stencil2d_gt.cc:32-40), which prints `file:line error code`
    and exits with status 2."""

    def __init__(self, where: str, code: int, detail: str):
        super().__init__(f"{where} error {code}: {detail}")
        self.where, self.code = where, code


def comm_check(fn, *args, **kwargs):
    """CHECK(x) for library communication calls: re-raise failures as CommError carrying the call
    site, so that the CLI programs can print the reference's line and exit(2)."""
    import inspect
    try:
        return fn(*args, **kwargs)
    except CommError:
        raise
    except Exception as e:  # noqa: BLE001
        fr = inspect.stack()[1]
        raise CommError(f"{fr.filename.rsplit('/', 1)[-1]}:{fr.lineno}", 1, repr(e)) from e


def exchange_pairs(sends: list, recvs: list, group=None) -> None:
    """Post all receives and sends (MPI_Irecv/MPI_Isend), then wait for all (MPI_Waitall).
    sends/recvs: lists of (tensor, peer_rank)."""
    ops = [dist.P2POp(dist.irecv, t, p, group=group) for (t, p) in recvs]
    ops += [dist.P2POp(dist.isend, t, p, group=group) for (t, p) in sends]
    if not ops:
        return

    def run():
        for w in dist.batch_isend_irecv(ops):
            w.wait()

    comm_check(run)


class BaselineStencil1D:
    """1-D slab array with in-array ghosts; exchange directly on the array like the reference
    (16-byte messages, mpi_stencil_gt.cc:96-106)."""

    def __init__(self, rt: "_rt.Runtime", n_local: int, n_global: int | None = None, variant: str = "vector"):
        self.rt, self.n = rt, n_local
        self.n_global = n_global if n_global is not None else n_local * rt.world_size
        self.slab = decomp.Slab1D(self.n_global, rt.world_size, rt.rank)
        self.dx = DOMAIN / self.n_global
        self.scale = self.n_global / DOMAIN
        self.device = rt.device
        self.variant = variant
        r = rt.rank
        self.z = {r: torch.empty(n_local + 4, dtype=F64, device=self.device)}
        self.out = {r: torch.empty(n_local, dtype=F64, device=self.device)}
        self.launches_per_step = 1

    def init_analytic(self, poison: float = 0.0) -> None:
        r, W = self.rt.rank, self.rt.world_size
        local.init_stencil1d(self.z[r], None, self.n, r * self.n * self.dx, self.dx, r == 0, r == W - 1)

    def exchange(self) -> None:
        y = self.z[self.rt.rank]
        n = self.n
        sends, recvs = [], []
        if self.slab.lo >= 0:
            recvs.append((y[0:2], self.slab.lo))
            sends.append((y[2:4], self.slab.lo))
        if self.slab.hi >= 0:
            recvs.append((y[n + 2:n + 4], self.slab.hi))
            sends.append((y[n:n + 2], self.slab.hi))
        exchange_pairs(sends, recvs)
        _sync(self.device)

    def stencil(self) -> None:
        r = self.rt.rank
        local.stencil1d(self.z[r], self.scale, out=self.out[r], variant=self.variant)

    def step(self) -> None:
        self.exchange()
        self.stencil()

    def analytic(self, rank: int) -> torch.Tensor:
        xs = (rank * self.n + torch.arange(self.n, dtype=F64, device=self.device)) * self.dx
        return 3.0 * xs ** 2


class BaselineStencil2D:
    def __init__(self, rt: "_rt.Runtime", px: int, py: int, nx: int, ny: int, gx: int, gy: int,
                 stage: str = "device", mem_space: str = "device", variant: str = "vector",
                 dx: float | None = None, dy: float | None = None):
        if px * py != rt.world_size:
            raise ValueError(f"grid {px}x{py} does not match world {rt.world_size}")
        if stage not in ("device", "none", "host"):
            raise ValueError("stage must be device, none or host")
        self.rt, self.px, self.py, self.nx, self.ny, self.gx, self.gy = rt, px, py, nx, ny, gx, gy
        self.stage, self.mem_space, self.variant = stage, mem_space, variant
        self.device = rt.device
        self.grid = decomp.Grid2D(px, py, rt.rank)
        # slab decompositions use the reference's single spacing (mpi_stencil2d_gt.cc:427-448)
        dflt_x = DOMAIN / (nx * px) if (gx or not gy) else DOMAIN / (ny * py)
        dflt_y = DOMAIN / (ny * py) if (gy or not gx) else DOMAIN / (nx * px)
        self.dx = dx if dx is not None else dflt_x
        self.dy = dy if dy is not None else dflt_y
        self.sx = 1.0 / self.dx if gx else 0.0
        self.sy = 1.0 / self.dy if gy else 0.0
        r = rt.rank
        self.ldz = nx + 2 * gx
        self.z = {r: ga.empty((ny + 2 * gy, self.ldz), mem_space, self.device)}
        self.out = {r: ga.empty((ny, nx), mem_space, self.device)}
        z = self.z[r]
        # boundary (send) and ghost (recv) views per side, column-major torch layout (n_dim1, n_dim0)
        self.send_view, self.recv_view = {}, {}
        if gx:
            rows = slice(gy, gy + ny)
            self.send_view[0], self.recv_view[0] = z[rows, gx:gx + 2], z[rows, 0:2]
            self.send_view[1], self.recv_view[1] = z[rows, nx:nx + 2], z[rows, gx + nx:gx + nx + 2]
        if gy:
            cols = slice(gx, gx + nx)
            self.send_view[2], self.recv_view[2] = z[gy:gy + 2, cols], z[0:2, cols]
            self.send_view[3], self.recv_view[3] = z[ny:ny + 2, cols], z[gy + ny:gy + ny + 2, cols]
        self.neighbours = {s: self.grid.neighbour(s) for s in self.send_view}
        self.active = [s for s, nb in self.neighbours.items() if nb >= 0]
        # zero-copy is only possible where the strips are contiguous in memory
        self.zero_copy = {s: (stage == "none" and self.send_view[s].is_contiguous()) for s in self.active}
        self.sbuf, self.rbuf, self.h_sbuf, self.h_rbuf = {}, {}, {}, {}
        for s in self.active:
            if self.zero_copy[s]:
                continue
            shp = tuple(self.send_view[s].shape)
            self.sbuf[s] = torch.empty(shp, dtype=F64, device=self.device)
            self.rbuf[s] = torch.empty(shp, dtype=F64, device=self.device)
            if stage == "host":
                pin = self.device.type == "cuda"
                self.h_sbuf[s] = torch.empty(shp, dtype=F64, pin_memory=pin)
                self.h_rbuf[s] = torch.empty(shp, dtype=F64, pin_memory=pin)
        n_staged = sum(1 for s in self.active if not self.zero_copy[s])
        self.launches_per_step = 1 + 2 * n_staged  # pack + unpack per staged side, + stencil

    def init_analytic(self, poison: float = 0.0) -> None:
        r = self.rt.rank
        local.init_stencil2d(self.z[r], None, self.nx, self.ny, self.gx, self.gy, self.grid.cx * self.nx * self.dx,
                             self.grid.cy * self.ny * self.dy, self.dx, self.dy, 0, self.sx, self.sy,
                             self.grid.physical_mask(), poison)

    def exchange(self) -> None:
        """One boundary exchange, every phase followed by the sync the reference has."""
        dev = self.device
        host = self.stage == "host"
        # pack (view-assign kernels, mpi_stencil2d_gt.cc:166,172,289,292)
        for s in self.active:
            if not self.zero_copy[s]:
                local.copy2d(self.send_view[s], self.sbuf[s])
                if host:
                    self.h_sbuf[s].copy_(self.sbuf[s], non_blocking=True)
        _sync(dev)  # send buffers ready (mpi_stencil2d_gt.cc:202,321)
        sends, recvs = [], []
        for s in self.active:
            nb = self.neighbours[s]
            if self.zero_copy[s]:
                sends.append((self.send_view[s], nb))
                recvs.append((self.recv_view[s], nb))
            elif host:
                sends.append((self.h_sbuf[s], nb))
                recvs.append((self.h_rbuf[s], nb))
            else:
                sends.append((self.sbuf[s], nb))
                recvs.append((self.rbuf[s], nb))
        group = _host_group() if (host and self.rt.backend == "nccl") else None
        if self.rt.world_size > 1:
            exchange_pairs(sends, recvs, group)
        # unpack (mpi_stencil2d_gt.cc:236-252,356-369)
        for s in self.active:
            if not self.zero_copy[s]:
                if host:
                    self.rbuf[s].copy_(self.h_rbuf[s], non_blocking=True)
                local.copy2d(self.rbuf[s], self.recv_view[s])
        _sync(dev)

    def stencil(self) -> None:
        r = self.rt.rank
        z, out = self.z[r], self.out[r]
        if self.gx and self.gy:
            local.stencil2d_xy(z, self.sx, self.sy, out=out)
        elif self.gx:
            local.stencil2d(z, 0, self.sx, out=out, variant=self.variant)
        else:
            local.stencil2d(z, 1, self.sy, out=out, variant=self.variant)

    def step(self) -> None:
        self.exchange()
        self.stencil()

    def analytic(self, rank: int) -> torch.Tensor:
        g = decomp.Grid2D(self.px, self.py, rank)
        xs = (g.cx * self.nx + torch.arange(self.nx, dtype=F64, device=self.device)) * self.dx
        ys = (g.cy * self.ny + torch.arange(self.ny, dtype=F64, device=self.device)) * self.dy
        a = torch.zeros(self.ny, self.nx, dtype=F64, device=self.device)
        if self.gx:
            a += 3.0 * xs[None, :] ** 2
        if self.gy:
            a += 2.0 * ys[:, None]
        return a
