"""Fused halo-exchange + stencil plans: the Python owner of the arrays, inboxes, flags and
epoch counters that the kernels in csrc/stencil_fused.cu work on.

Reference counterparts: boundary_exchange (mpi_stencil_gt.cc:83-122), boundary_exchange_x /
boundary_exchange_y (mpi_stencil2d_gt.cc:136-373) followed by the stencil assignment
(mpi_stencil_gt.cc:206, mpi_stencil2d_gt.cc:530-532).  There, every call allocates staging
buffers, packs, calls MPI, unpacks and syncs the device several times.  Here a plan is built
once; `step()` is a single kernel launch per rank with no host synchronisation, safe to
capture in a CUDA graph (the epoch lives in device memory).

A plan works on a `PeerSpace` (parallel/symm.py): in a real job each process hosts one rank and
the neighbours' inboxes are NVLink peer mappings; under `VirtualSpace` one process hosts all
ranks on one GPU and `step()` runs them in one cooperative launch.
"""
from __future__ import annotations

import torch

from ..ops import local
from ..ops._lib import Halo1dRank, Halo2dRank, lib
from .decomp import Grid2D, Slab1D
from .symm import PeerSpace, SymmHeap

F64 = torch.float64
DOMAIN_LENGTH = 8.0


class HaloTimeout(RuntimeError):
    """A bounded in-kernel spin expired: a neighbour never delivered its boundary."""


def _structs_to_device(structs, device) -> torch.Tensor:
    raw = b"".join(bytes(s) for s in structs)
    return torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(device)


class _PlanBase:
    def __init__(self, space: PeerSpace, with_err: bool, spin_limit: int):
        self.space, self.with_err, self.spin_limit = space, bool(with_err), int(spin_limit)
        self.device = space.device
        self.ranks = list(space.ranks_here)
        n = len(self.ranks)
        self._epoch = torch.zeros(n, 8, dtype=torch.int64, device=self.device)
        self._done = torch.zeros(n, 8, dtype=torch.int32, device=self.device)
        self._status = torch.zeros(n, 8, dtype=torch.int32, device=self.device)
        self._errs = torch.zeros(n, 8, dtype=F64, device=self.device)
        self._scratch = torch.zeros(n, int(lib.gmt_reduce_scratch_bytes()), dtype=torch.uint8, device=self.device)
        self.steps = 0

    def _slot(self, rank: int) -> int:
        return self.ranks.index(rank)

    def epochs(self) -> list:
        return self._epoch[:, 0].tolist()

    def check_status(self) -> None:
        """Raise if any in-kernel wait timed out (synchronises)."""
        st = self._status[:, 0].tolist()
        if any(st):
            raise HaloTimeout(f"halo wait expired: status per hosted rank {st} (1+side of the silent neighbour)")

    def err_sumsq(self, rank: int) -> torch.Tensor:
        """Device scalar: sum((numeric - analytic)^2) of the last step (with_err plans only)."""
        return self._errs[self._slot(rank), 0:1]

    def err_norm(self, rank: int) -> float:
        return float(self.err_sumsq(rank).item()) ** 0.5


class FusedStencil1D(_PlanBase):
    """1-D slab-decomposed array y = x^3 on [0, 8) and its fused derivative step."""

    def __init__(self, space: PeerSpace, n_local: int, with_err: bool = True, spin_limit: int = 0,
                 n_global: int | None = None, variant: str = "vector"):
        """variant: "vector" (256-bit LSU loads, the default and the faster one) or "tma" (windows
        streamed into shared memory by cp.async.bulk + mbarriers; no in-kernel error norm)."""
        super().__init__(space, with_err, spin_limit)
        if variant not in ("vector", "tma"):
            raise ValueError("variant must be vector or tma")
        if variant == "tma" and with_err:
            raise ValueError("the TMA variant has no in-kernel error norm: with_err=False")
        self.variant = variant
        if n_local % 2:
            raise ValueError("fused 1-D exchange needs an even n_local (16-byte halo stores)")
        W = space.world
        self.n = n_local
        self.n_global = n_global if n_global is not None else n_local * W
        self.dx = DOMAIN_LENGTH / self.n_global
        self.scale = self.n_global / DOMAIN_LENGTH
        self.slabs = {r: Slab1D(self.n_global, W, r) for r in self.ranks}
        self.z = {r: torch.empty(n_local + 4, dtype=F64, device=self.device) for r in self.ranks}
        self.out = {r: torch.empty(n_local, dtype=F64, device=self.device) for r in self.ranks}
        # symmetric part: 4 inboxes of 32 B and 2 flag words on separate 256-byte lines
        self.heap = SymmHeap(space, 8 * 256)
        self.off_inbox = [[self.heap.take(32) for _p in range(2)] for _s in range(2)]
        self.off_flag = [self.heap.take(8) for _s in range(2)]
        structs = []
        buf = self.heap.buf
        for r in self.ranks:
            k = self._slot(r)
            s = Halo1dRank()
            s.z, s.out, s.n, s.scale = self.z[r].data_ptr(), self.out[r].data_ptr(), n_local, self.scale
            nbr = [self.slabs[r].lo, self.slabs[r].hi]
            for side in range(2):
                nb = nbr[side]
                s.has[side] = int(nb >= 0)
                for p in range(2):
                    s.my_inbox[side][p] = buf.ptr(r, self.off_inbox[side][p])
                    s.peer_inbox[side][p] = buf.ptr(nb, self.off_inbox[1 - side][p]) if nb >= 0 else 0
                s.my_flag[side] = buf.ptr(r, self.off_flag[side])
                s.peer_flag[side] = buf.ptr(nb, self.off_flag[1 - side]) if nb >= 0 else 0
            s.epoch, s.done, s.status = (self._epoch[k].data_ptr(), self._done[k].data_ptr(),
                                         self._status[k].data_ptr())
            s.err_scratch, s.err_sumsq = self._scratch[k].data_ptr(), self._errs[k].data_ptr()
            s.x0, s.dx, s.spin_limit = r * n_local * self.dx, self.dx, self.spin_limit
            structs.append(s)
        self._structs = structs
        self._ranks_dev = _structs_to_device(structs, self.device)

    def init_analytic(self) -> None:
        """y = x^3 with analytic ghosts at the physical ends (mpi_stencil_gt.cc:177-196)."""
        W = self.space.world
        for r in self.ranks:
            local.init_stencil1d(self.z[r], None, self.n, r * self.n * self.dx, self.dx, r == 0, r == W - 1)

    def step(self) -> None:
        st = torch.cuda.current_stream().cuda_stream
        if self.variant == "tma":
            lib.gmt_fused_stencil1d_tma_f64(self._ranks_dev.data_ptr(), len(self.ranks), self.n, st)
        else:
            lib.gmt_fused_stencil1d_f64(self._ranks_dev.data_ptr(), len(self.ranks), self.n, int(self.with_err), st)
        self.steps += 1

    def exchange(self) -> None:
        """The halo exchange alone (boundary_exchange, mpi_stencil_gt.cc:83-122): afterwards z's
        ghost cells hold the neighbours' boundary values; `out` is untouched.  One small kernel
        per rank over the same peer inboxes / flags / epoch as step(); the two can be mixed."""
        lib.gmt_halo_exchange1d_f64(self._ranks_dev.data_ptr(), len(self.ranks),
                                    torch.cuda.current_stream().cuda_stream)
        self.steps += 1


class FusedStencil2D(_PlanBase):
    """2-D array z = x^3 + y^2 on a px x py process grid with ghosts of width gx, gy and the
    fused step out = sx*d/dx + sy*d/dy.

    gx=2, gy=0 on a (world x 1) grid is the reference's dim-0 test, gx=0, gy=2 on (1 x world)
    its dim-1 test (mpi_stencil2d_gt.cc:385-572); gx=gy=2 is the 4-neighbour case of BASELINE
    config 3.  nx, ny are per-rank interior extents (weak scaling, mpi_stencil2d_gt.cc:395).
    periodic=True wraps the process grid, so every rank exchanges all of its used faces (an
    8-GPU 4 x 2 grid then has four live faces on every rank; without it at most three).
    """

    def __init__(self, space: PeerSpace, px: int, py: int, nx: int, ny: int, gx: int = 2, gy: int = 2,
                 with_err: bool = True, spin_limit: int = 0, dx: float | None = None, dy: float | None = None,
                 wx: float = 1.0, wy: float = 1.0, mem_space: str = "device", host_fed: bool = False,
                 periodic: bool = False, land_ghosts: bool = True):
        super().__init__(space, with_err, spin_limit)
        # land_ghosts=False (dim-0 slabs only): step() consumes the x-face ghost values straight from
        # the inbox and leaves z's own ghost cells alone -- same `out`, no strided writes into the array
        self.land_ghosts = bool(land_ghosts) or not (gx and not gy)
        if periodic and with_err:
            raise ValueError("periodic grids have no analytic error norm (x^3 + y^2 does not wrap): with_err=False")
        if px * py != space.world:
            raise ValueError(f"grid {px}x{py} does not match world {space.world}")
        if nx % 4:
            raise ValueError("fused 2-D exchange needs nx % 4 == 0 (256-bit rows)")
        self.px, self.py, self.nx, self.ny, self.gx, self.gy = px, py, nx, ny, gx, gy
        # slab decompositions (one differentiated dim) are the reference's tests: ONE spacing,
        # 8 / n_global_deriv, for both coordinates (mpi_stencil2d_gt.cc:427-448)
        dflt_x = DOMAIN_LENGTH / (nx * px) if (gx or not gy) else DOMAIN_LENGTH / (ny * py)
        dflt_y = DOMAIN_LENGTH / (ny * py) if (gy or not gx) else DOMAIN_LENGTH / (nx * px)
        self.dx = dx if dx is not None else dflt_x
        self.dy = dy if dy is not None else dflt_y
        self.sx = wx / self.dx if gx else 0.0
        self.sy = wy / self.dy if gy else 0.0
        self.periodic = bool(periodic)
        self.grids = {r: Grid2D(px, py, r, self.periodic) for r in self.ranks}
        self.ldz = nx + 2 * gx
        # only inboxes and flags are touched by peers; the arrays themselves may live in any
        # memory space the GPU can address (device, or managed for the TEST_MANAGED flavours)
        from . import array as ga
        self.z = {r: ga.empty((ny + 2 * gy, self.ldz), mem_space, self.device) for r in self.ranks}
        self.out = {r: ga.empty((ny, nx), mem_space, self.device) for r in self.ranks}
        nfx = int(lib.gmt_halo2d_flag_count(0, nx, ny))
        nfy = int(lib.gmt_halo2d_flag_count(2, nx, ny))
        box_x, box_y = 2 * ny * 8, 2 * nx * 8
        total = 4 * (box_x + 256) * bool(gx) + 4 * (box_y + 256) * bool(gy) + 2 * (nfx + nfy + 64) * 8 + 4096
        self.heap = SymmHeap(space, total)
        self.off_inbox = [[0, 0] for _ in range(4)]
        self.off_flag = [0] * 4
        for side in range(4):
            used = gx if side < 2 else gy
            if not used:
                continue
            for p in range(2):
                self.off_inbox[side][p] = self.heap.take(box_x if side < 2 else box_y)
            self.off_flag[side] = self.heap.take((nfx if side < 2 else nfy) * 8)
        buf = self.heap.buf
        # host-fed physical sides ("the host is one more neighbour"): their ghost values arrive by
        # a host-to-device copy into a local inbox while the kernel is already sweeping the interior
        self.host_fed = bool(host_fed)
        self.host_sides = {}     # (rank, side) -> (offset in doubles into hostbox[rank], count, flag tensor)
        self.hostbox = {}
        self._hostflags = {}
        self._hostflags_pinned = {}
        structs = []
        for r in self.ranks:
            k = self._slot(r)
            g = self.grids[r]
            s = Halo2dRank()
            s.z, s.out = self.z[r].data_ptr(), self.out[r].data_ptr()
            s.nx, s.ny, s.ldz, s.ldo, s.gx, s.gy, s.sx, s.sy = nx, ny, self.ldz, nx, gx, gy, self.sx, self.sy
            fed = [side for side in range(4) if (gx if side < 2 else gy) and g.neighbour(side) < 0] if host_fed else []
            total = sum(2 * (ny if side < 2 else nx) for side in fed)
            if fed:
                # one inbox per epoch parity, like the peers' inboxes: the upload for step s+1 may then
                # run while the kernel of step s is still reading step s's values
                self.hostbox[r] = torch.zeros(2, total, dtype=F64, device=self.device)
                self._hostflags[r] = torch.zeros(4, max(nfx, nfy), dtype=torch.int64, device=self.device)
                # two pinned sources: a pipelined caller enqueues step s+1 before step s has run
                self._hostflags_pinned[r] = [torch.zeros(4, max(nfx, nfy), dtype=torch.int64, pin_memory=True)
                                             for _ in range(2)]
            off = 0
            for side in range(4):
                used = gx if side < 2 else gy
                nb = g.neighbour(side) if used else -1
                s.has[side] = int(nb >= 0)
                opp = side ^ 1
                for p in range(2):
                    s.my_inbox[side][p] = buf.ptr(r, self.off_inbox[side][p]) if used else 0
                    s.peer_inbox[side][p] = buf.ptr(nb, self.off_inbox[opp][p]) if nb >= 0 else 0
                s.my_flag[side] = buf.ptr(r, self.off_flag[side]) if used else 0
                s.peer_flag[side] = buf.ptr(nb, self.off_flag[opp]) if nb >= 0 else 0
                if side in fed:
                    cnt = 2 * (ny if side < 2 else nx)
                    s.has[side] = 2
                    for p in range(2):
                        s.my_inbox[side][p] = self.hostbox[r][p][off:off + cnt].data_ptr()
                    nflags = nfx if side < 2 else nfy
                    s.my_flag[side] = self._hostflags[r][side].data_ptr()
                    self.host_sides[(r, side)] = (off, cnt, self._hostflags[r][side][:nflags])
                    off += cnt
            s.epoch, s.done, s.status = (self._epoch[k].data_ptr(), self._done[k].data_ptr(),
                                         self._status[k].data_ptr())
            s.err_scratch, s.err_sumsq = self._scratch[k].data_ptr(), self._errs[k].data_ptr()
            s.x0, s.y0, s.dx, s.dy = g.cx * nx * self.dx, g.cy * ny * self.dy, self.dx, self.dy
            s.spin_limit = self.spin_limit
            structs.append(s)
        self._structs = structs
        self._ranks_dev = _structs_to_device(structs, self.device)

    def origin(self, rank: int) -> tuple[float, float]:
        g = self.grids[rank]
        return g.cx * self.nx * self.dx, g.cy * self.ny * self.dy

    def init_analytic(self, poison: float = 0.0) -> None:
        """z = x^3 + y^2; ghosts analytic on physical sides only (mpi_stencil2d_gt.cc:445-497)."""
        for r in self.ranks:
            x0, y0 = self.origin(r)
            mask = self.grids[r].physical_mask()
            local.init_stencil2d(self.z[r], None, self.nx, self.ny, self.gx, self.gy, x0, y0, self.dx, self.dy,
                                 0, self.sx, self.sy, mask, poison)

    def analytic(self, rank: int) -> torch.Tensor:
        """The exact derivative the numeric result is compared with (host-free)."""
        x0, y0 = self.origin(rank)
        xs = x0 + torch.arange(self.nx, dtype=F64, device=self.device) * self.dx
        ys = y0 + torch.arange(self.ny, dtype=F64, device=self.device) * self.dy
        a = torch.zeros(self.ny, self.nx, dtype=F64, device=self.device)
        if self.gx:
            a += (self.sx * self.dx) * 3.0 * xs[None, :] ** 2
        if self.gy:
            a += (self.sy * self.dy) * 2.0 * ys[:, None]
        return a

    def step(self, result_slot: int | None = None) -> None:
        """One fused exchange + stencil step.  result_slot (0/1, after host_result_slots()): the error
        word of this step goes straight into that pinned host word instead of device memory."""
        desc = self._ranks_dev if result_slot is None else self._ranks_dev_slot[result_slot]
        lib.gmt_fused_stencil2d_f64(desc.data_ptr(), len(self.ranks), self.nx, self.ny, self.gx, self.gy,
                                    int(self.with_err) | (0 if self.land_ghosts else 2),
                                    torch.cuda.current_stream().cuda_stream)
        self.steps += 1

    def host_result_slots(self) -> list:
        """Two pinned host tensors (one fp64 word per hosted rank each) and, for each, a copy of the
        rank descriptors whose err_sumsq points at it: pinned memory is mapped into the device's
        address space (unified addressing), so the kernel's last CTA stores the step's error word
        straight to the host and no device-to-host copy has to be queued behind the kernel.  Two
        slots, because a pipelined caller reads step s's word while step s+1 is already running."""
        if getattr(self, "_result_host", None) is None:
            self._result_host = [torch.zeros(len(self.ranks), 8, dtype=F64, pin_memory=True) for _ in range(2)]
            self._ranks_dev_slot = []
            for slot in range(2):
                structs = []
                for k, s0 in enumerate(self._structs):
                    s = type(s0).from_buffer_copy(bytes(s0))   # ctypes structures with pointers cannot be copy.copy()ed
                    s.err_sumsq = self._result_host[slot][k].data_ptr()
                    structs.append(s)
                self._ranks_dev_slot.append(_structs_to_device(structs, self.device))
        return self._result_host

    def exchange(self) -> None:
        """The halo exchange alone (boundary_exchange_x / _y, mpi_stencil2d_gt.cc:136-373): afterwards
        z's ghost cells of every side with a neighbour hold that neighbour's boundary rows / columns;
        `out` is untouched, so any compute step of the caller's can follow.  One small kernel per
        rank over the same peer inboxes / flags / epoch as step(); the two can be mixed."""
        lib.gmt_halo_exchange2d_f64(self._ranks_dev.data_ptr(), len(self.ranks), self.nx, self.ny, self.gx, self.gy,
                                    torch.cuda.current_stream().cuda_stream)
        self.steps += 1

    # ---- host-fed sides ------------------------------------------------------------------
    def host_side_shape(self, side: int) -> tuple:
        """Shape of the host tensor feeding `side`: x faces (ny, 2) [2 ghost rows per column],
        y faces (2, nx) [2 ghost columns]."""
        return (self.ny, 2) if side < 2 else (2, self.nx)

    def ghost_view(self, rank: int, side: int) -> torch.Tensor:
        """The array's own ghost cells of a side as a strided view with host_side_shape."""
        z, gx, gy, nx, ny = self.z[rank], self.gx, self.gy, self.nx, self.ny
        return [z[gy:gy + ny, 0:2], z[gy:gy + ny, gx + nx:gx + nx + 2], z[0:2, gx:gx + nx],
                z[gy + ny:gy + ny + 2, gx:gx + nx]][side]

    def publish_host_sides_by_copy(self, rank: int, slot: int = 0) -> None:
        """Publish the host-fed sides' flags for the NEXT step with a host-to-device copy on the
        current stream (enqueue it right after the copy of the values).  A copy engine needs no SM,
        so this cannot be starved by the very kernel that is waiting for it -- a publishing kernel
        could be, once the fused kernel's CTAs occupy every SM.  The pinned source `slot` holds the
        epoch of the next step; it is rewritten by the host before every step (a pipelined caller
        alternates the two slots, so the source of a step still in the queue is never rewritten)."""
        if rank in self._hostflags:
            self._hostflags_pinned[rank][slot].fill_(self.steps + 1)
            self._hostflags[rank].copy_(self._hostflags_pinned[rank][slot], non_blocking=True)

    def refresh_host_epoch(self, slot: int = 0) -> None:
        """Before replaying a captured step: put the next epoch into the pinned flag source `slot`."""
        for rank in self._hostflags_pinned:
            self._hostflags_pinned[rank][slot].fill_(self.steps + 1)

    def next_parity(self) -> int:
        """Inbox parity (index into hostbox[rank]) the NEXT step or exchange will read."""
        return (self.steps + 1) & 1

    def publish_host_sides(self, rank: int) -> None:
        """Kernel flavour, for use on the SAME stream BEFORE the fused kernel (plain step() of a
        host-fed plan without new host data): the next step's inbox takes over what the last one
        held, and the host-fed sides of the step about to run are marked as arrived."""
        k = self._slot(rank)
        if rank in self.hostbox:
            p = self.next_parity()
            self.hostbox[rank][p].copy_(self.hostbox[rank][p ^ 1], non_blocking=True)
        st = torch.cuda.current_stream().cuda_stream
        for (r, side), (_off, _cnt, flags) in self.host_sides.items():
            if r == rank:
                lib.gmt_publish_epoch_flags(flags.data_ptr(), flags.numel(), self._epoch[k].data_ptr(), st)

    def halo_bytes_out(self, rank: int) -> int:
        """Bytes this rank stores into neighbours' inboxes per step."""
        g = self.grids[rank]
        b = 0
        for side in range(4):
            used = self.gx if side < 2 else self.gy
            if used and g.neighbour(side) >= 0:
                b += 2 * (self.ny if side < 2 else self.nx) * 8
        return b
