"""2-D stencil workload: the framework's flagship "model".

One object owns a rank's share of a 2-D fp64 field z = x^3 + y^2 on [0,8)^2, decomposed either
as the reference does (slabs along the differentiated dim, mpi_stencil2d_gt.cc:385-572) or on a
2-D process grid with four face neighbours (BASELINE config 3), and advances it by "steps":

    step()      exchange ghosts with the neighbours and apply the 5-point derivative(s)
                -- ONE fused kernel launch (parallel/halo.py), or the library-only baseline
                (baseline/exchange.py + un-fused kernels) when impl="baseline";
    step_e2e()  what a time-stepping user calls: upload this step's physical-boundary values
                from pinned host memory, step(), read the error norm back to the host.

The CPU/gloo flavour (no GPU) runs the same API on torch CPU tensors with the baseline
exchange.
"""
from __future__ import annotations

import torch

from ..ops import local
from ..parallel import decomp
from ..parallel import runtime as _rt
from ..parallel.halo import FusedStencil2D
from ..parallel.symm import PeerSpace, make_space

F64 = torch.float64
DOMAIN = 8.0


class Stencil2DWorkload:
    def __init__(self, rt: "_rt.Runtime | None" = None, nx: int = 16384, ny: int = 16384, mode: str = "grid",
                 impl: str = "fused", with_err: bool = True, space: PeerSpace | None = None,
                 grid: tuple | None = None, stage: str = "none", mem_space: str = "device",
                 spin_limit: int = 0, host_fed: bool = False):
        """mode: 'grid' (gx=gy=2, px x py ranks), 'dim0' (gx=2, world x 1), 'dim1' (gy=2, 1 x world).
        nx, ny: per-rank interior extents (weak scaling)."""
        self.rt = rt or _rt.current()
        self.mode, self.impl, self.with_err = mode, impl, with_err
        world = space.world if space is not None else self.rt.world_size
        if mode == "grid":
            self.px, self.py = grid if grid is not None else decomp.best_grid(world)
            self.gx = self.gy = 2
        elif mode == "dim0":
            self.px, self.py, self.gx, self.gy = world, 1, 2, 0
        elif mode == "dim1":
            self.px, self.py, self.gx, self.gy = 1, world, 0, 2
        else:
            raise ValueError("mode must be grid, dim0 or dim1")
        self.nx, self.ny = nx, ny
        self.device = self.rt.device if space is None else space.device
        self.stage, self.mem_space = stage, mem_space
        self.launches_per_step = 1
        if impl == "fused":
            if self.device.type != "cuda":
                raise RuntimeError("impl='fused' needs CUDA; use impl='baseline' on the CPU/gloo path")
            self.space = space if space is not None else make_space(self.rt)
            self.plan = FusedStencil2D(self.space, self.px, self.py, nx, ny, self.gx, self.gy, with_err=with_err,
                                       spin_limit=spin_limit, mem_space=mem_space, host_fed=host_fed)
            self.ranks = self.plan.ranks
            self.z, self.out = self.plan.z, self.plan.out
            self.dx, self.dy, self.sx, self.sy = self.plan.dx, self.plan.dy, self.plan.sx, self.plan.sy
        elif impl == "baseline":
            from ..baseline.exchange import BaselineStencil2D
            self.plan = BaselineStencil2D(self.rt, self.px, self.py, nx, ny, self.gx, self.gy, stage=stage,
                                          mem_space=mem_space)
            self.ranks = [self.rt.rank]
            self.z, self.out = self.plan.z, self.plan.out
            self.dx, self.dy, self.sx, self.sy = self.plan.dx, self.plan.dy, self.plan.sx, self.plan.sy
            self.launches_per_step = self.plan.launches_per_step
        else:
            raise ValueError("impl must be fused or baseline")
        self._host_bc = None
        self.host_fed = bool(host_fed) and impl == "fused"

    # ------------------------------------------------------------------ setup / step
    def setup(self, poison: float = 0.0) -> None:
        self.plan.init_analytic(poison=poison)
        if self.host_fed:
            # the kernel takes the host-fed ghosts from its inbox: start it off with the analytic values
            for (r, side), (off, cnt, _f) in self.plan.host_sides.items():
                self.plan.hostbox[r][off:off + cnt].view(self.plan.host_side_shape(side)).copy_(
                    self.plan.ghost_view(r, side))

    def step(self) -> None:
        if self.host_fed:
            for r in self.ranks:  # no new host data this step: re-announce what the inbox holds
                self.plan.publish_host_sides(r)
        self.plan.step()

    def points_per_rank(self) -> int:
        return self.nx * self.ny

    def hbm_bytes_per_step(self) -> int:
        """Algorithmic HBM traffic per rank: read z once, write out once."""
        return 16 * self.nx * self.ny

    # ------------------------------------------------------------------ end to end
    def _prepare_e2e_host_fed(self) -> None:
        """host_fed plans: a pinned buffer per hosted rank mirrors plan.hostbox (the kernel's inbox
        for the host-fed sides), so ONE H2D copy delivers all sides and no scatter kernel is needed."""
        plan = self.plan
        self._bc_views = {}
        self._fed = []   # (pinned host buffer, device inbox) per hosted rank that has host-fed sides
        total = 0
        for r in self.ranks:
            box = plan.hostbox.get(r)
            if box is None:
                continue
            h = torch.empty(box.numel(), dtype=F64, pin_memory=True)
            for (rr, side), (off, cnt, _f) in plan.host_sides.items():
                if rr == r:
                    # inbox layouts: x faces [2*j + r] == row-major (ny, 2); y faces [r*nx + i] == (2, nx)
                    h[off:off + cnt].view(plan.host_side_shape(side)).copy_(plan.ghost_view(rr, side))
            self._fed.append((h, box))
            total += box.numel()
        self._bc_total = total
        self._host_bc = self._fed[0][0] if self._fed else torch.empty(1, dtype=F64, pin_memory=True)
        self._err_dev = torch.zeros(len(self.ranks), dtype=F64, device=self.device)
        self._err_host = torch.zeros(len(self.ranks), dtype=F64, pin_memory=True)
        self._graph = None
        self._graph_failed = False
        self._side_stream = torch.cuda.Stream(device=self.device)

    def _prepare_e2e(self) -> None:
        """Per-step host input and output: ONE pinned buffer holding this rank's physical-boundary
        ghost values (all sides back to back), its device staging twin, and a pinned scalar for the
        per-step result."""
        if self.host_fed:
            return self._prepare_e2e_host_fed()
        pin = self.device.type == "cuda"
        self._bc_views = {}
        total = 0
        for r in self.ranks:
            g = decomp.Grid2D(self.px, self.py, r)
            z = self.z[r]
            gx, gy, nx, ny = self.gx, self.gy, self.nx, self.ny
            views = {}
            if gx and g.neighbour(0) < 0:
                views[0] = z[gy:gy + ny, 0:2]
            if gx and g.neighbour(1) < 0:
                views[1] = z[gy:gy + ny, gx + nx:gx + nx + 2]
            if gy and g.neighbour(2) < 0:
                views[2] = z[0:2, gx:gx + nx]
            if gy and g.neighbour(3) < 0:
                views[3] = z[gy + ny:gy + ny + 2, gx:gx + nx]
            for side, v in views.items():
                self._bc_views[(r, side)] = (total, v)
                total += v.numel()
        self._bc_total = total
        self._host_bc = torch.empty(max(total, 1), dtype=F64, pin_memory=pin)
        self._dev_bc = torch.empty(max(total, 1), dtype=F64, device=self.device)
        for (r, side), (off, v) in self._bc_views.items():
            self._host_bc[off:off + v.numel()].view(v.shape).copy_(v)
        self._err_dev = torch.zeros(len(self.ranks), dtype=F64, device=self.device)
        self._err_host = torch.zeros(len(self.ranks), dtype=F64, pin_memory=pin)
        self._graph = None
        self._graph_failed = False

    def e2e_bytes(self) -> tuple[int, int]:
        """(host->device, device->host) bytes per step, counted from the tensors copied."""
        if self._host_bc is None:
            self._prepare_e2e()
        return self._bc_total * 8, self._err_host.numel() * 8

    def _e2e_body(self) -> None:
        """Stream-ordered body of one end-to-end step (also what gets captured in a CUDA graph)."""
        if self.host_fed:
            # two branches: the fused kernel starts at once on the main stream and sweeps the
            # interior, while the side stream uploads the boundary values straight into the kernel's
            # inbox and then publishes their flags; the kernel's boundary tiles (processed last) pick
            # them up like a neighbour's.  The upload is hidden behind ~95% of the sweep.
            main = torch.cuda.current_stream()
            side = self._side_stream
            side.wait_stream(main)
            with torch.cuda.stream(side):
                for h, box in self._fed:
                    box.copy_(h, non_blocking=True)
                for r in self.ranks:
                    self.plan.publish_host_sides_by_copy(r)   # copy engine: needs no SM slot
            self.plan.step()
            main.wait_stream(side)
            for k, r in enumerate(self.ranks):   # straight from the kernel's result word to pinned memory
                self._err_host[k:k + 1].copy_(self.err_sumsq(r), non_blocking=True)
            return
        if self._bc_total:
            self._dev_bc.copy_(self._host_bc, non_blocking=True)          # one H2D for all sides
            for (r, side), (off, v) in self._bc_views.items():
                local.copy2d(self._dev_bc[off:off + v.numel()].view(v.shape), v)  # scatter into the ghosts
        self.step()
        for k, r in enumerate(self.ranks):
            self._err_dev[k:k + 1].copy_(self.err_sumsq(r))
        self._err_host.copy_(self._err_dev, non_blocking=True)             # one D2H

    def step_e2e(self, use_graph: bool = True) -> float:
        """What a time-stepping user calls: upload this step's boundary values from pinned host
        memory, run the fused step, read the error norm back.  Returns the largest err_norm over
        the ranks hosted by this process (host float: synchronises every call).  The stream work
        of a step is captured once in a CUDA graph (a single-rank plan launches no cooperative
        kernel, so it is capturable) and replayed: one launch per step instead of seven."""
        if self._host_bc is None:
            self._prepare_e2e()
        cuda = self.device.type == "cuda"
        graphable = (use_graph and cuda and self.impl == "fused" and len(self.ranks) == 1
                     and not self._graph_failed)
        if graphable and self._graph is None:
            try:
                self._e2e_body()                      # warm-up outside capture
                torch.cuda.synchronize(self.device)
                g = torch.cuda.CUDAGraph()
                steps_before = self.plan.steps
                with torch.cuda.graph(g):
                    self._e2e_body()
                self.plan.steps = steps_before   # capture launched nothing: the device epoch did not move
                self._graph = g
            except Exception:  # noqa: BLE001 - capture is an optimisation, never a requirement
                self._graph_failed = True
                self._graph = None
                torch.cuda.synchronize(self.device)
        if graphable and self._graph is not None:
            if self.host_fed:
                self.plan.refresh_host_epoch()
            self._graph.replay()
            self.plan.steps += 1
        else:
            self._e2e_body()
        if cuda:
            torch.cuda.current_stream().synchronize()
        if len(self.ranks) == 1:
            return float(self._err_host[0]) ** 0.5
        return float(self._err_host.max().item()) ** 0.5

    # ------------------------------------------------------------------ validation
    def analytic(self, rank: int) -> torch.Tensor:
        return self.plan.analytic(rank)

    def err_sumsq(self, rank: int) -> torch.Tensor:
        """Device scalar sum((numeric-analytic)^2): from the fused kernel's epilogue when the plan
        computes it, else from the device reduction kernel."""
        if self.impl == "fused" and self.with_err:
            return self.plan.err_sumsq(rank)
        return local.sumsq_diff(self.out[rank], self.analytic(rank))

    def err_norm(self, rank: int) -> float:
        return float(self.err_sumsq(rank).item()) ** 0.5

    def check(self) -> None:
        if hasattr(self.plan, "check_status"):
            self.plan.check_status()
