"""In-kernel collectives over peer memory: the product-path replacements for the reference's
MPI_Allgather(MPI_IN_PLACE) (mpi_daxpy_nvtx.cc:285, mpigatherinplace.f90:39), MPI_Allgather
(mpi_daxpy_nvtx.cc:288), MPI_Allreduce(MPI_IN_PLACE) (mpi_stencil2d_gt.cc:615,624),
MPI_Barrier (mpi_daxpy_nvtx.cc:277), plus MPI_Gather-in-place (BASELINE config 5) and the two
fused compute+collective ops (daxpy + global sum, axis-sum + all-reduce).

A `PeerCollectives` object owns, per hosted rank: a symmetric user buffer of `buf_doubles`
doubles (the gathered array lives here), a symmetric staging area for reductions, a flag block
and an epoch counter; every method is one kernel launch per rank (csrc/collectives.cu) with no
host synchronisation.  Under `VirtualSpace` all ranks run in one cooperative launch.
"""
from __future__ import annotations

import torch

from ..ops._lib import CollRank, lib
from .symm import PeerSpace

F64 = torch.float64
FLAG_WORDS = 256
MAX_WORLD = 32


class CollectiveTimeout(RuntimeError):
    pass


def _dev_ptr_array(ptrs, device) -> torch.Tensor:
    return torch.tensor([int(p) for p in ptrs], dtype=torch.int64, device=device)


class PeerCollectives:
    def __init__(self, space: PeerSpace, buf_doubles: int = 0, stage_cap: int = 8192, spin_limit: int = 0):
        if space.world > MAX_WORLD:
            raise ValueError(f"world {space.world} > {MAX_WORLD}")
        self.space, self.device = space, space.device
        self.world, self.ranks = space.world, list(space.ranks_here)
        # at least gmt_coll_min_stage_cap() cells per source rank: daxpy + global sum stores one per CTA
        self.buf_doubles, self.stage_cap = int(buf_doubles), max(int(stage_cap), int(lib.gmt_coll_min_stage_cap()))
        self.spin_limit = int(spin_limit)
        self.ll_gather_max = min(self.stage_cap, int(lib.gmt_coll_ll_gather_max()))
        self._buf = space.alloc(max(self.buf_doubles, 4) * 8)
        self._stage = space.alloc(int(lib.gmt_coll_stage_bytes(space.world, self.stage_cap)))   # vectors + LL cells, zeroed
        self._flags = space.alloc(FLAG_WORDS * 8)
        n = len(self.ranks)
        self._epoch = torch.zeros(n, 8, dtype=torch.int64, device=self.device)
        self._done = torch.zeros(n, 8, dtype=torch.int32, device=self.device)
        self._status = torch.zeros(n, 8, dtype=torch.int32, device=self.device)
        self._scratch = torch.zeros(n, int(lib.gmt_reduce_scratch_bytes()), dtype=torch.uint8, device=self.device)
        self._tab_buf = _dev_ptr_array(self._buf.ptrs, self.device)
        self._tab_stage = _dev_ptr_array(self._stage.ptrs, self.device)
        self._tab_flag = _dev_ptr_array(self._flags.ptrs, self.device)
        self.has_multicast = bool(self._buf.mc_ptr) and bool(self._stage.mc_ptr)
        self._operands = {r: (0, 0, 0) for r in self.ranks}
        self._ranks_dev = None
        self._rebuild()

    # ------------------------------------------------------------------ storage
    def buffer(self, rank: int) -> torch.Tensor:
        """The symmetric user buffer of a hosted rank as a 1-D fp64 tensor."""
        return self._buf.view(rank, 0, self.buf_doubles, F64)

    def _rebuild(self) -> None:
        structs = []
        for k, r in enumerate(self.ranks):
            s = CollRank()
            s.rank, s.world = r, self.world
            s.peer_buf, s.mc_buf = self._tab_buf.data_ptr(), self._buf.mc_ptr
            s.peer_stage, s.mc_stage = self._tab_stage.data_ptr(), self._stage.mc_ptr
            s.stage_cap = self.stage_cap
            s.peer_flag = self._tab_flag.data_ptr()
            s.epoch, s.done, s.status = self._epoch[k].data_ptr(), self._done[k].data_ptr(), self._status[k].data_ptr()
            s.scratch = self._scratch[k].data_ptr()
            s.inp, s.out, s.y = self._operands[r]
            s.spin_limit = self.spin_limit
            structs.append(s)
        raw = b"".join(bytes(s) for s in structs)
        host = torch.frombuffer(bytearray(raw), dtype=torch.uint8)
        if self._ranks_dev is None:
            self._ranks_dev = host.to(self.device)
        else:
            self._ranks_dev.copy_(host, non_blocking=False)

    def bind(self, operands: dict) -> None:
        """Set the local operand pointers {rank: (in, out, y)} (tensors or None).  Re-binding costs
        a small H2D copy; benchmarks bind once and launch many times."""
        new = {}
        for r in self.ranks:
            t = operands.get(r, (None, None, None))
            new[r] = tuple(0 if x is None else int(x.data_ptr()) for x in t)
        if new != self._operands:
            self._operands = new
            self._rebuild()

    def _stream(self) -> int:
        return torch.cuda.current_stream().cuda_stream

    def check_status(self) -> None:
        st = self._status[:, 0].tolist()
        if any(v & 0x1000 for v in st):
            raise CollectiveTimeout(f"a collective needed more LL cells than stage_cap={self.stage_cap}: status {st}")
        if any(st):
            raise CollectiveTimeout(f"collective wait expired: status per hosted rank {st} (1 + silent rank)")

    def epochs(self) -> list:
        return self._epoch[:, 0].tolist()

    # ------------------------------------------------------------------ collectives
    GATHER_MODES = {"p2p": 0, "multimem": 1, "ll": 2, "multimem_weak": 3, "mc_store": 4}

    def _gather_mode(self, n: int, multicast, mode: str | None) -> int:
        """mode: "p2p" (world-1 peer stores per 32 bytes), "ll" (epoch-carrying cells, slices up to
        64 KiB: one one-way trip, no signalling rounds), "multimem" / "multimem_weak" / "mc_store"
        (NVLS: the switch replicates one store); None = ll when the slice fits, else p2p."""
        if mode is None:
            if multicast:
                mode = "multimem"
            else:
                mode = "ll" if (self.world > 1 and n <= self.ll_gather_max) else "p2p"
        m = self.GATHER_MODES[mode]
        if m in (1, 3, 4) and not self.has_multicast:
            m = 0
        if m == 2 and n > self.ll_gather_max:
            raise ValueError(f"LL allgather takes slices of up to {self.ll_gather_max} doubles, got {n}")
        return m

    def allgather_inplace(self, n: int, multicast: bool = False, mode: str | None = None) -> None:
        """Every rank's slot [r*n, (r+1)*n) of buffer(r) ends up in every rank's buffer."""
        self._check_gather(n)
        self.bind({})
        lib.gmt_allgather_f64(self._ranks_dev.data_ptr(), len(self.ranks), n, 0,
                              self._gather_mode(n, multicast, mode), self._stream())

    def allgather(self, n: int, src: dict, multicast: bool = False, mode: str | None = None) -> None:
        """Out-of-place source: slot r of every buffer <- src[r] (n doubles)."""
        self._check_gather(n)
        self.bind({r: (src[r], None, None) for r in self.ranks})
        lib.gmt_allgather_f64(self._ranks_dev.data_ptr(), len(self.ranks), n, 1,
                              self._gather_mode(n, multicast, mode), self._stream())

    def gather_inplace(self, n: int, root: int = 0) -> None:
        self._check_gather(n)
        if not (0 <= root < self.world):
            raise ValueError("bad root")
        self.bind({})
        lib.gmt_gather_f64(self._ranks_dev.data_ptr(), len(self.ranks), n, root, self._stream())

    def _check_gather(self, n: int) -> None:
        if n * self.world > self.buf_doubles:
            raise ValueError(f"buffer of {self.buf_doubles} doubles cannot hold {self.world} slices of {n}")

    ALGOS = {"one_shot": 0, "two_shot": 1, "multimem": 2, "ll": 3}
    LL_MAX = 1 << 18   # doubles x ranks: beyond this the 16-byte-per-double-per-peer cells stop paying
    #                    (2 MiB on 2 GPUs: 15.7 us LL vs 30 us one-shot; on 4 GPUs: 34 vs 33 us)

    def allreduce(self, vec: dict, out: dict | None = None, algo: str = "auto") -> None:
        """out[r] <- sum over ranks of vec[r]; in place when out is None.  algo: "ll" (epoch-carrying
        16-byte cells, no rendezvous), "one_shot", "two_shot", "multimem"; "auto" = ll while
        n * world <= LL_MAX, two_shot beyond 2^17 doubles on 8+ ranks, else one_shot."""
        out = out or vec
        n = next(iter(vec.values())).numel()
        if algo == "auto":
            algo = "ll" if n * self.world <= self.LL_MAX else ("two_shot" if (n > (1 << 17) and self.world >= 8) else "one_shot")
        if n > self.stage_cap:
            raise ValueError(f"all-reduce of {n} doubles exceeds the staging capacity {self.stage_cap}")
        if algo == "multimem" and not self.has_multicast:
            raise RuntimeError("multimem all-reduce needs a multicast-capable peer space")
        self.bind({r: (vec[r], out[r], None) for r in self.ranks})
        lib.gmt_allreduce_f64(self._ranks_dev.data_ptr(), len(self.ranks), n, self.ALGOS[algo], self._stream())

    def daxpy_allreduce(self, a: float, x: dict, y: dict, out: dict) -> None:
        """y[r] <- a*x[r] + y[r]; out[r][0] <- sum over all ranks of sum(y_new)."""
        n = next(iter(x.values())).numel()
        if any(x[r].numel() != n or y[r].numel() != n for r in self.ranks):
            raise ValueError("daxpy_allreduce: every rank must pass the same n")
        self.bind({r: (x[r], out[r], y[r]) for r in self.ranks})
        lib.gmt_daxpy_allreduce_f64(self._ranks_dev.data_ptr(), len(self.ranks), n, float(a), self._stream())

    def axis_sum_allreduce(self, z: dict, axis: int, out: dict, n_out: int | None = None) -> None:
        """Axis sum of each rank's column-major array (torch shape (ny, ld>=nx)) followed by the
        all-reduce of the first n_out entries, one kernel."""
        z0 = next(iter(z.values()))
        ny, nx = z0.shape
        ld = z0.stride(0) if ny > 1 else nx
        n_out = (ny if axis == 0 else nx) if n_out is None else n_out
        if n_out > self.stage_cap:
            raise ValueError(f"all-reduce of {n_out} doubles exceeds the staging capacity {self.stage_cap}")
        self.bind({r: (z[r], out[r], None) for r in self.ranks})
        lib.gmt_axis_sum_allreduce_f64(self._ranks_dev.data_ptr(), len(self.ranks), nx, ny, ld, int(axis), n_out,
                                       self._stream())

    def barrier(self) -> None:
        lib.gmt_barrier(self._ranks_dev.data_ptr(), len(self.ranks), self._stream())
