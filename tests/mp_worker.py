"""Worker launched by torchrun from the multi-process tests (GPU: NCCL + peer memory;
CPU: gloo + baseline exchange).  Usage: mp_worker.py <case> [args...]; exits non-zero on failure."""
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from gpu_mpi_tests_b200.parallel import runtime


def case_fused(rt, args):
    from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
    from gpu_mpi_tests_b200.parallel.halo import FusedStencil1D
    from gpu_mpi_tests_b200.parallel.symm import make_space
    space = make_space(rt)
    if rt.rank == 0:
        print("peer space:", space.kind, "multicast:", space.has_multicast, flush=True)
    p1 = FusedStencil1D(space, 1 << 20, with_err=True, spin_limit=1 << 26)
    p1.init_analytic()
    r = rt.rank
    if r > 0:
        p1.z[r][:2] = 1e30
    if r < rt.world_size - 1:
        p1.z[r][-2:] = 1e30
    for _ in range(7):
        p1.step()
    torch.cuda.synchronize()
    p1.check_status()
    e = p1.err_norm(r)
    assert e < 1e-2, f"1-D err_norm {e}"
    for mode, nx, ny in (("grid", 2048, 1024), ("dim0", 1024, 4096), ("dim1", 4096, 512)):
        wl = Stencil2DWorkload(rt, nx, ny, mode=mode, impl="fused", with_err=True, space=space, spin_limit=1 << 26)
        wl.setup(poison=1e30)
        for _ in range(6):
            wl.step()
        torch.cuda.synchronize()
        wl.check()
        e = wl.err_norm(r)
        ref = float((wl.out[r] - wl.analytic(r)).abs().max())
        assert e < 1e-3 and ref < 1e-6, f"{mode}: err_norm {e} max {ref}"
        v = wl.step_e2e()
        assert v < 1e-3, f"{mode}: e2e err {v}"
    rt.barrier()
    if rt.rank == 0:
        print("fused ok", flush=True)


def case_baseline(rt, args):
    """Library-only exchange (NCCL or gloo) with every staging knob; also the CPU path."""
    from gpu_mpi_tests_b200.baseline.exchange import BaselineStencil1D
    from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
    from gpu_mpi_tests_b200.ops import local
    r = rt.rank
    b1 = BaselineStencil1D(rt, 1 << 14)
    b1.init_analytic()
    b1.step()
    e = local.diff_norm(b1.out[r], b1.analytic(r))
    assert e < 1e-4, f"baseline 1-D {e}"
    combos = [("dim0", "device", "device"), ("dim0", "host", "device"), ("dim1", "device", "device"),
              ("dim1", "none", "device"), ("grid", "device", "device")]
    if rt.is_cuda:
        combos += [("dim0", "device", "managed"), ("dim1", "none", "managed")]
    for mode, stage, mem in combos:
        wl = Stencil2DWorkload(rt, 64, 48, mode=mode, impl="baseline", stage=stage, mem_space=mem)
        wl.setup(poison=1e30)
        wl.step()
        wl.step()
        e = wl.err_norm(r)
        assert e < 1e-6, f"baseline {mode}/{stage}/{mem}: {e}"
    rt.barrier()
    if rt.rank == 0:
        print("baseline ok", flush=True)


def case_collectives(rt, args):
    """The in-kernel collectives across real GPUs (P2P and, when available, NVLS multicast)."""
    from gpu_mpi_tests_b200.parallel.collectives import PeerCollectives
    from gpu_mpi_tests_b200.parallel.symm import make_space
    from gpu_mpi_tests_b200.utils import oracle
    space = make_space(rt)
    W, r, dev = rt.world_size, rt.rank, rt.device
    n = 1 << 18
    pc = PeerCollectives(space, buf_doubles=W * n, stage_cap=1 << 16, spin_limit=1 << 26)
    variants = [False] + ([True] if pc.has_multicast else [])
    for mc in variants:
        for it in range(3):
            b = pc.buffer(r)
            b.fill_(-1.0)
            b[r * n:(r + 1) * n] = oracle.gather_slot_values(r + 1, n, dev) + it
            pc.allgather_inplace(n, multicast=mc)
            torch.cuda.synchronize()
            for p in range(W):
                ref = oracle.gather_slot_values(p + 1, n, dev) + it
                assert torch.equal(b[p * n:(p + 1) * n], ref), f"allgather mc={mc} it={it} slot {p}"
            rt.barrier()
    b = pc.buffer(r)
    b.fill_(-1.0)
    b[r * n:(r + 1) * n] = float(r + 1)
    pc.gather_inplace(n, 0)
    torch.cuda.synchronize()
    rt.barrier()
    if r == 0:
        for p in range(W):
            assert float(b[p * n]) == p + 1 and float(b[(p + 1) * n - 1]) == p + 1
    algos = ["ll", "one_shot", "two_shot"] + (["multimem"] if pc.has_multicast else [])
    for algo in algos:
        for m in (1, 1024, 1 << 16):
            for it in range(3):
                v = torch.full((m,), float(r + 1 + it), dtype=torch.float64, device=dev)
                pc.allreduce({r: v}, algo=algo)
                torch.cuda.synchronize()
                expect = sum(p + 1 + it for p in range(W))
                assert float((v - expect).abs().max()) < 1e-12, f"allreduce {algo} m={m}"
    x, y = oracle.daxpy_init_scaled(1 << 20, dev)
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    pc.daxpy_allreduce(2.0, {r: x}, {r: y}, {r: out})
    torch.cuda.synchronize()
    expect = oracle.daxpy_scaled_expected_allsum(1 << 20, W)
    assert abs(out.item() - expect) < 1e-6 * expect, (out.item(), expect)
    z = torch.full((64, 128), float(r + 1), dtype=torch.float64, device=dev)
    o = torch.zeros(64, dtype=torch.float64, device=dev)
    pc.axis_sum_allreduce({r: z}, 0, {r: o})
    torch.cuda.synchronize()
    assert float((o - 128 * sum(p + 1 for p in range(W))).abs().max()) < 1e-9
    for _ in range(10):
        pc.barrier()
    torch.cuda.synchronize()
    pc.check_status()
    rt.barrier()
    if r == 0:
        print("collectives ok (multicast: %s)" % pc.has_multicast, flush=True)


def case_darray(rt, args):
    """A user's own compute step on distributed arrays: arbitrary data (a function of the GLOBAL
    indices, nothing analytic about it), DistArray.exchange() / .apply(fn) with fn a torch expression
    that reaches two cells into the ghosts, checked against the same expression on the global field.
    Runs on CPUs (library exchange over gloo) and on GPUs (peer-memory exchange kernel) alike."""
    from gpu_mpi_tests_b200.parallel import decomp
    from gpu_mpi_tests_b200.parallel.darray import DistArray1D, DistArray2D
    r, W = rt.rank, rt.world_size
    dev = rt.device
    f64 = torch.float64

    def field(gi, gj):   # any function of the global indices
        return torch.sin(0.37 * gi)[None, :] * torch.cos(0.11 * gj)[:, None] + 1e-3 * gi[None, :] + 2e-3 * gj[:, None]

    def wide_cross(z, gx, gy, nx, ny):   # reaches 2 cells in every decomposed direction
        c = z[gy:gy + ny, gx:gx + nx]
        out = -4.0 * c
        if gx:
            out = out + z[gy:gy + ny, 0:nx] + z[gy:gy + ny, 2 * gx:2 * gx + nx]
        if gy:
            out = out + z[0:ny, gx:gx + nx] + z[2 * gy:2 * gy + ny, gx:gx + nx]
        return out

    nx, ny = 64, 48
    for dims in ((0,), (1,), (0, 1)):
        if dims == (0, 1):
            px, py = decomp.best_grid(W)
        else:
            px, py = (W, 1) if dims == (0,) else (1, W)
        a = DistArray2D(rt, nx, ny, dims=dims)
        gx, gy = a.gx, a.gy
        cx, cy = r % px, r // px
        gi = torch.arange(-gx, nx + gx, dtype=f64, device=dev) + cx * nx     # global i of every local column incl. ghosts
        gj = torch.arange(-gy, ny + gy, dtype=f64, device=dev) + cy * ny
        full = field(gi, gj)
        a.local.copy_(full)
        # poison the ghosts that face a neighbour: only the exchange may fill them
        g = decomp.Grid2D(px, py, r)
        for side in range(4):
            if (gx if side < 2 else gy) and g.neighbour(side) >= 0:
                a.ghost(side).fill_(1e30)
        got = a.apply(lambda z: wide_cross(z, gx, gy, nx, ny))
        want = wide_cross(full, gx, gy, nx, ny)
        err = float((got - want).abs().max())
        assert err < 1e-12, f"dims {dims}: custom op after exchange differs by {err}"
        assert torch.equal(a.local, full), f"dims {dims}: exchange changed something other than the live ghosts"
    a1 = DistArray1D(rt, 1 << 12)
    n = a1.slab.n_local
    gi = torch.arange(-2, n + 2, dtype=f64, device=dev) + r * n
    full = torch.sin(0.37 * gi) + 1e-3 * gi
    a1.local.copy_(full)
    if r > 0:
        a1.local[:2] = 1e30
    if r < W - 1:
        a1.local[-2:] = 1e30
    got = a1.apply(lambda y: y[:-4] + y[4:] - 2.0 * y[2:-2])
    want = full[:-4] + full[4:] - 2.0 * full[2:-2]
    assert float((got - want).abs().max()) < 1e-12
    rt.barrier()
    if r == 0:
        print("darray ok", flush=True)


CASES = {"darray": case_darray, "fused": case_fused, "baseline": case_baseline, "collectives": case_collectives}

if __name__ == "__main__":
    name = sys.argv[1]
    rt = runtime.init()
    try:
        CASES[name](rt, sys.argv[2:])
    finally:
        runtime.finalize()
