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


CASES = {"fused": case_fused, "baseline": case_baseline, "collectives": case_collectives}

if __name__ == "__main__":
    name = sys.argv[1]
    rt = runtime.init()
    try:
        CASES[name](rt, sys.argv[2:])
    finally:
        runtime.finalize()
