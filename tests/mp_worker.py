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


CASES = {"fused": case_fused, "baseline": case_baseline}

if __name__ == "__main__":
    name = sys.argv[1]
    rt = runtime.init()
    try:
        CASES[name](rt, sys.argv[2:])
    finally:
        runtime.finalize()
