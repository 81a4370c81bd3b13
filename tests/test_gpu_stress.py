"""Protocol stress: many iterations of the fused
exchange with pseudo-random delays injected before every flag store, the field changing
every step, checked against the neighbours' interiors -- on virtual ranks in one cooperative
launch, so every flag, epoch and parity path is exercised on a single GPU."""
import pytest
import torch

from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.ops._lib import lib
from gpu_mpi_tests_b200.parallel.halo import FusedStencil1D, FusedStencil2D
from gpu_mpi_tests_b200.parallel.symm import VirtualSpace

pytestmark = pytest.mark.gpu
F64 = torch.float64


@pytest.fixture
def delays(native):
    lib.gmt_debug_set_push_delay(40_000)   # up to 40 us before each flag
    yield
    lib.gmt_debug_set_push_delay(0)


def test_stress_2d_grid_with_injected_delays(dev, delays):
    px, py, nx, ny = 4, 2, 2048, 384
    plan = FusedStencil2D(VirtualSpace(px * py, dev), px, py, nx, ny, 2, 2, with_err=False, spin_limit=1 << 26)
    plan.init_analytic(poison=1e30)
    steps = 120
    for it in range(steps):
        for r in plan.ranks:   # the field changes every step (stream-ordered with the kernels)
            plan.z[r][2:-2, 2:-2].mul_(1.0 + 1e-3 * ((it + r) % 5))
        plan.step()
        if it % 40 == 39 or it == steps - 1:
            torch.cuda.synchronize()
            plan.check_status()
            for r in plan.ranks:
                g = plan.grids[r]
                z = plan.z[r]
                for side, (mine, theirs) in {0: (z[2:-2, :2], lambda o: o[2:-2, -4:-2]), 1: (z[2:-2, -2:], lambda o: o[2:-2, 2:4]),
                                             2: (z[:2, 2:-2], lambda o: o[-4:-2, 2:-2]), 3: (z[-2:, 2:-2], lambda o: o[2:4, 2:-2])}.items():
                    nb = g.neighbour(side)
                    if nb >= 0:
                        assert torch.equal(mine, theirs(plan.z[nb])), f"step {it} rank {r} side {side}"
                ref = local.stencil2d_xy(z, plan.sx, plan.sy)
                assert float((plan.out[r] - ref).abs().max()) <= 1e-10 * (float(ref.abs().max()) + 1.0)
    assert plan.epochs() == [steps] * (px * py)


@pytest.mark.parametrize("land", [True, False])
@pytest.mark.parametrize("gx,gy,px,py,nx,ny", [(2, 0, 4, 1, 1024, 4096), (0, 2, 1, 4, 4096, 512)])
def test_stress_2d_slabs_with_injected_delays(dev, delays, gx, gy, px, py, nx, ny, land):
    if not land and gy:
        pytest.skip("land_ghosts=False exists for dim-0 slabs only")
    plan = FusedStencil2D(VirtualSpace(px * py, dev), px, py, nx, ny, gx, gy, with_err=True, spin_limit=1 << 26,
                          land_ghosts=land)
    plan.init_analytic(poison=1e30)
    for _ in range(60):
        plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    for r in plan.ranks:
        assert float((plan.out[r] - plan.analytic(r)).abs().max()) < 1e-6


def test_stress_four_live_faces_and_exchange_only_with_injected_delays(dev, delays):
    """3 x 3 grid: the centre rank has four live faces.  Fused steps and exchange-only calls are mixed
    on the same plan with delayed flags, the field changing before every call."""
    px = py = 3
    nx, ny = 1028, 300
    plan = FusedStencil2D(VirtualSpace(px * py, dev), px, py, nx, ny, 2, 2, with_err=False, spin_limit=1 << 26)
    plan.init_analytic(poison=1e30)
    calls = 60
    for it in range(calls):
        for r in plan.ranks:
            plan.z[r][2:-2, 2:-2].mul_(1.0 + 1e-3 * ((it + r) % 5))
        if it % 3 == 1:
            plan.exchange()
        else:
            plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [calls] * (px * py)
    c = plan.grids[4]
    assert all(c.neighbour(s) >= 0 for s in range(4))
    for r in plan.ranks:
        g, z = plan.grids[r], plan.z[r]
        for side, (mine, theirs) in {0: (z[2:-2, :2], lambda o: o[2:-2, -4:-2]), 1: (z[2:-2, -2:], lambda o: o[2:-2, 2:4]),
                                     2: (z[:2, 2:-2], lambda o: o[-4:-2, 2:-2]), 3: (z[-2:, 2:-2], lambda o: o[2:4, 2:-2])}.items():
            nb = g.neighbour(side)
            if nb >= 0:
                assert torch.equal(mine, theirs(plan.z[nb])), f"rank {r} side {side}"
        ref = local.stencil2d_xy(z, plan.sx, plan.sy)
        assert float((plan.out[r] - ref).abs().max()) <= 1e-10 * (float(ref.abs().max()) + 1.0)


def test_stress_1d_with_injected_delays(dev, delays):
    world, n = 8, 1 << 16
    plan = FusedStencil1D(VirtualSpace(world, dev), n, with_err=False, spin_limit=1 << 26)
    plan.init_analytic()
    for it in range(200):
        for r in plan.ranks:
            plan.z[r][2:-2].mul_(1.0 + 1e-4 * ((it + r) % 3))
        plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    for r in plan.ranks:
        if r > 0:
            assert torch.equal(plan.z[r][:2], plan.z[r - 1][-4:-2])
        if r < world - 1:
            assert torch.equal(plan.z[r][-2:], plan.z[r + 1][2:4])
    assert plan.epochs() == [200] * world
