"""The fused halo-exchange + stencil kernels, all ranks emulated on one GPU in a single
cooperative launch (VirtualSpace): protocol correctness (flags, epochs, parity double
buffering), ghost landing, fused error norm -- against the analytic derivative and against
the un-fused kernels fed with an oracle-exchanged array."""
import pytest
import torch

from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.parallel.halo import FusedStencil1D, FusedStencil2D
from gpu_mpi_tests_b200.parallel.symm import VirtualSpace

pytestmark = pytest.mark.gpu
F64 = torch.float64


def close(a, b, rtol=1e-10):
    """Same arithmetic, separately compiled kernels: equal up to FMA-contraction round-off,
    which the 1/dx scale amplifies -- compare relative to the field's magnitude."""
    return float((a - b).abs().max()) <= rtol * (float(b.abs().max()) + 1.0)


@pytest.mark.parametrize("world,n", [(1, 4096), (2, 64), (3, 1000), (4, 8192), (8, 65536), (5, 4096 * 3 + 2)])
def test_fused_stencil1d(dev, world, n):
    plan = FusedStencil1D(VirtualSpace(world, dev), n, with_err=True, spin_limit=1 << 24)
    plan.init_analytic()
    for r in plan.ranks:  # interior-facing ghosts must come from the neighbour, not from init
        if r > 0:
            plan.z[r][:2] = 1e30
        if r < world - 1:
            plan.z[r][-2:] = 1e30
    steps = 5
    for _ in range(steps):
        plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [steps] * world
    xs = torch.arange(n * world, dtype=F64, device=dev) * plan.dx
    for r in plan.ranks:
        act = 3 * xs[r * n:(r + 1) * n] ** 2
        assert float((plan.out[r] - act).abs().max()) < 1e-6
        assert plan.err_norm(r) == pytest.approx(float(((plan.out[r] - act) ** 2).sum()) ** 0.5, rel=0.1, abs=1e-9)  # both sides are pure round-off
        # ghosts landed in the array: same state as after the reference's boundary_exchange
        if r > 0:
            assert torch.equal(plan.z[r][:2], plan.z[r - 1][-4:-2])
        if r < world - 1:
            assert torch.equal(plan.z[r][-2:], plan.z[r + 1][2:4])
        # identical to the plain kernel on the exchanged array
        # (not bit-identical: the two kernels are compiled separately and may contract FMAs differently)
        assert close(plan.out[r], local.stencil1d(plan.z[r], plan.scale))


@pytest.mark.parametrize("world,n", [(1, 4096), (2, 64), (3, 1000), (4, 8192), (8, 65536), (5, 4096 * 3 + 2), (2, 1 << 20)])
def test_fused_stencil1d_tma_variant(dev, world, n):
    """The fused 1-D step on the TMA data path (cp.async.bulk ring): same result as the LSU kernel,
    ghosts landed, and the two variants can be mixed on one plan (same inboxes / flags / epoch)."""
    plan = FusedStencil1D(VirtualSpace(world, dev), n, with_err=False, spin_limit=1 << 24, variant="tma")
    plan.init_analytic()
    for r in plan.ranks:
        if r > 0:
            plan.z[r][:2] = 1e30
        if r < world - 1:
            plan.z[r][-2:] = 1e30
    for it in range(4):
        for r in plan.ranks:
            plan.z[r][2:-2].mul_(1.0 + 1e-4 * ((it + r) % 3))
        plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [4] * world
    for r in plan.ranks:
        if r > 0:
            assert torch.equal(plan.z[r][:2], plan.z[r - 1][-4:-2])
        if r < world - 1:
            assert torch.equal(plan.z[r][-2:], plan.z[r + 1][2:4])
        assert close(plan.out[r], local.stencil1d(plan.z[r], plan.scale))
    plan.variant = "vector"      # same plan, LSU kernel: the epoch and the inbox parities carry over
    plan.step()
    plan.variant = "tma"
    plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [6] * world
    for r in plan.ranks:
        assert close(plan.out[r], local.stencil1d(plan.z[r], plan.scale))


CASES_2D = [
    # px, py, nx, ny, gx, gy
    (1, 1, 2048, 512, 2, 2),
    (4, 1, 1024, 256, 2, 0),     # reference dim-0 slab test
    (2, 1, 4096, 100, 2, 0),     # several row tiles per column
    (3, 1, 64, 3000, 2, 0),      # several flag chunks per face
    (1, 4, 2048, 256, 0, 2),     # reference dim-1 slab test
    (1, 3, 1024, 130, 0, 2),
    (2, 2, 2048, 256, 2, 2),     # BASELINE config 3 shape: 4 neighbours
    (4, 2, 1024, 384, 2, 2),
    (3, 2, 36, 50, 2, 2),
    (3, 3, 1028, 300, 2, 2),     # the centre rank has FOUR live faces
    (3, 4, 2048, 130, 2, 2),     # two ranks with four live faces, several row tiles
    (2, 1, 64, 66000, 2, 0),     # 65 x-face chunks per side
]


@pytest.mark.parametrize("px,py,nx,ny,gx,gy", CASES_2D)
def test_fused_stencil2d(dev, px, py, nx, ny, gx, gy):
    plan = FusedStencil2D(VirtualSpace(px * py, dev), px, py, nx, ny, gx, gy, with_err=True, spin_limit=1 << 24)
    plan.init_analytic(poison=1e30)
    steps = 4
    for _ in range(steps):
        plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [steps] * (px * py)
    for r in plan.ranks:
        act = plan.analytic(r)
        assert float((plan.out[r] - act).abs().max()) < 1e-6, f"rank {r}"
        assert plan.err_norm(r) == pytest.approx(float(((plan.out[r] - act) ** 2).sum()) ** 0.5, rel=0.1, abs=1e-9)  # both sides are pure round-off
        z = plan.z[r]
        if gx and gy:
            ref = local.stencil2d_xy(z, plan.sx, plan.sy)
        elif gx:
            ref = local.stencil2d(z, 0, plan.sx)
        else:
            ref = local.stencil2d(z, 1, plan.sy)
        assert close(plan.out[r], ref)


@pytest.mark.parametrize("px,nx,ny", [(4, 1024, 256), (2, 4096, 100), (3, 64, 3000), (2, 64, 66000), (1, 1024, 64)])
def test_fused_dim0_ghosts_straight_from_the_inbox(dev, px, nx, ny, monkeypatch):
    """land_ghosts=False (dim-0 slabs): the inbox is the ghost storage.  Same `out` as the landing
    flavour, z's own ghost cells stay untouched (still poisoned), and exchange() fills them on demand."""
    if ny >= 60000:
        monkeypatch.setenv("GMT_PUSH_GROUP", "4")     # the tuning knob: several chunks per fence
        monkeypatch.setenv("GMT_PUSH_CTAS", "16")
    plan = FusedStencil2D(VirtualSpace(px, dev), px, 1, nx, ny, 2, 0, with_err=True, spin_limit=1 << 24,
                          land_ghosts=False)
    ref = FusedStencil2D(VirtualSpace(px, dev), px, 1, nx, ny, 2, 0, with_err=True, spin_limit=1 << 24)
    assert not plan.land_ghosts and ref.land_ghosts
    plan.init_analytic(poison=1e30)
    ref.init_analytic(poison=1e30)
    for it in range(4):
        for r in plan.ranks:     # the field changes every step: both inbox parities carry fresh values
            plan.z[r][:, 2:-2].mul_(1.0 + 1e-3 * (it + r))
            ref.z[r][:, 2:-2].mul_(1.0 + 1e-3 * (it + r))
        plan.step()
        ref.step()
    torch.cuda.synchronize()
    plan.check_status()
    ref.check_status()
    for r in plan.ranks:
        assert close(plan.out[r], ref.out[r]), f"rank {r}"            # same arithmetic on the same values
        assert float((plan.out[r] - ref.out[r]).abs().max()) < 1e20    # no poisoned ghost leaked into a result
        assert plan.err_norm(r) == pytest.approx(ref.err_norm(r), rel=1e-3, abs=1e-9)
        if r > 0:
            assert float(plan.z[r][:, :2].min()) == 1e30              # not landed
        if r < px - 1:
            assert float(plan.z[r][:, -2:].min()) == 1e30
    plan.exchange()
    torch.cuda.synchronize()
    plan.check_status()
    for r in plan.ranks:
        assert torch.equal(plan.z[r], ref.z[r])


def _ghost_pairs(plan, r):
    """(ghost view of rank r, matching boundary view of the neighbour) for every live side."""
    z, gx, gy, nx, ny = plan.z[r], plan.gx, plan.gy, plan.nx, plan.ny
    g = plan.grids[r]
    ys, xs = slice(gy, gy + ny), slice(gx, gx + nx)
    out = []
    for side in range(4):
        if not (gx if side < 2 else gy):
            continue
        nb = g.neighbour(side)
        if nb < 0:
            continue
        o = plan.z[nb]
        if side == 0:
            out.append((side, z[ys, 0:2], o[ys, nx:nx + 2]))
        elif side == 1:
            out.append((side, z[ys, gx + nx:gx + nx + 2], o[ys, gx:gx + 2]))
        elif side == 2:
            out.append((side, z[0:2, xs], o[ny:ny + 2, xs]))
        else:
            out.append((side, z[gy + ny:gy + ny + 2, xs], o[gy:gy + 2, xs]))
    return out


EXCH_2D = CASES_2D + [
    (2, 2, 16384, 16384 // 8, 2, 2),   # flagship tile width, many row tiles per face
    (1, 1, 1024, 256, 2, 2),
]


@pytest.mark.parametrize("px,py,nx,ny,gx,gy", EXCH_2D)
@pytest.mark.parametrize("periodic", [False, True])
def test_halo_exchange2d_alone(dev, px, py, nx, ny, gx, gy, periodic):
    """The exchange as a callable of its own (boundary_exchange_x/_y, mpi_stencil2d_gt.cc:136-373):
    poisoned ghosts become the neighbours' boundary values, nothing else in the array and nothing
    in `out` changes, and it can be mixed with fused steps on the same plan."""
    if periodic and nx * ny > (1 << 22):
        pytest.skip("one big case is enough")
    plan = FusedStencil2D(VirtualSpace(px * py, dev), px, py, nx, ny, gx, gy, with_err=False, spin_limit=1 << 24,
                          periodic=periodic)
    plan.init_analytic(poison=1e30)
    for r in plan.ranks:
        plan.out[r].fill_(-7.0)
    before = {r: plan.z[r].clone() for r in plan.ranks}
    for it in range(3):
        for r in plan.ranks:   # the field changes between calls: both inbox parities are exercised
            plan.z[r][gy:gy + ny, gx:gx + nx].add_(float(it + r))
            before[r][gy:gy + ny, gx:gx + nx].add_(float(it + r))
        plan.exchange()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [3] * (px * py)
    for r in plan.ranks:
        assert float(plan.out[r].min()) == float(plan.out[r].max()) == -7.0     # no sweep happened
        pairs = _ghost_pairs(plan, r)
        if periodic:
            assert len(pairs) == 2 * (bool(gx) + bool(gy))
        for side, mine, theirs in pairs:
            assert torch.equal(mine, theirs), f"rank {r} side {side}"
        # everything that is not a live ghost cell is untouched
        z, b = plan.z[r], before[r]
        mask = torch.ones_like(z, dtype=torch.bool)
        g = plan.grids[r]
        if gx and g.neighbour(0) >= 0:
            mask[gy:gy + ny, 0:2] = False
        if gx and g.neighbour(1) >= 0:
            mask[gy:gy + ny, gx + nx:] = False
        if gy and g.neighbour(2) >= 0:
            mask[0:2, gx:gx + nx] = False
        if gy and g.neighbour(3) >= 0:
            mask[gy + ny:, gx:gx + nx] = False
        assert torch.equal(z[mask], b[mask]), f"rank {r}: exchange touched cells outside the live ghost strips"
    # mixed with a fused step on the same plan (same epoch counter and inbox parities)
    plan.step()
    plan.exchange()
    plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [6] * (px * py)
    if not periodic:
        for r in plan.ranks:
            ref = (local.stencil2d_xy(plan.z[r], plan.sx, plan.sy) if gx and gy else
                   local.stencil2d(plan.z[r], 0 if gx else 1, plan.sx if gx else plan.sy))
            assert close(plan.out[r], ref)


@pytest.mark.parametrize("world,n", [(1, 4096), (2, 64), (5, 4096 * 3 + 2), (8, 65536)])
def test_halo_exchange1d_alone(dev, world, n):
    """boundary_exchange (mpi_stencil_gt.cc:83-122) as a callable: 16 bytes per side."""
    plan = FusedStencil1D(VirtualSpace(world, dev), n, with_err=False, spin_limit=1 << 24)
    plan.init_analytic()
    for r in plan.ranks:
        plan.out[r].fill_(-7.0)
        if r > 0:
            plan.z[r][:2] = 1e30
        if r < world - 1:
            plan.z[r][-2:] = 1e30
    for it in range(4):
        for r in plan.ranks:
            plan.z[r][2:-2].add_(float(it + r))
        plan.exchange()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [4] * world
    for r in plan.ranks:
        assert float(plan.out[r].min()) == -7.0
        if r > 0:
            assert torch.equal(plan.z[r][:2], plan.z[r - 1][-4:-2])
        if r < world - 1:
            assert torch.equal(plan.z[r][-2:], plan.z[r + 1][2:4])
    plan.step()
    plan.exchange()
    torch.cuda.synchronize()
    plan.check_status()
    assert plan.epochs() == [6] * world


def test_periodic_grid_gives_every_rank_four_live_faces(dev):
    """Grid2D(periodic=True): a 4 x 2 grid (the 8-GPU shape) exchanges all four faces on every rank;
    ghosts equal the wrapped neighbours' boundary values after fused steps with a changing field."""
    px, py, nx, ny = 4, 2, 1024, 256
    plan = FusedStencil2D(VirtualSpace(px * py, dev), px, py, nx, ny, 2, 2, with_err=False, spin_limit=1 << 24,
                          periodic=True)
    plan.init_analytic(poison=1e30)
    assert all(plan.halo_bytes_out(r) == 2 * 8 * 2 * (nx + ny) for r in plan.ranks)
    for it in range(4):
        for r in plan.ranks:
            plan.z[r][2:-2, 2:-2].mul_(1.0 + 1e-3 * ((it + r) % 5))
        plan.step()
    torch.cuda.synchronize()
    plan.check_status()
    for r in plan.ranks:
        pairs = _ghost_pairs(plan, r)
        assert len(pairs) == 4
        for side, mine, theirs in pairs:
            assert torch.equal(mine, theirs), f"rank {r} side {side}"
        assert close(plan.out[r], local.stencil2d_xy(plan.z[r], plan.sx, plan.sy))


def test_fused_values_change_between_steps(dev):
    """Parity double buffering: the array changes every step; ghosts must track it."""
    plan = FusedStencil2D(VirtualSpace(4, dev), 2, 2, 1024, 128, 2, 2, with_err=False, spin_limit=1 << 24)
    plan.init_analytic(poison=1e30)
    for it in range(6):
        for r in plan.ranks:
            plan.z[r][2:-2, 2:-2] += float(it + 1)  # interior only; ghosts refreshed by the exchange
        plan.step()
        torch.cuda.synchronize()
        for r in plan.ranks:
            g = plan.grids[r]
            if g.neighbour(0) >= 0:
                assert torch.equal(plan.z[r][2:-2, :2], plan.z[g.neighbour(0)][2:-2, -4:-2])
            if g.neighbour(3) >= 0:
                assert torch.equal(plan.z[r][-2:, 2:-2], plan.z[g.neighbour(3)][2:4, 2:-2])
    plan.check_status()


def test_fused_without_err_and_world_of_one_matches_plain(dev):
    plan = FusedStencil2D(VirtualSpace(1, dev), 1, 1, 4096, 300, 2, 2, with_err=False)
    plan.init_analytic()
    plan.step()
    torch.cuda.synchronize()
    assert close(plan.out[0], local.stencil2d_xy(plan.z[0], plan.sx, plan.sy))


def test_dead_neighbour_times_out_instead_of_hanging(dev):
    """Rank 0 believes it has a high neighbour that never launches: the bounded spin must expire
    and be reported, not hang the GPU."""
    from gpu_mpi_tests_b200.parallel.halo import HaloTimeout
    plan = FusedStencil1D(VirtualSpace(2, dev), 8192, with_err=False, spin_limit=2000)
    plan.init_analytic()
    # launch only rank 0's entry: its neighbour's flag never arrives
    from gpu_mpi_tests_b200.ops._lib import lib
    lib.gmt_fused_stencil1d_f64(plan._ranks_dev.data_ptr(), 1, plan.n, 0, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    with pytest.raises(HaloTimeout):
        plan.check_status()


@pytest.mark.parametrize("px,py", [(1, 1), (2, 2)])
def test_host_fed_sides_behave_like_a_neighbour(dev, px, py):
    """Physical-boundary ghosts delivered by the host while the kernel is already running
    (has == 2): same result as ghosts that were in the array beforehand, and new host values
    show up in the array's ghost cells after the step."""
    from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
    from gpu_mpi_tests_b200.parallel import runtime
    rt = runtime.init()
    nx, ny = 2048, 384
    ref = Stencil2DWorkload(rt, nx, ny, mode="grid", impl="fused", space=VirtualSpace(px * py, dev), grid=(px, py),
                            spin_limit=1 << 26)
    fed = Stencil2DWorkload(rt, nx, ny, mode="grid", impl="fused", space=VirtualSpace(px * py, dev), grid=(px, py),
                            spin_limit=1 << 26, host_fed=True)
    ref.setup()
    fed.setup(poison=0.0)
    for r in fed.ranks:  # the in-array physical ghosts must NOT be what makes the result right
        for (rr, side) in fed.plan.host_sides:
            if rr == r:
                fed.plan.ghost_view(r, side).fill_(1e30)
    for _ in range(3):
        ref.step()
        fed.step()
    e_ref = ref.step_e2e()
    e_fed = fed.step_e2e()
    e_fed = fed.step_e2e()  # second call replays the captured graph (single hosted rank) or re-runs eagerly
    torch.cuda.synchronize()
    fed.check()
    assert e_fed == pytest.approx(e_ref, rel=1e-6, abs=1e-9)
    for r in fed.ranks:
        assert close(fed.out[r], ref.out[r])
        for (rr, side) in fed.plan.host_sides:
            if rr == r:  # the kernel landed the host values in the array's own ghost cells
                assert torch.equal(fed.plan.ghost_view(r, side), ref.plan.ghost_view(r, side))
    # new boundary values from the host reach the array
    for h in fed.host_inputs(0):
        h.add_(1.0)
    fed.step_e2e()
    torch.cuda.synchronize()
    r0 = fed.ranks[0]
    side0 = [s for (rr, s) in fed.plan.host_sides if rr == r0][0]
    assert torch.allclose(fed.plan.ghost_view(r0, side0), ref.plan.ghost_view(r0, side0) + 1.0)


def test_pipelined_e2e_lags_one_step_and_tracks_new_inputs(dev):
    """step_e2e_pipelined(): the result returned by call k is step k-1's; inputs written into the
    buffers host_inputs() hands out reach exactly the step they were written for (two alternating
    sets of pinned buffers / flag sources / result words)."""
    from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
    from gpu_mpi_tests_b200.parallel import runtime
    rt = runtime.init()
    nx, ny = 2048, 384
    sync = Stencil2DWorkload(rt, nx, ny, mode="grid", impl="fused", space=VirtualSpace(1, dev), grid=(1, 1),
                             spin_limit=1 << 26, host_fed=True)
    pipe = Stencil2DWorkload(rt, nx, ny, mode="grid", impl="fused", space=VirtualSpace(1, dev), grid=(1, 1),
                             spin_limit=1 << 26, host_fed=True)
    sync.setup()
    pipe.setup()
    base = [h.clone() for h in pipe.host_inputs(0)]
    want, got = [], []
    for k in range(7):
        bump = 0.125 * k                      # a different boundary value every step: err_norm changes with it
        for h, b in zip(sync.host_inputs(0), base):
            h.copy_(b + bump)
        want.append(sync.step_e2e())
        for h, b in zip(pipe.host_inputs(), base):
            h.copy_(b + bump)
        got.append(pipe.step_e2e_pipelined())
    got.append(pipe.e2e_flush())
    torch.cuda.synchronize()
    pipe.check()
    assert got[0] is None
    assert got[1:] == pytest.approx(want, rel=1e-6, abs=1e-12)
    assert want[-1] > 10 * want[0]            # the bumped boundary values really reached the kernel
    assert pipe.plan.epochs() == [pipe.plan.steps]


@pytest.mark.parametrize("oneshot", ["0", "1"])
def test_host_fed_on_a_full_gpu(dev, monkeypatch, oneshot):
    """The host-fed flags must arrive although the fused kernel's CTAs occupy every SM slot
    (persistent mode never frees one before the end): they are published by a copy engine, not by
    a kernel that would need a slot.  A starved publisher would show up as HaloTimeout."""
    from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
    from gpu_mpi_tests_b200.parallel import runtime
    monkeypatch.setenv("GMT_FUSED_ONESHOT", oneshot)
    rt = runtime.init()
    n = 8192
    fed = Stencil2DWorkload(rt, n, n, mode="grid", impl="fused", space=VirtualSpace(1, dev), grid=(1, 1),
                            spin_limit=1 << 24, host_fed=True, with_err=True)
    fed.setup()
    fed.e2e_bytes()   # builds the pinned mirror of the boundary values from the (still analytic) ghosts
    for (rr, side) in fed.plan.host_sides:
        fed.plan.ghost_view(rr, side).fill_(1e30)
    errs = [fed.step_e2e() for _ in range(4)]
    torch.cuda.synchronize()
    fed.check()
    assert max(errs) < 1e-4 and errs[0] == pytest.approx(errs[-1], rel=0.1, abs=1e-9)  # atomics: order varies
