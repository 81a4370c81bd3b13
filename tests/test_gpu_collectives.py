"""In-kernel collectives, all ranks emulated on one GPU in one cooperative launch
(VirtualSpace): allgather (in place and from a separate source), gather to root, all-reduce
(one-shot and two-shot), fused daxpy + global sum, fused axis-sum + all-reduce, barrier --
against plain PyTorch fp64 references, repeated so that epochs and staging parity cycle."""
import pytest
import torch

from gpu_mpi_tests_b200.parallel.collectives import PeerCollectives
from gpu_mpi_tests_b200.parallel.symm import VirtualSpace
from gpu_mpi_tests_b200.utils import oracle

pytestmark = pytest.mark.gpu
F64 = torch.float64


@pytest.mark.parametrize("mode", ["ll", "p2p"])
@pytest.mark.parametrize("world,n", [(1, 1000), (2, 128), (4, 4096), (8, 1 << 16), (3, 1001), (8, 7), (8, 8192)])
def test_allgather_inplace_and_gather(dev, world, n, mode):
    pc = PeerCollectives(VirtualSpace(world, dev), buf_doubles=world * n, spin_limit=1 << 24)
    if mode == "ll" and n > pc.ll_gather_max:
        with pytest.raises(ValueError):
            pc.allgather_inplace(n, mode="ll")
        mode = None   # the default picks the flag-based kernel for slices beyond 64 KiB
    for it in range(3):
        slices = [torch.rand(n, dtype=F64, device=dev) + r + 10 * it for r in range(world)]
        for r in range(world):
            b = pc.buffer(r)
            b.fill_(-1.0)
            b[r * n:(r + 1) * n] = slices[r]
        pc.allgather_inplace(n, mode=mode)
        torch.cuda.synchronize()
        ref = torch.cat(slices)
        for r in range(world):
            assert torch.equal(pc.buffer(r), ref), f"iter {it} rank {r}"
    # gather to root 1 (or 0): only the root's buffer is complete, nobody else's is touched
    root = min(1, world - 1)
    slices = [torch.rand(n, dtype=F64, device=dev) for r in range(world)]
    for r in range(world):
        b = pc.buffer(r)
        b.fill_(-1.0)
        b[r * n:(r + 1) * n] = slices[r]
    pc.gather_inplace(n, root)
    torch.cuda.synchronize()
    assert torch.equal(pc.buffer(root), torch.cat(slices))
    for r in range(world):
        if r != root:
            others = torch.cat([pc.buffer(r)[:r * n], pc.buffer(r)[(r + 1) * n:]])
            assert others.numel() == 0 or float(others.max()) == -1.0
    pc.check_status()
    assert pc.epochs() == [4] * world


@pytest.mark.parametrize("mode", ["ll", "p2p"])
def test_allgather_from_separate_source(dev, mode):
    world, n = 4, 2048
    pc = PeerCollectives(VirtualSpace(world, dev), buf_doubles=world * n, spin_limit=1 << 24)
    for _ in range(3):
        src = {r: torch.rand(n, dtype=F64, device=dev) for r in range(world)}
        pc.allgather(n, src, mode=mode)
        torch.cuda.synchronize()
        ref = torch.cat([src[r] for r in range(world)])
        for r in range(world):
            assert torch.equal(pc.buffer(r), ref)
    pc.check_status()


def test_too_small_staging_area_is_reported_not_overrun(dev):
    """The raw entry points cannot see stage_cap from the host: a call that needs more LL cells than
    the descriptor provides must write nothing out of bounds and say so in the status word."""
    from gpu_mpi_tests_b200.ops._lib import lib
    from gpu_mpi_tests_b200.parallel.collectives import CollectiveTimeout
    world, n = 2, 4096
    pc = PeerCollectives(VirtualSpace(world, dev), buf_doubles=world * n, stage_cap=2048, spin_limit=1 << 20)
    for r in range(world):
        pc.buffer(r).fill_(-1.0)
    lib.gmt_allgather_f64(pc._ranks_dev.data_ptr(), world, n, 0, 2, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    with pytest.raises(CollectiveTimeout, match="stage_cap"):
        pc.check_status()
    assert all(float(pc.buffer(r).max()) == -1.0 for r in range(world))
    # the all-reduce entry points say so as well instead of overrunning the staging area
    for algo in (0, 1, 3):
        pc2 = PeerCollectives(VirtualSpace(world, dev), stage_cap=2048, spin_limit=1 << 20)
        vec = {r: torch.ones(4096, dtype=F64, device=dev) for r in range(world)}
        pc2.bind({r: (vec[r], vec[r], None) for r in range(world)})
        lib.gmt_allreduce_f64(pc2._ranks_dev.data_ptr(), world, 4096, algo, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        with pytest.raises(CollectiveTimeout, match="stage_cap"):
            pc2.check_status()
        assert all(float(vec[r].min()) == float(vec[r].max()) == 1.0 for r in range(world))
        assert pc2.epochs() == [1] * world          # the ranks stay in step


@pytest.mark.parametrize("algo", ["ll", "one_shot", "two_shot"])
@pytest.mark.parametrize("world,n", [(2, 1024), (8, 1024), (4, 1), (3, 4097), (8, 1 << 17)])
def test_allreduce(dev, algo, world, n):
    pc = PeerCollectives(VirtualSpace(world, dev), stage_cap=max(n, 8), spin_limit=1 << 24)
    for it in range(4):  # parity of the staging area alternates
        vec = {r: torch.rand(n, dtype=F64, device=dev) * (r + 1) for r in range(world)}
        ref = torch.zeros(n, dtype=F64, device=dev)
        for r in range(world):
            ref += vec[r]
        pc.allreduce(vec, algo=algo)  # in place
        torch.cuda.synchronize()
        for r in range(world):
            assert float((vec[r] - ref).abs().max()) <= 1e-13 * float(ref.abs().max())
            assert torch.equal(vec[r], vec[0])  # rank-ordered adds: identical bits on every rank
    pc.check_status()


@pytest.mark.parametrize("misalign", [0, 1])
@pytest.mark.parametrize("world,n", [(1, 1024), (2, 1 << 20), (8, (1 << 18) + 6), (4, 1000), (1, 1 << 24), (3, 3), (8, 5 * 4096 * 37 + 13)])
def test_daxpy_allreduce(dev, world, n, misalign):
    """One partial per CTA travels as an LL cell to every rank; the cases cover one chunk per CTA,
    several chunks per CTA (a cooperative launch of 8 ranks leaves 37 CTAs each), the scalar tail,
    n smaller than a vector, and operands that are not 32-byte aligned (scalar path)."""
    if misalign and n > (1 << 22):
        pytest.skip("the scalar path is covered by the smaller cases")
    pc = PeerCollectives(VirtualSpace(world, dev), spin_limit=1 << 24)
    xs, ys, outs = {}, {}, {}
    for r in range(world):
        x, y = oracle.daxpy_init_scaled(n, dev)
        if misalign:   # 8-byte offset views: not 32-byte aligned
            xb, yb = torch.empty(n + 1, dtype=F64, device=dev), torch.empty(n + 1, dtype=F64, device=dev)
            xb[1:].copy_(x)
            yb[1:].copy_(y)
            x, y = xb[1:], yb[1:]
        xs[r], ys[r] = x, y
        outs[r] = torch.zeros(1, dtype=F64, device=dev)
    pc.daxpy_allreduce(2.0, xs, ys, outs)
    torch.cuda.synchronize()
    expect = oracle.daxpy_scaled_expected_allsum(n, world)  # ALLSUM of mpi_daxpy_nvtx.cc:310
    for r in range(world):
        assert outs[r].item() == pytest.approx(expect, rel=1e-12)
        assert outs[r].item() == pytest.approx(world * float(torch.sum(xs[r])), rel=1e-12)   # torch fp64 reference
        assert torch.equal(outs[r], outs[0])                 # fixed summation order: same bits on every rank
        assert float((ys[r] - xs[r]).abs().max()) < 1e-15  # y = 2x - x = x
    # second call on the updated y: y = 2x + x = 3x
    pc.daxpy_allreduce(2.0, xs, ys, outs)
    torch.cuda.synchronize()
    for r in range(world):
        assert outs[r].item() == pytest.approx(3 * expect, rel=1e-12)
    pc.check_status()


@pytest.mark.parametrize("axis", [0, 1])
def test_axis_sum_allreduce(dev, axis):
    world, nx, ny = 4, 96, 80
    pc = PeerCollectives(VirtualSpace(world, dev), stage_cap=256, spin_limit=1 << 24)
    z = {r: torch.rand(ny, nx, dtype=F64, device=dev) for r in range(world)}
    n_out = 64  # the reference all-reduces n_local_deriv entries, not the full surviving axis
    out = {r: torch.zeros(n_out, dtype=F64, device=dev) for r in range(world)}
    for _ in range(3):
        pc.axis_sum_allreduce(z, axis, out, n_out)
    torch.cuda.synchronize()
    ref = sum((z[r].sum(dim=1) if axis == 0 else z[r].sum(dim=0))[:n_out] for r in range(world))
    for r in range(world):
        assert float((out[r] - ref).abs().max()) <= 1e-12 * float(ref.abs().max())
    pc.check_status()


@pytest.mark.parametrize("axis", [0, 1])
def test_axis_sum_allreduce_large_takes_the_one_shot_path(dev, axis):
    """n_out x world beyond 2^18 cells: the kernel itself switches from LL cells to the staged one-shot
    (the C entry point used to keep LL up to 32 Ki doubles whatever the world size)."""
    world = 8
    n_out = (1 << 15) + 8            # x 8 ranks > 2^18
    nx, ny = (4, n_out) if axis == 0 else (n_out, 3)
    pc = PeerCollectives(VirtualSpace(world, dev), stage_cap=n_out, spin_limit=1 << 24)
    z = {r: torch.rand(ny, nx, dtype=F64, device=dev) for r in range(world)}
    out = {r: torch.zeros(n_out, dtype=F64, device=dev) for r in range(world)}
    for _ in range(2):
        pc.axis_sum_allreduce(z, axis, out, n_out)
    torch.cuda.synchronize()
    ref = sum((z[r].sum(dim=1) if axis == 0 else z[r].sum(dim=0))[:n_out] for r in range(world))
    for r in range(world):
        assert float((out[r] - ref).abs().max()) <= 1e-12 * float(ref.abs().max())
        assert torch.equal(out[r], out[0])
    pc.check_status()


@pytest.mark.parametrize("axis", [0, 1])
def test_axis_sum_allreduce_does_not_need_its_whole_grid_resident(dev, axis):
    """The LL flavour must not depend on co-residency: whoever waits for a cell waits for another
    RANK's push only, never for a sibling CTA (ranks sharing a GPU across processes, or another stream
    holding SMs, leave only part of a non-cooperative grid resident).  The launch is given 16 times
    more CTAs than fit on the GPU at once; a CTA that waited for a cell a not-yet-resident CTA has to
    push would expire its (short) bounded spin and report it in the status word."""
    from gpu_mpi_tests_b200.ops._lib import lib
    n_out = 1 << 17        # 8192 CTAs' worth of columns (axis 0) / 4096 (axis 1); x 1 rank <= 2^18: LL path
    nx, ny = (8, n_out) if axis == 0 else (n_out, 4)
    pc = PeerCollectives(VirtualSpace(1, dev), stage_cap=n_out, spin_limit=200_000)
    z = {0: torch.rand(ny, nx, dtype=F64, device=dev)}
    out = {0: torch.zeros(n_out, dtype=F64, device=dev)}
    lib.gmt_debug_set_grid_factor(16)
    try:
        for _ in range(3):
            pc.axis_sum_allreduce(z, axis, out, n_out)
        torch.cuda.synchronize()
    finally:
        lib.gmt_debug_set_grid_factor(1)
    pc.check_status()
    ref = (z[0].sum(dim=1) if axis == 0 else z[0].sum(dim=0))[:n_out]
    assert float((out[0] - ref).abs().max()) <= 1e-12 * float(ref.abs().max())


def test_barrier_and_timeout(dev):
    pc = PeerCollectives(VirtualSpace(4, dev), spin_limit=1 << 24)
    for _ in range(5):
        pc.barrier()
    torch.cuda.synchronize()
    pc.check_status()
    assert pc.epochs() == [5] * 4
    # a rank whose peers never show up gives up instead of hanging
    from gpu_mpi_tests_b200.ops._lib import lib
    from gpu_mpi_tests_b200.parallel.collectives import CollectiveTimeout
    lone = PeerCollectives(VirtualSpace(2, dev), spin_limit=2000)
    lib.gmt_barrier(lone._ranks_dev.data_ptr(), 1, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    with pytest.raises(CollectiveTimeout):
        lone.check_status()
