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


@pytest.mark.parametrize("world,n", [(1, 1000), (2, 128), (4, 4096), (8, 1 << 16), (3, 1001), (8, 7)])
def test_allgather_inplace_and_gather(dev, world, n):
    pc = PeerCollectives(VirtualSpace(world, dev), buf_doubles=world * n, spin_limit=1 << 24)
    for it in range(3):
        slices = [torch.rand(n, dtype=F64, device=dev) + r + 10 * it for r in range(world)]
        for r in range(world):
            b = pc.buffer(r)
            b.fill_(-1.0)
            b[r * n:(r + 1) * n] = slices[r]
        pc.allgather_inplace(n)
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


def test_allgather_from_separate_source(dev):
    world, n = 4, 2048
    pc = PeerCollectives(VirtualSpace(world, dev), buf_doubles=world * n, spin_limit=1 << 24)
    src = {r: torch.rand(n, dtype=F64, device=dev) for r in range(world)}
    pc.allgather(n, src)
    torch.cuda.synchronize()
    ref = torch.cat([src[r] for r in range(world)])
    for r in range(world):
        assert torch.equal(pc.buffer(r), ref)


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


@pytest.mark.parametrize("world,n", [(1, 1024), (2, 1 << 20), (8, (1 << 18) + 6), (4, 1000)])
def test_daxpy_allreduce(dev, world, n):
    pc = PeerCollectives(VirtualSpace(world, dev), spin_limit=1 << 24)
    xs, ys, outs = {}, {}, {}
    for r in range(world):
        xs[r], ys[r] = oracle.daxpy_init_scaled(n, dev)
        outs[r] = torch.zeros(1, dtype=F64, device=dev)
    pc.daxpy_allreduce(2.0, xs, ys, outs)
    torch.cuda.synchronize()
    expect = oracle.daxpy_scaled_expected_allsum(n, world)  # ALLSUM of mpi_daxpy_nvtx.cc:310
    for r in range(world):
        assert outs[r].item() == pytest.approx(expect, rel=1e-12)
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
