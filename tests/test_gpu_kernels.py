"""Single-GPU numerics: every sm_100a kernel against a plain PyTorch fp64 reference of the
same op, including shapes that do not divide the tile sizes."""
import pytest
import torch

from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.utils import oracle

pytestmark = pytest.mark.gpu
F64 = torch.float64
TOL = 1e-12


def rel(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-300))


SIZES_1D = [1, 2, 3, 4, 7, 1000, 1024, 4096, 4096 * 3 + 8, (1 << 20) + 6, 5 * (1 << 20)]


@pytest.mark.parametrize("n", SIZES_1D)
def test_daxpy(dev, n):
    g = torch.Generator(device="cuda").manual_seed(n)
    x = torch.rand(n, dtype=F64, device=dev, generator=g)
    y = torch.rand(n, dtype=F64, device=dev, generator=g)
    ref = 2.0 * x + y
    got = local.daxpy(2.0, x, y.clone())
    assert rel(got, ref) < TOL


@pytest.mark.parametrize("n", SIZES_1D)
def test_daxpy_sum_and_reductions(dev, n):
    g = torch.Generator(device="cuda").manual_seed(n + 1)
    x = torch.rand(n, dtype=F64, device=dev, generator=g)
    y = torch.rand(n, dtype=F64, device=dev, generator=g)
    ref = 2.0 * x + y
    y2 = y.clone()
    for _ in range(2):  # scratch must be reusable call after call
        y2.copy_(y)
        s = local.daxpy_sum(2.0, x, y2)
        assert rel(y2, ref) < TOL
        assert abs(s.item() - ref.sum().item()) <= 1e-12 * abs(ref.sum().item()) + 1e-300
    assert abs(local.total(x).item() - x.sum().item()) <= 1e-12 * x.sum().item()
    r = ((x - y) ** 2).sum().item()
    assert abs(local.sumsq_diff(x, y).item() - r) <= 1e-12 * r + 1e-300


def test_daxpy_unaligned_views(dev):
    x = torch.rand(10001, dtype=F64, device=dev)
    y = torch.rand(10001, dtype=F64, device=dev)
    xs, ys = x[1:], y[1:]  # 8-byte aligned only -> scalar path
    ref = 3.0 * xs + ys
    s = local.daxpy_sum(3.0, xs, ys)
    assert rel(ys, ref) < TOL and abs(s.item() - ref.sum().item()) < 1e-9


def test_reference_daxpy_values(dev):
    """daxpy.cu:55-87: x=i+1, y=-i-1, a=2 -> y=i+1, SUM = 524800."""
    x, y = oracle.daxpy_init_small(1024, dev)
    s = local.daxpy_sum(2.0, x, y)
    assert torch.equal(y, torch.arange(1, 1025, dtype=F64, device=dev))
    assert s.item() == 524800.0


@pytest.mark.parametrize("variant", ["generic", "vector", "tma"])
@pytest.mark.parametrize("n", [4, 6, 1000, 4096, 4098, 3 * 4096 + 10, (1 << 20) + 2])
def test_stencil1d(dev, variant, n):
    y = torch.rand(n + 4, dtype=F64, device=dev)
    ref = oracle.stencil1d_5(y, 3.5)
    got = local.stencil1d(y, 3.5, variant=variant)
    assert rel(got, ref) < TOL


@pytest.mark.parametrize("variant", ["generic", "vector", "tma"])
@pytest.mark.parametrize("shape", [(8, 8), (1024, 64), (2048, 130), (12, 7), (4096, 257), (1028, 33), (10, 5)])
def test_stencil2d(dev, variant, shape):
    nx, ny = shape
    z0 = torch.rand(ny, nx + 4, dtype=F64, device=dev)
    assert rel(local.stencil2d(z0, 0, 2.5, variant=variant), oracle.stencil2d_1d_5(z0, 0, 2.5)) < TOL
    z1 = torch.rand(ny + 4, nx, dtype=F64, device=dev)
    assert rel(local.stencil2d(z1, 1, 2.5, variant=variant), oracle.stencil2d_1d_5(z1, 1, 2.5)) < TOL
    zz = torch.rand(ny + 4, nx + 4, dtype=F64, device=dev)
    assert rel(local.stencil2d_xy(zz, 2.5, 1.5), oracle.stencil2d_xy(zz, 2.5, 1.5)) < TOL


@pytest.mark.parametrize("shape", [(1024, 64), (2048, 130), (4096, 257), (1028, 33), (8, 8), (12, 7), (5120, 16)])
def test_stencil2d_dim1_tma_cluster_variant(dev, shape):
    """2-D tensor-map TMA loads + 2-CTA clusters with DSMEM halo sharing vs the fp64 reference."""
    nx, ny = shape
    z1 = torch.rand(ny + 4, nx, dtype=F64, device=dev)
    ref = oracle.stencil2d_1d_5(z1, 1, 2.5)
    out = torch.full((ny, nx), -7.0, dtype=F64, device=dev)
    local.stencil2d(z1, 1, 2.5, out=out, variant="tma")
    assert rel(out, ref) < TOL
    # twice in a row on the same stream (barrier phases, stage reuse)
    local.stencil2d(z1, 1, 2.5, out=out, variant="tma")
    assert rel(out, ref) < TOL


def test_stencil2d_padded_leading_dimension(dev):
    nx, ny = 64, 20
    big = torch.rand(ny, nx + 4 + 12, dtype=F64, device=dev)
    z = big[:, :nx + 4]  # ld = nx + 16
    outbig = torch.zeros(ny, nx + 8, dtype=F64, device=dev)
    out = outbig[:, :nx]
    local.stencil2d(z, 0, 1.0, out=out)
    assert rel(out, oracle.stencil2d_1d_5(z, 0, 1.0)) < TOL
    assert float(outbig[:, nx:].abs().max()) == 0.0


@pytest.mark.parametrize("shape", [(2, 1000), (1000, 2), (17, 13), (2, 524288 // 64)])
def test_copy2d_pack_unpack(dev, shape):
    rows, cols = shape
    src_full = torch.rand(cols + 3, rows + 6, dtype=F64, device=dev)
    src = src_full[1:1 + cols, 2:2 + rows]
    dst = torch.zeros(cols, rows, dtype=F64, device=dev)
    local.copy2d(src, dst)
    assert torch.equal(dst, src)
    back = torch.zeros_like(src_full)
    local.copy2d(dst, back[1:1 + cols, 2:2 + rows])
    assert torch.equal(back[1:1 + cols, 2:2 + rows], src)
    assert float(back.sum() - src.sum()) == pytest.approx(0.0, abs=1e-9)


@pytest.mark.parametrize("shape", [(1024, 96), (100, 37), (33, 1000)])
def test_axis_sum(dev, shape):
    nx, ny = shape
    z = torch.rand(ny, nx, dtype=F64, device=dev)
    a0 = local.axis_sum(z, 0)
    a1 = local.axis_sum(z, 1)
    assert rel(a0, z.sum(dim=1)) < TOL and rel(a1, z.sum(dim=0)) < TOL
    part = local.axis_sum(z, 1, n_out=min(nx, 16))
    assert rel(part, z.sum(dim=0)[:min(nx, 16)]) < TOL


def test_fill_iota_sumsq2d(dev):
    x = torch.empty(100003, dtype=F64, device=dev)
    local.fill(x, 3.25)
    assert float(x.min()) == 3.25 and float(x.max()) == 3.25
    local.iota(x, 1.0, 1.0, 1.0 / x.numel())
    assert rel(x, (torch.arange(x.numel(), dtype=F64, device=dev) + 1) / x.numel()) < TOL
    a = torch.rand(50, 70, dtype=F64, device=dev)
    b = torch.rand(50, 70, dtype=F64, device=dev)
    got = local.sumsq_diff(a[:, :64], b[:, :64]).item()
    assert got == pytest.approx(((a[:, :64] - b[:, :64]) ** 2).sum().item(), rel=1e-12)


def test_device_side_init_matches_oracle(dev):
    n, world = 4096, 4
    for rank in range(world):
        y_ref, act_ref, scale = oracle.init_stencil1d(n, rank, world)
        y = torch.empty(n + 4, dtype=F64, device=dev)
        act = torch.empty(n, dtype=F64, device=dev)
        dx = 8.0 / (n * world)
        local.init_stencil1d(y, act, n, rank * n * dx, dx, rank == 0, rank == world - 1)
        assert rel(y.cpu(), y_ref) < 1e-14 and rel(act.cpu(), act_ref) < 1e-14
    for dim in (0, 1):
        for rank in range(3):
            z_ref, act_ref, scale = oracle.init_stencil2d(64, 40, dim, rank, 3)
            nx, ny = (64, 40) if dim == 0 else (40, 64)
            gx, gy = (2, 0) if dim == 0 else (0, 2)
            delta = 8.0 / (64 * 3)
            z = torch.empty(ny + 2 * gy, nx + 2 * gx, dtype=F64, device=dev)
            act = torch.empty(ny, nx, dtype=F64, device=dev)
            x0, y0 = (rank * 64 * delta, 0.0) if dim == 0 else (0.0, rank * 64 * delta)
            lo, hi = (1, 2) if dim == 0 else (4, 8)
            mask = (lo if rank == 0 else 0) | (hi if rank == 2 else 0)
            local.init_stencil2d(z, act, nx, ny, gx, gy, x0, y0, delta, delta, 1 if dim == 0 else 2, 1.0, 1.0, mask, 0.0)
            assert rel(z.cpu(), z_ref) < 1e-14 and rel(act.cpu(), act_ref) < 1e-14


def test_errors_are_loud(native):
    from gpu_mpi_tests_b200.ops._lib import GmtError
    with pytest.raises(GmtError):
        native.gmt_stencil1d_f64(16, 0, 0, 1.0, 1, 0)
    with pytest.raises(GmtError):
        native.gmt_axis_sum_f64(4, 4, 2, 8, 0, 4, 8, 0)
