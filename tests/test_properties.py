"""Property tests (hypothesis): decomposition arithmetic, slice rule, closed forms, stencil
exactness on polynomials -- and, on a GPU, the kernels against the CPU oracle over shapes that
are not multiples of any tile or vector width."""
import math

import pytest
import torch
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

from gpu_mpi_tests_b200.parallel import decomp
from gpu_mpi_tests_b200.parallel.array import normalize_slice
from gpu_mpi_tests_b200.parallel.runtime import RankDeviceError, map_rank_to_device
from gpu_mpi_tests_b200.utils import oracle

CPU = settings(max_examples=60, deadline=None)


@CPU
@given(world=st.integers(1, 64), n_local=st.integers(1, 1 << 20))
def test_slabs_tile_the_axis(world, n_local):
    slabs = [decomp.weak_slab(n_local, world, r) for r in range(world)]
    assert sum(s.n_local for s in slabs) == n_local * world
    assert [s.start for s in slabs] == [r * n_local for r in range(world)]
    assert slabs[0].lo == -1 and slabs[-1].hi == -1
    for a, b in zip(slabs, slabs[1:]):
        assert a.hi == b.rank and b.lo == a.rank


@CPU
@given(world=st.integers(2, 64), n_global=st.integers(1, 1 << 24))
def test_indivisible_domain_is_an_error(world, n_global):
    if n_global % world:
        with pytest.raises(decomp.DecompositionError, match="must be divisor of domain size"):
            decomp.Slab1D(n_global, world, 0)
    else:
        assert decomp.Slab1D(n_global, world, world - 1).n_local * world == n_global


@CPU
@given(world=st.integers(1, 256))
def test_grid_neighbours_are_mutual(world):
    px, py = decomp.best_grid(world)
    assert px * py == world and px >= py
    for r in range(world):
        g = decomp.Grid2D(px, py, r)
        for side in range(4):
            nb = g.neighbour(side)
            if nb >= 0:
                assert decomp.Grid2D(px, py, nb).neighbour(side ^ 1) == r
            assert bool(g.physical_mask() >> side & 1) == (nb < 0)


@CPU
@given(n_ranks=st.integers(1, 96), n_dev=st.integers(1, 16))
def test_rank_device_rule(n_ranks, n_dev):
    """mpi_daxpy.cc:41-55: block mapping; more ranks than devices only in whole multiples."""
    if n_ranks > n_dev and n_ranks % n_dev:
        with pytest.raises(RankDeviceError, match="not a multiple of number of GPUs"):
            map_rank_to_device(n_ranks, 0, n_dev)
        return
    devs = [map_rank_to_device(n_ranks, r, n_dev)[0] for r in range(n_ranks)]
    assert devs == sorted(devs) and all(0 <= d < n_dev for d in devs)
    per = max(1, n_ranks // n_dev)
    assert all(devs.count(d) == per for d in set(devs))


@CPU
@given(extent=st.integers(1, 100), start=st.integers(-100, 100), end=st.integers(-100, 100))
def test_slice_rule_matches_python_slicing(extent, start, end):
    """mpi_stencil2d_sycl_oo.cc:174-181: negatives count from the end, end == 0 means 'to the end'."""
    data = list(range(extent))
    try:
        s, e = normalize_slice(start, end, extent)
    except IndexError:
        return
    py_end = None if (end == 0 and (start > 0 or start < 0)) else end
    if -extent <= start < extent:
        assert data[s:e] == data[start:py_end]
    assert 0 <= s < e <= extent


@CPU
@given(rank=st.integers(0, 40), n=st.integers(1, 5000))
def test_gather_slot_closed_form(rank, n):
    assert float(oracle.gather_slot_values(rank, n).sum()) == oracle.gather_slot_sum(rank, n)


@CPU
@given(n=st.integers(8, 400), x0=st.floats(-4, 4), dx=st.floats(1e-3, 0.5),
       c=st.lists(st.floats(-2, 2), min_size=5, max_size=5))
def test_stencil_is_exact_for_quartics(n, x0, dx, c):
    """The 5-point 4th-order first derivative differentiates polynomials up to degree 4 exactly."""
    x = x0 + dx * torch.arange(-2, n + 2, dtype=torch.float64)
    y = sum(ck * x ** k for k, ck in enumerate(c))
    dy = sum(k * ck * x[2:-2] ** (k - 1) for k, ck in enumerate(c) if k)
    got = oracle.stencil1d_5(y, 1.0 / dx)
    scale = 1.0 + float(y.abs().max()) / dx
    assert float((got - dy).abs().max()) <= 1e-11 * scale


from gpu_mpi_tests_b200.ops.local import VARIANTS

VARIANT_NAMES = sorted(VARIANTS)
GPU = settings(max_examples=25, deadline=None, suppress_health_check=[HealthCheck.too_slow])


def _close(a, b, rtol=1e-11):
    return bool(((a - b).abs() <= rtol * (1.0 + b.abs())).all())


@pytest.mark.gpu
@GPU
@given(n=st.integers(1, 70000), a=st.floats(-3, 3), off=st.integers(0, 3))
def test_gpu_daxpy_sum_any_length_and_alignment(n, a, off):
    from gpu_mpi_tests_b200.ops import local
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n + off, dtype=torch.float64, generator=g).to(dev)[off:]
    y = torch.randn(n + off, dtype=torch.float64, generator=g).to(dev)[off:]
    want = a * x.cpu() + y.cpu()
    s = local.daxpy_sum(a, x, y)
    assert _close(y.cpu(), want)
    assert float(s) == pytest.approx(float(want.sum()), rel=1e-10, abs=1e-9 * math.sqrt(n))


@pytest.mark.gpu
@GPU
@given(n=st.integers(1, 50000), off=st.integers(0, 3), variant=st.sampled_from(VARIANT_NAMES))
def test_gpu_stencil1d_any_length(n, off, variant):
    from gpu_mpi_tests_b200.ops import local
    dev = torch.device("cuda", 0)
    y = torch.randn(n + 4 + off, dtype=torch.float64, generator=torch.Generator().manual_seed(n)).to(dev)[off:]
    got = local.stencil1d(y, 3.5, variant=variant)
    assert _close(got.cpu(), oracle.stencil1d_5(y.cpu(), 3.5))


@pytest.mark.gpu
@GPU
@given(nx=st.integers(1, 300), ny=st.integers(1, 300), dim=st.sampled_from([0, 1]), pad=st.integers(0, 5),
       variant=st.sampled_from(VARIANT_NAMES))
def test_gpu_stencil2d_any_shape(nx, ny, dim, pad, variant):
    from gpu_mpi_tests_b200.ops import local
    dev = torch.device("cuda", 0)
    gx, gy = (2, 0) if dim == 0 else (0, 2)
    g = torch.Generator().manual_seed(nx * 1000 + ny)
    full = torch.randn(ny + 2 * gy, nx + 2 * gx + pad, dtype=torch.float64, generator=g).to(dev)
    z = full[:, :nx + 2 * gx]                      # leading dimension larger than the row length
    got = local.stencil2d(z, dim, 0.75, variant=variant)
    assert _close(got.cpu(), oracle.stencil2d_1d_5(z.cpu(), dim, 0.75))


@pytest.mark.gpu
@GPU
@given(nx=st.integers(1, 200), ny=st.integers(1, 200), axis=st.sampled_from([0, 1]))
def test_gpu_axis_sum_and_diff_norm_any_shape(nx, ny, axis):
    from gpu_mpi_tests_b200.ops import local
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(nx * 977 + ny)
    a = torch.randn(ny, nx, dtype=torch.float64, generator=g).to(dev)
    b = torch.randn(ny, nx, dtype=torch.float64, generator=g).to(dev)
    got = local.axis_sum(a, axis)
    want = a.cpu().sum(dim=1 if axis == 0 else 0)   # torch [j, i]: axis 0 sums over i
    assert _close(got.cpu(), want, rtol=1e-10)
    assert local.diff_norm(a, b) == pytest.approx(float(((a - b) ** 2).sum().sqrt()), rel=1e-10)
