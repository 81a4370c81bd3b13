"""CPU unit tests (no GPU, no process group): decomposition arithmetic, rank->device mapping
incl. the error message, slice normalisation, stencil math against closed forms, closed-form
sums, golden output strings."""
import math

import pytest
import torch

from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.parallel import array as ga
from gpu_mpi_tests_b200.parallel import decomp, runtime
from gpu_mpi_tests_b200.utils import oracle, report

F64 = torch.float64


# ------------------------------------------------------------------ rank -> device
def test_rank_device_mapping_identity_and_oversubscription():
    assert runtime.map_rank_to_device(4, 3, 8) == (3, 1)
    # 42 procs on 6 GPUs (daxpy.cu:6-7): 7 consecutive ranks share a device
    assert [runtime.map_rank_to_device(42, r, 6)[0] for r in (0, 6, 7, 41)] == [0, 0, 1, 5]
    assert runtime.map_rank_to_device(42, 0, 6)[1] == 7


def test_rank_device_mapping_error_message():
    with pytest.raises(runtime.RankDeviceError) as e:
        runtime.map_rank_to_device(7, 0, 6)
    assert str(e.value) == "ERROR: Number of ranks (7) not a multiple of number of GPUs (6)"


def test_vendor_id_packing():
    assert runtime.pack_vendor_id(0, 0x1B, 0) == 0x00001B00
    assert report.err_norm_line(1, 4, 1, runtime.pack_vendor_id(1, 0xAB, 3), 0.5) == \
        "1/4 [1:0x0001ab03] err_norm = 0.50000000"


def test_world_of_one_runtime(monkeypatch):
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    rt = runtime.init(use_cuda=False)
    assert (rt.world_size, rt.rank, rt.node_count, rt.distributed) == (1, 0, 1, False)
    assert rt.reduce_scalar(3.5) == 3.5
    assert rt.rank_device_line().startswith("RANK[1/1] => DEVICE[1/1] mem=")
    rt.barrier()
    runtime.finalize()


# ------------------------------------------------------------------ decomposition
def test_slab_decomposition():
    s = decomp.Slab1D(32 * 1024 * 1024, 8, 3)
    assert (s.n_local, s.n_local_with_ghost, s.start, s.lo, s.hi) == (4 * 1024 * 1024, 4 * 1024 * 1024 + 4,
                                                                      12 * 1024 * 1024, 2, 4)
    assert decomp.Slab1D(64, 4, 0).lo == -1 and decomp.Slab1D(64, 4, 3).hi == -1
    with pytest.raises(decomp.DecompositionError) as e:
        decomp.Slab1D(33554432, 3, 2)
    assert str(e.value) == "2 nmpi (3) must be divisor of domain size (33554432), exiting"
    assert decomp.weak_slab(1024, 4, 1).n_global == 4096


def test_grid_decomposition():
    assert decomp.best_grid(8) == (4, 2) and decomp.best_grid(4) == (2, 2) and decomp.best_grid(2) == (2, 1)
    assert decomp.best_grid(1) == (1, 1) and decomp.best_grid(6) == (3, 2)
    g = decomp.Grid2D(4, 2, 5)
    assert (g.cx, g.cy) == (1, 1)
    assert g.neighbours() == [4, 6, 1, -1]
    assert g.physical_mask() == 8
    assert decomp.Grid2D(1, 1, 0).physical_mask() == 15
    assert decomp.slab_as_grid(0, 4, 2).neighbours() == [1, 3, -1, -1]
    assert decomp.slab_as_grid(1, 4, 0).neighbours() == [-1, -1, -1, 1]
    # every face is shared symmetrically
    for r in range(8):
        g = decomp.Grid2D(4, 2, r)
        for s in range(4):
            nb = g.neighbour(s)
            if nb >= 0:
                assert decomp.Grid2D(4, 2, nb).neighbour(s ^ 1) == r


# ------------------------------------------------------------------ slices / arrays
def test_slice_normalisation_rules():
    """mpi_stencil2d_sycl_oo.cc:174-181: negatives count from the end, end == 0 means to-end."""
    assert ga.normalize_slice(2, 4, 12) == (2, 4)
    assert ga.normalize_slice(-4, -2, 12) == (8, 10)
    assert ga.normalize_slice(-2, 0, 12) == (10, 12)
    assert ga.normalize_slice(0, 2, 12) == (0, 2)
    with pytest.raises(IndexError):
        ga.normalize_slice(5, 3, 12)
    with pytest.raises(IndexError):
        ga.normalize_slice(0, 13, 12)


def test_array2d_is_column_major_and_slice_copies():
    """The 6x6(+ghost) pattern of the reference's disabled test_buf_view
    (mpi_stencil2d_sycl.cc:118-159): recognisable values through pack and unpack."""
    n, nb = 6, 2
    a = ga.Array2D(n + 2 * nb, n, device=torch.device("cpu"))
    for col in range(n):
        for row in range(n + 2 * nb):
            a.set(row, col, 100.0 * row + col)
    assert a(3, 4) == 304.0
    assert a.t[4, 3] == 304.0 and a.t.stride(1) == 1  # first index fastest
    with pytest.raises(IndexError):
        a(n + 2 * nb, 0)
    sbuf = ga.Array2D(nb, n, device=torch.device("cpu"))
    ga.copy_src_slice(a, sbuf, 0, nb, 2 * nb)            # rows [2,4): left interior boundary
    assert sbuf(0, 5) == 205.0 and sbuf(1, 0) == 300.0
    ga.copy_src_slice(a, sbuf, 0, -2 * nb, -nb)          # rows [6,8): right interior boundary
    assert sbuf(0, 1) == 601.0
    ga.copy_dest_slice(sbuf, a, 0, -nb, 0)               # into the right ghost rows [8,10)
    assert a(8, 1) == 601.0 and a(9, 2) == 702.0
    cbuf = ga.Array2D(n + 2 * nb, 2, device=torch.device("cpu"))
    ga.copy_src_slice(a, cbuf, 1, -2, 0)                 # last two columns
    assert cbuf(3, 1) == 305.0
    with pytest.raises(ValueError):
        ga.copy_src_slice(a, cbuf, 0, 0, 2)


def test_memory_spaces_on_cpu():
    for space in ga.SPACES:
        t = ga.zeros((3, 4), space, torch.device("cpu"))
        assert t.shape == (3, 4) and t.dtype == F64 and float(t.sum()) == 0.0
    with pytest.raises(ValueError):
        ga.empty(4, "texture")
    assert float(ga.full(5, math.pi, device=torch.device("cpu")).sum()) == pytest.approx(5 * math.pi)


# ------------------------------------------------------------------ math
def test_daxpy_closed_forms():
    x, y = oracle.daxpy_init_small(1024)
    local.daxpy(2.0, x, y)
    assert float(y.sum()) == oracle.daxpy_small_expected_sum(1024) == 524800.0
    assert report.daxpy_sum(float(y.sum())) == "SUM = 524800.000000"
    n = 1 << 20
    x, y = oracle.daxpy_init_scaled(n)
    s = local.daxpy_sum(2.0, x, y)
    assert float(s) == pytest.approx(oracle.daxpy_scaled_expected_sum(n), rel=1e-12)
    assert report.rank_sum(0, 2, 524288.5) == "0/2 SUM = 524288.500000"
    assert report.rank_allsum(1, 2, oracle.daxpy_scaled_expected_allsum(n, 2)) == "1/2 ALLSUM = 1048577.000000"


def test_stencil_is_exact_for_cubics_and_err_norm_is_roundoff():
    n, world = 1 << 16, 4
    total = 0.0
    ys = [oracle.init_stencil1d(n, r, world) for r in range(world)]
    for r in range(world):  # oracle exchange
        y = ys[r][0]
        if r > 0:
            y[:2] = ys[r - 1][0][-4:-2]
        if r < world - 1:
            y[-2:] = ys[r + 1][0][2:4]
    for r in range(world):
        y, act, scale = ys[r]
        e = oracle.err_norm(local.stencil1d(y, scale), act)
        assert e < oracle.err_norm_pass_bound(n)
        total += e
    assert total < 1e-5
    # a stale ghost cell is unmistakable
    y, act, scale = oracle.init_stencil1d(n, 1, world)
    assert oracle.err_norm(local.stencil1d(y, scale), act) > 1e4


@pytest.mark.parametrize("dim", [0, 1])
def test_stencil2d_oracle_single_rank(dim):
    z, act, scale = oracle.init_stencil2d(64, 48, dim, 0, 1)
    e = oracle.err_norm(local.stencil2d(z, dim, scale), act)
    assert e < 1e-8
    out = local.stencil2d(z, dim, scale, variant="generic")
    assert out.shape == act.shape


def test_cpu_init_helpers_match_oracle():
    n, world = 512, 3
    for r in range(world):
        y_ref, act_ref, _ = oracle.init_stencil1d(n, r, world)
        y = torch.empty(n + 4, dtype=F64)
        act = torch.empty(n, dtype=F64)
        dx = 8.0 / (n * world)
        local.init_stencil1d(y, act, n, r * n * dx, dx, r == 0, r == world - 1)
        assert torch.allclose(y, y_ref, rtol=0, atol=1e-12) and torch.allclose(act, act_ref, rtol=0, atol=1e-12)
    z = torch.empty(10 + 4, 8 + 4, dtype=F64)
    act = torch.empty(10, 8, dtype=F64)
    local.init_stencil2d(z, act, 8, 10, 2, 2, 0.0, 0.0, 0.5, 0.25, 3, 2.0, 4.0, 1 | 8, -7.0)
    assert float(z[5, 0]) == pytest.approx((-1.0) ** 3 + (0.75) ** 2)   # low-x ghost analytic
    assert float(z[5, 11]) == -7.0                                      # high-x ghost poisoned
    assert float(z[0, 5]) == -7.0 and float(z[13, 5]) == pytest.approx(1.5 ** 3 + 2.75 ** 2)
    assert float(z[0, 0]) == -7.0                                       # corner never analytic
    assert float(act[2, 3]) == pytest.approx(2.0 * 3 * 1.5 ** 2 + 4.0 * 2 * 0.5)


def test_gather_closed_form():
    n = 1000
    for r in range(6):
        assert float(oracle.gather_slot_values(r, n).sum()) == oracle.gather_slot_sum(r, n)
    assert oracle.gather_all_sum(4, n) == sum(float(oracle.gather_slot_values(r, n).sum()) for r in range(4))


def test_axis_sum_and_reductions_cpu():
    z = torch.rand(7, 5, dtype=F64)
    assert torch.allclose(local.axis_sum(z, 0), z.sum(dim=1))
    assert torch.allclose(local.axis_sum(z, 1, n_out=3), z.sum(dim=0)[:3])
    a, b = torch.rand(100, dtype=F64), torch.rand(100, dtype=F64)
    assert local.diff_norm(a, b) == pytest.approx(float(((a - b) ** 2).sum()) ** 0.5)
    assert float(local.total(a)) == pytest.approx(float(a.sum()))


# ------------------------------------------------------------------ golden strings
def test_golden_output_strings():
    assert report.rank_device(0, 4, 0, 8, 191_000_000_000) == "RANK[1/4] => DEVICE[1/8] mem=191000000000"
    assert report.memory_per_core(None) == "MEMORY_PER_CORE is not set"
    assert report.memory_per_core("4096") == "MEMORY_PER_CORE=4096"
    assert report.rank_sum_gt(2, 4, 2, 0x1B00, 524800.0) == "2/4 [2:0x00001b00] SUM = 524800.000000"
    assert report.daxpy_nvtx_banner(1, 8, 6291456, 50331648) == "1 nodes, 8 ranks, 6291456 elements each, total 50331648"
    assert report.gpu_memory(1 << 30, 2 << 30) == "GPU memory 1024.000 / 2048.000 (1024.000) MB"
    assert report.time_line(3, 8, "total", 1.23456) == "3/8 TIME total  : 1.235"
    assert report.time_line(3, 8, "kernel", 0.001) == "3/8 TIME kernel : 0.001"
    assert report.time_line(3, 8, "barrier", 0.0) == "3/8 TIME barrier: 0.000"
    assert report.time_line(3, 8, "gather", 0.5) == "3/8 TIME gather : 0.500"
    assert report.meminfo("d_x", -2) == "CUDA PreferredLocation of 'd_x' is UNMANAGED"
    assert report.meminfo("m_x", 0) == "CUDA PreferredLocation of 'm_x' is DEVICE (0)"
    assert report.meminfo("m_x", -1) == "CUDA PreferredLocation of 'm_x' is CPU (-1)"
    assert report.meminfo("h", -3) == "CUDA PreferredLocation of 'h' is NOT CUDA"
    assert report.stencil1d_banner(8, 33554432, 4194304) == ["n procs  = 8", "n_global = 33554432", "n_local  = 4194304"]
    assert report.exchange_time_s(0, 2, 0.000123) == "0/2 exchange time 0.00012300"
    assert report.stencil2d_banner(4, 4096, 524288, 1000, 10)[1:3] == ["n_global_deriv = 4096", "n_global_other = 524288"]
    assert report.test_deriv_line(0, "device", True, 1.5, 2e-7) == "TEST dim:0, device , buf:1; 1.50000000, err=0.00000020"
    assert report.test_deriv_line(1, "managed", False, 0.25, 0.0) == "TEST dim:1, managed, buf:0; 0.25000000, err=0.00000000"
    assert report.test_sum_line(1, "device", 0.003) == "TEST dim:1, device , buf:0; allreduce=0.00300000"
    assert report.oo_exchange_time(1, 0.5) == "1: exchange time 0.50000000 ms"
    assert report.oo_err_norm(1, 0x8086, 1e-6) == "1: [0x00008086] err_norm = 0.00000100"
    assert report.dev_bytes(8.6e9).endswith("GB") and report.dev_bytes(5e6).endswith("MB")
    assert report.cuda_error("x.cu", 12, "malloc", 2, "out of memory") == "x.cu(12): CUDA Error (malloc) 2: out of memory"
    assert report.mpi_error("f.cc", 35, 13) == "f.cc:35 error 13"
    assert "MEMORY_PER_CORE=" in report.mpienv_line(3, 4096)
    assert report.sycl_banner(2, 0, 524288, 2048, 1024, 100, 5, False)[-1] == "stage_host = 0"


def test_dist_arrays_world_of_one_cpu(monkeypatch):
    """The user-facing distributed-array objects on the CPU path (world of one)."""
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    from gpu_mpi_tests_b200.parallel.darray import DistArray1D, DistArray2D
    rt = runtime.init(use_cuda=False)
    a = DistArray1D(rt, 1 << 12)
    a.fill_analytic()
    a.exchange()
    d = a.derivative()
    assert d.shape == (1 << 12,) and a.err_norm() < 1e-6
    assert a.interior.shape == (1 << 12,) and a.local.shape == ((1 << 12) + 4,)
    for dims, ghosts in (((0,), (2, 0)), ((1,), (0, 2)), ((0, 1), (2, 2))):
        b = DistArray2D(rt, 64, 48, dims=dims)
        assert (b.gx, b.gy) == ghosts
        b.fill_analytic()
        b.exchange()
        out = b.derivative()
        assert out.shape == (48, 64) and b.err_norm() < 1e-8
        assert b.interior.shape == (48, 64)
        if 0 in dims:
            assert b.ghost(0).shape == (48, 2)
        if 1 in dims:
            assert b.ghost(3).shape == (2, 64)
    runtime.finalize()


@pytest.mark.parametrize("dim", [0, 1])
def test_slab_modes_use_the_reference_single_spacing(monkeypatch, dim):
    """dim0 / dim1 workloads use ONE grid spacing, 8 / n_global_deriv, for both coordinates like
    test_deriv (mpi_stencil2d_gt.cc:427-448): the initial field and the analytic derivative are those
    of the CPU oracle built the reference's way, and err= is the oracle's."""
    for k in ("WORLD_SIZE", "RANK", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
        monkeypatch.delenv(k, raising=False)
    from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
    rt = runtime.init(use_cuda=False)
    n_deriv, n_other = 32, 4096
    nx, ny = (n_deriv, n_other) if dim == 0 else (n_other, n_deriv)
    wl = Stencil2DWorkload(rt, nx, ny, mode=f"dim{dim}", impl="baseline", with_err=False)
    wl.setup()
    z_ref, actual, scale = oracle.init_stencil2d(n_deriv, n_other, dim, 0, 1)
    assert wl.dx == wl.dy == 8.0 / n_deriv
    assert torch.allclose(wl.z[0], z_ref, rtol=1e-14, atol=0.0)
    assert torch.allclose(wl.analytic(0), actual, rtol=1e-14, atol=0.0)
    wl.step()
    want = oracle.err_norm(oracle.stencil2d_1d_5(z_ref, dim, scale), actual)
    got = wl.err_norm(0)
    # the other axis reaches 4096 * 0.25 = 1024, so rounding against y^2 (x^3) dominates: both see it
    assert want > 1e-9 and abs(got - want) <= 0.5 * want + 1e-12
    runtime.finalize()


def test_native_engine_device_rule():
    """gmt_engine_default_device is the reference's set_rank_device rule (mpi_daxpy.cc:36-62);
    it makes no CUDA call, so it can be checked here through the C ABI."""
    from gpu_mpi_tests_b200.ops._lib import GmtError, have_native
    if not have_native():
        pytest.skip("library not built")
    from gpu_mpi_tests_b200.parallel.engine import NativeEngine2D
    from gpu_mpi_tests_b200.parallel.runtime import map_rank_to_device
    for n_ranks, n_dev in [(8, 8), (4, 8), (16, 8), (42, 6), (1, 1)]:
        for r in range(n_ranks):
            assert NativeEngine2D.default_device(n_ranks, r, n_dev) == map_rank_to_device(n_ranks, r, n_dev)[0]
    with pytest.raises(GmtError, match="not a multiple"):
        NativeEngine2D.default_device(9, 0, 8)


def test_check_header_overloads_compile(tmp_path):
    """include/gmt_check.hpp: CHECK/WARN for gmt codes and cudaError_t, and the compile-out switch
    (the reference's GPU_NO_CHECK_CALLS, cuda_error.h:16-26) -- compile-only, no GPU needed."""
    import shutil
    import subprocess
    from pathlib import Path
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not Path(nvcc).exists():
        pytest.skip("nvcc not found")
    src = tmp_path / "t.cpp"
    src.write_text('#include <cuda_runtime_api.h>\n#include "gmt_check.hpp"\n#include "gmt_b200.hpp"\n'
                   'int f(double* p) {\n  GMT_CHECK("gmt call", gmt_set_device(0));\n'
                   '  int w = GMT_WARN("cuda call", cudaSetDevice(0));\n  GMT_PTRINFO(p);\n  GMT_MEMINFO(p, 8);\n'
                   '  gmt::NvtxRange r("x");\n  return w;\n}\n')
    inc = str(Path(__file__).resolve().parents[1] / "include")
    for extra in ([], ["-DGMT_NO_CHECK_CALLS"]):
        res = subprocess.run([nvcc, "-std=c++17", "-I", inc, *extra, "-c", str(src), "-o", str(tmp_path / "t.o")],
                             capture_output=True, text=True)
        assert res.returncode == 0, res.stderr


def test_native_mpienv_program():
    """native/gmt_mpienv.cpp needs no GPU: same line as cli.mpienv (mpienv.f90:32)."""
    import os
    import subprocess
    from pathlib import Path
    from gpu_mpi_tests_b200.utils import report
    exe = Path(__file__).resolve().parents[1] / "gpu_mpi_tests_b200" / "_native" / "bin" / "gmt_mpienv"
    if not exe.exists():
        pytest.skip("native programs not built")
    res = subprocess.run([str(exe), "--ranks", "3"], capture_output=True, text=True, env=dict(os.environ, MEMORY_PER_CORE="12"))
    assert res.returncode == 0
    assert res.stdout.splitlines() == [report.mpienv_line(r, 12) for r in range(3)]


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` must print one JSON line {"impl": "reference", "unavailable": ...}
    and exit 0 (the reference is a flat tree of MPI programs that cannot be installed here)."""
    import json
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    res = subprocess.run([sys.executable, str(root / "bench.py"), "--impl", "reference", "--gpus", "1", "--steps", "2",
                          "--warmup", "3"], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stderr
    lines = [ln for ln in res.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and isinstance(rec["unavailable"], str) and "\n" not in rec["unavailable"]


def test_bench_json_keys_are_the_contract():
    """Static check of the keys bench.py emits (the GPU run itself is the driver's job)."""
    from pathlib import Path
    src = (Path(__file__).resolve().parents[1] / "bench.py").read_text()
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
                "vs_baseline", "dtype", "data", "config", "clocks", "e2e", "h2d_bytes_per_step", "d2h_bytes_per_step",
                "gpu_launches", "sm_mhz", "sm_max_mhz", "reasons"):
        assert f'"{key}"' in src, key


def test_public_header_is_plain_c(tmp_path):
    """include/gmt_b200.h is the C ABI: it must compile as C99 with no CUDA header around."""
    import shutil
    import subprocess
    from pathlib import Path
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("gcc not found")
    src = tmp_path / "abi.c"
    src.write_text('#include "gmt_b200.h"\nint main(void) { gmt_halo2d_rank r; gmt_coll_rank c; (void)r; (void)c; '
                   'return gmt_version() > 0 ? 0 : 1; }\n')
    inc = str(Path(__file__).resolve().parents[1] / "include")
    res = subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", inc, "-fsyntax-only", str(src)],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stderr


def test_ctypes_structs_match_the_c_header(tmp_path):
    """The Python layer fills gmt_halo1d_rank / gmt_halo2d_rank / gmt_coll_rank / gmt_device_props through
    ctypes mirrors; a field added on one side only would corrupt memory silently.  Compile a C program
    that prints sizeof and every offsetof, and compare with the ctypes layout."""
    import shutil
    import subprocess
    from pathlib import Path
    from gpu_mpi_tests_b200.ops import _lib
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("gcc not found")
    pairs = {"gmt_halo1d_rank": _lib.Halo1dRank, "gmt_halo2d_rank": _lib.Halo2dRank, "gmt_coll_rank": _lib.CollRank,
             "gmt_device_props": _lib.DeviceProps}
    rename = {"inp": "in"}   # `in` is a keyword-ish name on the Python side
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "gmt_b200.h"', 'int main(void) {']
    for cname, cls in pairs.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _t in cls._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {rename.get(fname, fname)}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "layout.c"
    src.write_text("\n".join(lines) + "\n")
    exe = tmp_path / "layout"
    inc = str(Path(__file__).resolve().parents[1] / "include")
    res = subprocess.run([gcc, "-std=c99", "-I", inc, str(src), "-o", str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    out = subprocess.run([str(exe)], capture_output=True, text=True).stdout.split("\n")
    got = {tuple(ln.split()[:2]): int(ln.split()[2]) for ln in out if ln.strip()}
    for cname, cls in pairs.items():
        assert got[(cname, "size")] == C_sizeof(cls), cname
        for fname, _t in cls._fields_:
            assert got[(cname, fname)] == getattr(cls, fname).offset, (cname, fname)


def C_sizeof(cls) -> int:
    import ctypes
    return ctypes.sizeof(cls)


def test_header_library_and_binding_agree_on_entry_points():
    """Every function the header declares is exported by the built library, and every name the
    ctypes table binds is declared in the header (argument counts included)."""
    import re
    import subprocess
    from pathlib import Path
    from gpu_mpi_tests_b200.ops import _lib
    if not _lib.have_native():
        pytest.skip("library not built")
    root = Path(__file__).resolve().parents[1]
    text = (root / "include" / "gmt_b200.h").read_text()
    text = re.sub(r"/\*.*?\*/", " ", text, flags=re.S)
    decls = {}
    for m in re.finditer(r"\b(gmt_[a-z0-9_]+)\s*\(([^;{]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        decls[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    assert len(decls) > 70
    exported = subprocess.run(["nm", "-D", "--defined-only", str(_lib.LIB_PATH)], capture_output=True, text=True).stdout
    exported = {ln.split()[-1] for ln in exported.splitlines() if ln.strip()}
    missing = sorted(set(decls) - exported)
    assert not missing, f"declared but not exported: {missing}"
    for name, (_res, argtypes) in _lib._SIGS.items():
        assert name in decls, f"{name} is bound in _lib.py but not declared in the header"
        assert len(argtypes) == decls[name], f"{name}: {len(argtypes)} ctypes arguments, {decls[name]} in the header"


def test_tile_order_is_a_bijection(tmp_path):
    """csrc/tile_order.h on the host: the fused 2-D kernel's counter -> tile map visits every tile
    exactly once for every grid shape / neighbour combination / order flag, never schedules a tile
    that needs no flag after one that does, and degenerates to memory order when nothing is deferred."""
    import shutil
    import subprocess
    from pathlib import Path
    gxx = shutil.which("g++")
    if not gxx:
        pytest.skip("g++ not found")
    root = Path(__file__).resolve().parents[1]
    exe = tmp_path / "tile_order"
    res = subprocess.run([gxx, "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I", str(root / "csrc"),
                          str(root / "tests" / "native" / "test_tile_order.cpp"), "-o", str(exe)],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    run = subprocess.run([str(exe)], capture_output=True, text=True)
    assert run.returncode == 0 and run.stdout.startswith("ok "), run.stdout


def test_pure_host_entry_points_work_without_a_gpu():
    """Entry points that only compute sizes must work on a machine with no GPU (plan set-up code
    calls them before touching a device)."""
    from gpu_mpi_tests_b200.ops import _lib
    if not _lib.have_native():
        pytest.skip("library not built")
    lib = _lib.lib
    assert lib.gmt_version() >= 100
    for nx, ny in [(1024, 524288), (16384, 16384), (8, 4), (4100, 1025)]:
        assert lib.gmt_halo2d_flag_count(0, nx, ny) == lib.gmt_halo2d_flag_count(1, nx, ny) == -(-ny // 1024)
        assert lib.gmt_halo2d_flag_count(2, nx, ny) == lib.gmt_halo2d_flag_count(3, nx, ny) == -(-nx // 1024)
    assert lib.gmt_reduce_scratch_bytes() >= 4096
    assert lib.gmt_noalloc_max_n() == 33_554_432 and lib.gmt_noalloc_scratch_bytes() == 65_792
    n = __import__("ctypes").c_int(-1)
    lib.gmt_device_count(__import__("ctypes").byref(n))    # no driver / no device -> 0, not an error
    assert n.value >= 0


def test_coverage_map_points_at_real_files_and_tests():
    """COVERAGE.md maps every inventory item of SURVEY.md section 2 to code and tests: every file path
    and every test it names must exist."""
    import re
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    text = (root / "COVERAGE.md").read_text()
    tests = {}
    for f in (root / "tests").glob("test_*.py"):
        tests[f.stem] = set(re.findall(r"^def (test_\w+)", f.read_text(), flags=re.M))
    current = None
    checked = 0
    for tok in re.findall(r"`([^`]+)`", text):
        m = re.fullmatch(r"(test_\w+)?::(test_[\w*]+|\*)(?: / _\w+)*", tok.strip())
        if m:
            current = m.group(1) or current
            assert current in tests, f"unknown test file in {tok!r}"
            name = m.group(2)
            if name == "*":
                continue
            if "*" in name:
                assert any(re.fullmatch(name.replace("*", r"\w*"), t) for t in tests[current]), tok
            else:
                assert name in tests[current], f"{current}::{name} does not exist"
            checked += 1
            continue
        for path in re.findall(r"\b((?:py|csrc|native|include|scripts|tests)/[\w./-]+\.(?:hpp|cuh|cpp|py|cu|sh|h)\b)", tok):
            real = path.replace("py/", "gpu_mpi_tests_b200/", 1) if path.startswith("py/") else path
            assert (root / real).exists(), f"{path} named in COVERAGE.md does not exist"
            checked += 1
    assert checked > 150


def test_reported_numbers_match_the_committed_records():
    """README.md / BASELINE.md quote the flagship numbers of profiles/bench_n{1,2,4,8}.json: the quoted
    GLUP/s must be the recorded ones (rounded), and the records must satisfy the bench contract, carry the
    same-box library-only arm (vs_baseline) and every component of BASELINE.json's metric."""
    import json
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    readme, baseline = (root / "README.md").read_text(), (root / "BASELINE.md").read_text()
    for n in (1, 2, 4, 8):
        rec = json.loads((root / "profiles" / f"bench_n{n}.json").read_text().strip().splitlines()[-1])
        assert rec["metric"] == "stencil2d_glups" and rec["n_gpus"] == n and rec["warmup"] >= 3 and rec["dtype"] == "fp64"
        assert rec["e2e"]["h2d_bytes_per_step"] > 0 and rec["e2e"]["d2h_bytes_per_step"] > 0
        assert rec["e2e"]["value"] < rec["value"]                      # end to end cannot beat the device time
        assert rec["e2e"]["value"] > 0.975 * rec["value"]              # ... and the pipelined step is within 2.5% of it
        assert rec["gpu_launches"] == rec["steps"]                     # one fused kernel per step
        assert not set(rec["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
        assert 0.9 < rec["roofline"]["frac_of_measured_copy_peak"] < 1.1
        assert rec["ghost_mismatch"] == 0
        arm = rec["baseline_arm"]
        assert rec["vs_baseline"] == pytest.approx(rec["value"] / arm["value"]) and arm["e2e_ratio"] > 4.0
        if n > 1:
            assert rec["vs_baseline"] > 1.1
        ex = rec["extra"]
        assert "errors" not in ex
        for key in ("halo_exchange_us", "daxpy_sum", "allreduce_8KiB_us", "barrier_us", "allgather", "stencil1d_2p27",
                    "stencil2d_ref_dim0", "stencil2d_ref_dim1", "stencil2d_periodic4"):
            assert key in ex, key
        quoted = f"{rec['value']:.0f}"
        assert quoted in baseline, f"{quoted} GLUP/s (N={n}) is not in BASELINE.md"
        if n in (1, 8):
            assert quoted in readme, f"{quoted} GLUP/s (N={n}) is not in README.md"


def test_baseline_tables_are_the_ones_generated_from_the_records():
    """BASELINE.md section 4a is generated by tools/baseline_rows.py from profiles/bench_n*.json and
    profiles/sweep_n*.jsonl: every generated table row must be in the document verbatim."""
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parents[1]
    out = subprocess.run([sys.executable, str(root / "tools" / "baseline_rows.py")], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    doc = (root / "BASELINE.md").read_text()
    rows = [ln for ln in out.stdout.splitlines() if ln.startswith("|")]
    assert len(rows) > 30
    missing = [ln for ln in rows if ln not in doc]
    assert not missing, "BASELINE.md section 4a is out of date with the records:\n" + "\n".join(missing[:5])


def test_periodic_grid_neighbours_are_mutual_and_everybody_has_four():
    from gpu_mpi_tests_b200.parallel.decomp import Grid2D
    for px, py in ((1, 1), (2, 1), (4, 2), (3, 3), (3, 4)):
        for r in range(px * py):
            g = Grid2D(px, py, r, periodic=True)
            nb = g.neighbours()
            assert all(0 <= n < px * py for n in nb) and g.physical_mask() == 0
            for side, n in enumerate(nb):
                assert Grid2D(px, py, n, periodic=True).neighbour(side ^ 1) == r
    g = Grid2D(4, 2, 0)
    assert g.neighbours() == [-1, 1, -1, 4] and g.physical_mask() == 0b0101
