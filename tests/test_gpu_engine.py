"""Native exchange engine (csrc/engine.cu) and the native program built on it."""
import re
import subprocess
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = Path(__file__).resolve().parents[1]
EXE = ROOT / "gpu_mpi_tests_b200" / "_native" / "bin" / "gmt_stencil2d"


@pytest.fixture(scope="module", autouse=True)
def _programs(native):
    """The native programs are build outputs (git-ignored): build them when the tree has none."""
    wanted = [p.stem for p in (ROOT / "native").glob("*.cpp")]
    if any(not (EXE.parent / name).exists() for name in wanted):
        from gpu_mpi_tests_b200._native.build import build_programs
        build_programs()


def close(a, b, rtol=1e-10):
    return bool(((a - b).abs() <= rtol * (1.0 + b.abs())).all())


@pytest.mark.parametrize("px,py,gx,gy", [(2, 2, 2, 2), (4, 1, 2, 0), (1, 3, 0, 2), (1, 1, 2, 2), (3, 3, 2, 2)])
def test_engine_matches_python_plan(px, py, gx, gy):
    """Same grid through the C++ engine (all ranks on GPU 0, one cooperative launch) and through
    the Python plan on virtual ranks: same result arrays, same landed ghosts, same error norm."""
    from gpu_mpi_tests_b200.parallel.engine import NativeEngine2D
    from gpu_mpi_tests_b200.parallel.halo import FusedStencil2D
    from gpu_mpi_tests_b200.parallel.symm import VirtualSpace
    dev = torch.device("cuda", 0)
    nx, ny = 1536, 320
    world = px * py
    eng = NativeEngine2D(px, py, nx, ny, gx, gy, with_err=True, devices=[0] * world, spin_limit=1 << 26)
    plan = FusedStencil2D(VirtualSpace(world, dev), px, py, nx, ny, gx, gy, with_err=True, spin_limit=1 << 26)
    eng.init_analytic(poison=1e30)
    plan.init_analytic(poison=1e30)
    for _ in range(3):   # both parities and a repeat
        eng.step()
        plan.step()
    eng.synchronize()
    torch.cuda.synchronize()
    eng.check()
    plan.check_status()
    assert eng.world == world and eng.n_device_groups == 1
    for r in range(world):
        assert close(eng.array(r, "out"), plan.out[r]), f"rank {r}"
        assert torch.equal(eng.array(r, "z"), plan.z[r]), f"rank {r}: ghosts"
        assert eng.err_norm(r) == pytest.approx(plan.err_norm(r), rel=0.1, abs=1e-9)
        assert eng.err_norm(r) < 1e-3
    # the exchange alone through the engine: ghosts poisoned again, then refreshed without a sweep
    for r in range(world):
        z = eng.array(r, "z")
        if gx:
            z[:, :2] = 1e30
            z[:, -2:] = 1e30
        if gy:
            z[:2, :] = 1e30
            z[-2:, :] = 1e30
    torch.cuda.synchronize()
    eng.exchange()
    eng.synchronize()
    eng.check()
    for r in range(world):
        z, ref = eng.array(r, "z"), plan.z[r]
        cx, cy = r % px, r // px
        ys, xs = slice(gy, gy + ny), slice(gx, gx + nx)
        if gx and cx > 0:
            assert torch.equal(z[ys, :2], ref[ys, :2])
        if gx and cx < px - 1:
            assert torch.equal(z[ys, -2:], ref[ys, -2:])
        if gy and cy > 0:
            assert torch.equal(z[:2, xs], ref[:2, xs])
        if gy and cy < py - 1:
            assert torch.equal(z[-2:, xs], ref[-2:, xs])
    assert eng.time_exchanges(3, 1) > 0.0
    eng.close()


def test_engine_timing_and_argument_checks():
    from gpu_mpi_tests_b200.ops._lib import GmtError
    from gpu_mpi_tests_b200.parallel.engine import NativeEngine2D
    eng = NativeEngine2D(2, 1, 2048, 1024, 2, 2, devices=[0, 0])
    eng.init_analytic()
    ms = eng.time_steps(5, 2)
    assert 0.0 < ms < 100.0
    assert max(eng.err_norm(r) for r in range(2)) < 1e-3
    eng.close()
    with pytest.raises(GmtError):
        NativeEngine2D(2, 1, 2048, 1024, 0, 2, devices=[0, 0])   # px > 1 without x ghosts
    with pytest.raises(GmtError):
        NativeEngine2D(1, 1, 2050, 1024, 2, 2)                  # nx % 4
    with pytest.raises(GmtError):
        NativeEngine2D(1, 1, 2048, 1024, 2, 2, devices=[99])


def test_engine_across_gpus():
    """One rank per GPU, peers wired through CUDA peer access inside one process."""
    from gpu_mpi_tests_b200.parallel.engine import NativeEngine2D
    n_dev = torch.cuda.device_count()
    if n_dev < 2:
        pytest.skip("needs 2 GPUs")
    world = 4 if n_dev >= 4 else 2
    px, py = (2, 2) if world == 4 else (2, 1)
    eng = NativeEngine2D(px, py, 4096, 2048, 2, 2)
    assert eng.n_device_groups == world and [eng.rank_device(r) for r in range(world)] == list(range(world))
    eng.init_analytic(poison=1e30)
    ms = eng.time_steps(10, 3)
    assert ms > 0
    for r in range(world):
        assert eng.err_norm(r) < 1e-3, f"rank {r}"     # a stale ghost would give ~1e30
    before = torch.cuda.current_device()
    eng.step()
    eng.synchronize()
    assert torch.cuda.current_device() == before
    eng.close()


def test_native_program_runs():
    assert EXE.exists(), f"{EXE} missing: run python -m gpu_mpi_tests_b200._native.build"
    res = subprocess.run([str(EXE), "256", "3", "--ranks", "4", "--other", "4096", "--tile", "1024"],
                         capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    out = res.stdout
    assert "n procs        = 4" in out and "n_global_deriv = 1024" in out
    tests = re.findall(r"^TEST (dim:0|dim:1|grid ), device , fused; ([0-9.]+), err=([0-9.eE+-]+)$", out, re.M)
    assert [t[0] for t in tests] == ["dim:0", "dim:1", "grid "], out
    for _label, secs, err in tests:
        assert float(secs) > 0 and float(err) < 1e-3


def test_native_program_rejects_uneven_oversubscription():
    n_dev = torch.cuda.device_count()
    if n_dev < 2:
        pytest.skip("with one GPU every rank count divides")
    res = subprocess.run([str(EXE), "256", "1", "--ranks", str(n_dev + 1)], capture_output=True, text=True, timeout=120)
    assert res.returncode == 1
    assert "not a multiple of number of GPUs" in res.stderr


def test_native_daxpy_program():
    """daxpy.cu's contract: n element lines then SUM = 524800.000000 (daxpy.cu:84-87)."""
    exe = EXE.parent / "gmt_daxpy"
    res = subprocess.run([str(exe)], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stdout + res.stderr
    lines = res.stdout.strip().splitlines()
    assert len(lines) == 1025 and lines[0] == "1.000000" and lines[1023] == "1024.000000"
    assert lines[-1] == "SUM = 524800.000000"
    res = subprocess.run([str(exe), "1048576", "--quiet"], capture_output=True, text=True, timeout=120)
    assert res.stdout.strip() == "SUM = %f" % (1048576 * 1048577 / 2)


def _coll_devices(world):
    """All ranks on GPU 0 on a 1-GPU box; spread by the block rule when there are more GPUs."""
    n_dev = torch.cuda.device_count()
    if n_dev >= 2 and world % 2 == 0:
        return None if world % n_dev == 0 or world <= n_dev else [r * 2 // world for r in range(world)]
    return [0] * world


@pytest.mark.parametrize("world", [1, 2, 4])
def test_native_collectives(world):
    from gpu_mpi_tests_b200.parallel.engine import NativeCollectives
    n = 3000
    eng = NativeCollectives(world, buf_doubles=n * world, stage_cap=2048, devices=_coll_devices(world),
                            spin_limit=1 << 26)
    devs = [torch.device("cuda", eng.rank_device(r)) for r in range(world)]
    gen = torch.Generator().manual_seed(5)
    slices = [torch.randn(n, dtype=torch.float64, generator=gen) for _ in range(world)]
    want = torch.cat(slices)
    # in-place allgather
    for r in range(world):
        with torch.cuda.stream(eng.stream(r)):
            b = eng.buffer(r)
            b.fill_(-7.0)
            b[r * n:(r + 1) * n].copy_(slices[r].to(devs[r]))
    eng.synchronize()
    eng.allgather(n)
    eng.allgather(n)   # twice: epochs / rounds must not interfere
    eng.synchronize()
    eng.check()
    for r in range(world):
        assert torch.equal(eng.buffer(r).cpu(), want), f"rank {r}"
    # out-of-place source + gather to root
    srcs = [(slices[r] * 2).to(devs[r]) for r in range(world)]
    for r in range(world):
        eng.bind(r, inp=srcs[r])
    eng.allgather(n, from_in=True)
    eng.synchronize()
    for r in range(world):
        assert torch.equal(eng.buffer(r).cpu(), want * 2), f"rank {r}"
    # all-reduce (both algorithms), identical bits on every rank
    vecs = [torch.randn(1024, dtype=torch.float64, generator=gen) for _ in range(world)]
    total = torch.stack(vecs).sum(0)
    for algo in (0, 1, 3):
        ins = [v.to(d) for v, d in zip(vecs, devs)]
        outs = [torch.zeros(1024, dtype=torch.float64, device=d) for d in devs]
        for r in range(world):
            eng.bind(r, inp=ins[r], out=outs[r])
        eng.allreduce(1024, algo)
        eng.synchronize()
        eng.check()
        for r in range(world):
            assert torch.allclose(outs[r].cpu(), total, rtol=1e-13, atol=1e-13)
            assert torch.equal(outs[r].cpu(), outs[0].cpu())
    # fused daxpy + global sum
    m = 1 << 18
    xs = [torch.randn(m, dtype=torch.float64, generator=gen).to(d) for d in devs]
    ys = [torch.randn(m, dtype=torch.float64, generator=gen).to(d) for d in devs]
    ref = sum(float((2.0 * x.cpu() + y.cpu()).sum()) for x, y in zip(xs, ys))
    sums = [torch.zeros(8, dtype=torch.float64, device=d) for d in devs]
    torch.cuda.synchronize()
    for r in range(world):
        eng.bind(r, inp=xs[r], out=sums[r], y=ys[r])
    eng.daxpy_allreduce(m, 2.0)
    eng.barrier()
    eng.synchronize()
    eng.check()
    for r in range(world):
        assert float(sums[r][0]) == pytest.approx(ref, rel=1e-11, abs=1e-8)
    ms = eng.timed(lambda: eng.barrier(), iters=10)
    assert 0 < ms < 50
    from gpu_mpi_tests_b200.ops._lib import GmtError
    with pytest.raises(GmtError):
        eng.allgather(n + 1)
    with pytest.raises(GmtError):
        eng.allreduce(4096, 0)
    eng.close()


@pytest.mark.parametrize("ranks", [1, 4])
def test_native_daxpy_gather_program(ranks):
    n_dev = torch.cuda.device_count()
    if ranks > n_dev and ranks % n_dev:
        pytest.skip("rank count does not divide")
    exe = EXE.parent / "gmt_daxpy_gather"
    n = 1 << 20
    res = subprocess.run([str(exe), "--n", str(n), "--ranks", str(ranks), "--barrier"], capture_output=True, text=True,
                         timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    out = res.stdout
    assert f"1 nodes, {ranks} ranks, {n} elements each, total {n * ranks}" in out
    for r in range(ranks):
        assert f"{r}/{ranks} SUM = {(n + 1) / 2:f}" in out
        assert f"{r}/{ranks} ALLSUM = {ranks * (n + 1) / 2:f}" in out
        assert re.search(rf"^{r}/{ranks} TIME gather : [0-9.]+$", out, re.M)


@pytest.mark.parametrize("args", [[], ["--gather", "1"]])
def test_native_gatherinplace_program(args):
    exe = EXE.parent / "gmt_gatherinplace"
    res = subprocess.run([str(exe), str(1 << 20), "--ranks", "4", *args], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr   # the program checks the closed-form sums itself
    rows = re.findall(r"^ (\d) / 4   ([0-9.]+)   ([0-9.]+)$", res.stdout, re.M)
    assert len(rows) == 4


def test_native_stencil1d_program():
    """Hand-wired descriptors against the C ABI (native/gmt_stencil1d.cpp), on every GPU."""
    exe = EXE.parent / "gmt_stencil1d"
    res = subprocess.run([str(exe), "8", "--iters", "5"], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr   # the program itself checks err_norm and status
    W = torch.cuda.device_count()
    assert f"n procs  = {W}\nn_global = {8 << 20}\nn_local  = {(8 << 20) // W}" in res.stdout
    for r in range(W):
        assert re.search(rf"^{r}/{W} exchange time [0-9.]+$", res.stdout, re.M)
        assert re.search(rf"^{r}/{W} \[{r}:0x[0-9a-f]{{8}}\] err_norm = [0-9.]+$", res.stdout, re.M)
    res = subprocess.run([str(exe), "--points", str(1000 * W + 1), "--ranks", str(W)], capture_output=True,
                         text=True, timeout=60)
    if W > 1:
        assert res.returncode == 1 and "must be divisor of domain size" in res.stdout


@pytest.mark.parametrize("ranks", [1, 4])
def test_native_daxpy_allreduce_program(ranks):
    """Fused daxpy + global sum through the C++ conveniences (include/gmt_b200.hpp)."""
    n_dev = torch.cuda.device_count()
    if ranks > n_dev and ranks % n_dev:
        pytest.skip("rank count does not divide")
    exe = EXE.parent / "gmt_daxpy_allreduce"
    n = 300000
    res = subprocess.run([str(exe), "--n", str(n), "--ranks", str(ranks), "--iters", "20"], capture_output=True,
                         text=True, timeout=300)
    assert res.returncode == 0, res.stdout + res.stderr
    for r in range(ranks):
        assert f"{r}/{ranks} ALLSUM = {ranks * (n + 1) / 2:f}" in res.stdout
    assert re.search(r"[0-9.]+ us per call", res.stdout)


def test_native_mpi_daxpy_program():
    """mpi_daxpy's contract (mpi_daxpy.cc:36-62,101-157): rank->device lines (1-based, memory share),
    MEMORY_PER_CORE echo, preferred-location report for device/managed/host pointers, SUM per rank."""
    import os
    exe = EXE.parent / "gmt_mpi_daxpy"
    n_dev = torch.cuda.device_count()
    ranks = 2 * n_dev
    env = dict(os.environ, MEMORY_PER_CORE="42")
    res = subprocess.run([str(exe), "--ranks", str(ranks)], capture_output=True, text=True, timeout=120, env=env)
    assert res.returncode == 0, res.stdout + res.stderr
    out = res.stdout
    assert "MEMORY_PER_CORE=42" in out
    for r in range(ranks):
        assert re.search(rf"^RANK\[{r + 1}/{ranks}\] => DEVICE\[{r // 2 + 1}/{n_dev}\] mem=\d+$", out, re.M)
        assert f"{r}/{ranks} SUM = 524800.000000" in out
    assert "CUDA PreferredLocation of 'd_x' is UNMANAGED" in out
    assert re.search(r"CUDA PreferredLocation of 'm_x' is (CPU|DEVICE|INVALID) \(-?\d+\)", out)
    assert "CUDA PreferredLocation of 'x' is NOT CUDA" in out
    assert re.search(r"CUDA pointer d_x \([0-9a-f]+\): Device", out) and re.search(r"CUDA pointer m_x \([0-9a-f]+\): Managed", out)
    if n_dev > 1:
        res = subprocess.run([str(exe), "--ranks", str(n_dev + 1)], capture_output=True, text=True, timeout=60)
        assert res.returncode == 1 and "not a multiple of number of GPUs" in res.stdout


def test_native_mpi_daxpy_gt_program():
    """mpi_daxpy_gt's line (mpi_daxpy_gt.cc:92): r/N [dev:0x<packed PCI id>] SUM = 524800.000000, two ranks per GPU."""
    exe = EXE.parent / "gmt_mpi_daxpy_gt"
    n_dev = torch.cuda.device_count()
    ranks = 2 * n_dev
    res = subprocess.run([str(exe), "--ranks", str(ranks)], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stdout + res.stderr
    lines = res.stdout.strip().splitlines()
    assert len(lines) == ranks
    for r, ln in enumerate(lines):
        assert re.fullmatch(rf"{r}/{ranks} \[{r // 2}:0x[0-9a-f]{{8}}\] SUM = 524800\.000000", ln), ln


@pytest.mark.parametrize("oo", [False, True])
def test_native_stencil2d_sycl_programs(oo):
    """The two SYCL programs' native counterpart: parameter block, per-rank exchange time (the exchange
    alone, device-timed) and err_norm lines in each program's own format (mpi_stencil2d_sycl.cc:422-434,530,550;
    mpi_stencil2d_sycl_oo.cc:575-581,679,700); two ranks (one GPU: both on it, cooperative launch)."""
    exe = EXE.parent / "gmt_stencil2d_sycl"
    F8 = r"\d+\.\d{8}"
    if oo:
        args = ["--oo", "2", "0", "6", "--ranks", "2"]
    else:
        args = ["512", "1", "6", "--ranks", "2", "--ny", "4096"]
    res = subprocess.run([str(exe), *args], capture_output=True, text=True, timeout=180)
    assert res.returncode == 0, res.stdout + res.stderr
    out = res.stdout
    assert "n procs    = 2" in out and "n_iter     = 6" in out and "n_warmup   = 5" in out
    if oo:
        assert "n_global   = 2048" in out and "n_local    = 1024" in out and "stage_host = 0" in out
        for r in range(2):
            assert re.search(rf"^{r}: exchange time {F8} ms$", out, re.M), out
            m = re.search(rf"^{r}: \[0x[0-9a-f]{{8}}\] err_norm = ({F8})$", out, re.M)
            assert m and float(m.group(1)) < 1e-3, out
    else:
        assert "ny         = 4096" in out and "nx_global  = 1024" in out and "nx_local   = 512" in out
        assert "stage_host = 1" in out and re.search(r"^dev bytes  = \d+\.\d{3} (MB|GB)$", out, re.M)
        for r in range(2):
            assert re.search(rf"^{r} [0-9a-f]{{4}}:[0-9a-f]{{2}}:[0-9a-f]{{2}}\.0\([0-9a-f]{{16}}\)$", out, re.M), out
            assert re.search(rf"^{r}/2 exchange time {F8} ms$", out, re.M), out
            m = re.search(rf"^{r}/2 \[\d+:0x[0-9a-f]{{8}}\] err_norm = ({F8})$", out, re.M)
            assert m and float(m.group(1)) < 1e-3, out
    # divisor check of the square-domain form
    if oo:
        res = subprocess.run([str(exe), "--oo", "1", "0", "2", "--ranks", "3"], capture_output=True, text=True, timeout=60)
        if torch.cuda.device_count() in (1, 3):
            assert res.returncode == 1 and "must be divisor of domain size" in res.stdout
