"""Golden-output tests: run each CLI program (CPU/gloo flavour here; the GPU flavour is run
by test_cli_gpu below) and match the exact line formats of the reference."""
import os
import re
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]
_port = [29800 + (os.getpid() % 150)]


def run_cli(module, *args, nproc=1, env_extra=None, cpu=True, timeout=600, expect_rc=0, with_stderr=False):
    env = dict(os.environ)
    if cpu:
        env["GMT_FORCE_CPU"] = "1"
    env.update(env_extra or {})
    if nproc == 1:
        for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "LOCAL_WORLD_SIZE"):
            env.pop(k, None)
        cmd = [sys.executable, "-m", f"gpu_mpi_tests_b200.cli.{module}", *map(str, args)]
    else:
        _port[0] += 1
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
               "--master-addr", "127.0.0.1", "--master-port", str(_port[0]), "-m",
               f"gpu_mpi_tests_b200.cli.{module}", *map(str, args)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=str(ROOT))
    assert res.returncode == expect_rc, f"{module}: rc={res.returncode}\n{res.stdout[-2000:]}\n{res.stderr[-3000:]}"
    lines = [ln for ln in res.stdout.splitlines() if ln.strip()]
    return (lines, res.stderr.splitlines()) if with_stderr else lines


def match_all(lines, patterns):
    """Every pattern matches at least one line, in order of first appearance not required."""
    for pat in patterns:
        assert any(re.fullmatch(pat, ln) for ln in lines), f"no line matches {pat!r} in\n" + "\n".join(lines)


F = r"-?\d+\.\d{6}"
F8 = r"\d+\.\d{8}"


def test_daxpy_prints_every_element_and_sum():
    lines = run_cli("daxpy")
    assert len(lines) == 1025
    assert lines[0] == "1.000000" and lines[1023] == "1024.000000"
    assert lines[-1] == "SUM = 524800.000000"
    assert run_cli("daxpy_nvtx", "--quiet") == ["SUM = 524800.000000"]


def test_mpi_daxpy_and_gt_world2():
    lines = run_cli("mpi_daxpy", nproc=2)
    match_all(lines, [r"MEMORY_PER_CORE is not set", r"RANK\[1/2\] => DEVICE\[1/2\] mem=\d+",
                      r"RANK\[2/2\] => DEVICE\[2/2\] mem=\d+", r"CUDA PreferredLocation of 'm_y' is .+",
                      r"0/2 SUM = 524800\.000000", r"1/2 SUM = 524800\.000000"])
    lines = run_cli("mpi_daxpy_gt", nproc=2)
    match_all(lines, [r"0/2 \[0:0x[0-9a-f]{8}\] SUM = 524800\.000000", r"1/2 \[1:0x[0-9a-f]{8}\] SUM = 524800\.000000"])


def test_mpi_daxpy_nvtx_world2_closed_forms():
    """BASELINE config 1: fp64 daxpy + global sum, n = 2^20 per rank, world_size = 2, gloo."""
    lines = run_cli("mpi_daxpy_nvtx", "--n-per-node", 2 * (1 << 20), "--barrier", nproc=2,
                    env_extra={"MEMORY_PER_CORE": "2048"})
    match_all(lines, [r"1 nodes, 2 ranks, 1048576 elements each, total 2097152", r"MEMORY_PER_CORE=2048",
                      r"GPU memory \d+\.\d{3} / \d+\.\d{3} \(\d+\.\d{3}\) MB",
                      r"0/2 SUM = 524288\.500000", r"1/2 SUM = 524288\.500000",
                      r"0/2 ALLSUM = 1048577\.000000", r"1/2 ALLSUM = 1048577\.000000",
                      r"0/2 TIME total  : \d+\.\d{3}", r"1/2 TIME kernel : \d+\.\d{3}",
                      r"0/2 TIME barrier: \d+\.\d{3}", r"1/2 TIME gather : \d+\.\d{3}"])


def test_mpienv():
    lines = run_cli("mpienv", nproc=2, env_extra={"MEMORY_PER_CORE": "4096"})
    match_all(lines, [r" rank\s+0\s+MEMORY_PER_CORE=\s+4096", r" rank\s+1\s+MEMORY_PER_CORE=\s+4096"])


def test_mpigatherinplace_world4():
    lines = run_cli("mpigatherinplace", 1000, nproc=4)
    from gpu_mpi_tests_b200.utils import oracle
    asum = oracle.gather_all_sum(4, 1000)
    for r in range(4):
        want = r"\s+%d /\s+4\s+%s\s+%s" % (r, re.escape("%.8E" % oracle.gather_slot_sum(r, 1000)), re.escape("%.8E" % asum))
        match_all(lines, [want])


def test_mpi_stencil_gt_world2_and_divisor_error():
    lines = run_cli("mpi_stencil_gt", 1, nproc=2)
    match_all(lines, [r"n procs  = 2", r"n_global = 1048576", r"n_local  = 524288",
                      rf"0/2 exchange time {F8}", rf"1/2 exchange time {F8}",
                      rf"0/2 \[0:0x[0-9a-f]{{8}}\] err_norm = {F8}", rf"1/2 \[1:0x[0-9a-f]{{8}}\] err_norm = {F8}"])
    for ln in lines:
        m = re.search(r"err_norm = (\S+)", ln)
        if m:
            assert float(m.group(1)) < 1e-3  # a stale ghost would give > 1e4
    lines = run_cli("mpi_stencil_gt", "--points", 1000, nproc=3, expect_rc=1)
    match_all(lines, [r"\d nmpi \(3\) must be divisor of domain size \(1000\), exiting"])


def test_mpi_stencil2d_gt_twelve_lines_world2():
    lines = run_cli("mpi_stencil2d_gt", 32, 3, "--n-other", 64, nproc=2)
    match_all(lines, [r"n procs        = 2", r"n_global_deriv = 64", r"n_global_other = 64", r"n_iter         = 3",
                      r"n_warmup       = 10"])
    tests = [ln for ln in lines if ln.startswith("TEST")]
    assert len(tests) == 12
    order = [(d, s, b) for d in (0, 1) for s in ("device ", "managed") for b in (1, 0)]
    for ln, (d, s, b) in zip(tests[:8], order):
        m = re.fullmatch(rf"TEST dim:{d}, {s}, buf:{b}; ({F8}), err=({F8})", ln)
        assert m, ln
        assert float(m.group(2)) < 1e-6
    for ln, (d, s) in zip(tests[8:], [(0, "device "), (0, "managed"), (1, "device "), (1, "managed")]):
        assert re.fullmatch(rf"TEST dim:{d}, {s}, buf:0; allreduce={F8}", ln), ln


def test_sycl_variants_world2():
    lines = run_cli("mpi_stencil2d_sycl", 16, 1, 3, "--ny", 64, nproc=2)
    match_all(lines, [r"0 \S+\([0-9a-f]+\)", r"n procs    = 2", r"nx_global  = 32", r"stage_host = 1", r"dev bytes  = \d+\.\d{3} (MB|GB)",
                      rf"0/2 exchange time {F8} ms", rf"1/2 \[1:0x[0-9a-f]{{8}}\] err_norm = {F8}"])
    lines = run_cli("mpi_stencil2d_sycl_oo", 1, 0, 2, nproc=2)
    match_all(lines, [r"n_global   = 1024", r"n_local    = 512", rf"0: exchange time {F8} ms", rf"1: \[0x[0-9a-f]{{8}}\] err_norm = {F8}"])


def test_sycl_oo_debug_dumps_world2():
    """--debug = the reference's DEBUG build: tiny domain, 1 iteration, boundary rows and received
    ghost rows dumped rank by rank on stderr (mpi_stencil2d_sycl_oo.cc:545-549,636-659,481-505)."""
    lines, err = run_cli("mpi_stencil2d_sycl_oo", 16, 0, 7, "--debug", nproc=2, with_stderr=True)
    match_all(lines, [r"n_global   = 16", r"n_local    = 8", r"n_iter     = 1", r"n_warmup   = 0"])
    rows = [ln for ln in err if re.match(r"\d: \[\d+, :\]", ln)]
    assert len(rows) == 8                                   # 2 + 2 boundary rows per rank
    assert [ln[0] for ln in rows] == ["0"] * 4 + ["1"] * 4   # rank 0's block first, then rank 1's
    assert all(len(ln.split()) == 3 + 16 for ln in rows)    # "r: [i, :]" + min(20, 16) values
    # rank 0 receives from its right neighbour only, rank 1 from its left neighbour only
    got = [ln for ln in err if "rbuf_" in ln]
    assert {ln.split("[")[0] for ln in got} == {"0: rbuf_r", "1: rbuf_l"}
    # what rank 1 received on its left is what rank 0 dumped as its last two interior rows
    sent = [float(v) for ln in rows[2:4] for v in ln.split()[3:]]              # rows 8, 9 of rank 0, 16 columns each
    recv = {}
    for ln in got:
        m = re.fullmatch(r"1: rbuf_l\[(\d+), (\d+)\] = (-?[0-9.]+)", ln)
        if m:
            recv[(int(m.group(1)), int(m.group(2)))] = float(m.group(3))
    assert len(recv) == 32
    for j in range(2):
        for i in range(16):
            assert recv[(j, i)] == pytest.approx(sent[j * 16 + i], abs=1e-6)


def test_jsonl_records_world2(tmp_path):
    """--jsonl PATH: one machine-readable record per rank next to the printed lines."""
    import json
    path = tmp_path / "rec.jsonl"
    run_cli("mpi_stencil_gt", "--points", 4096, "--jsonl", path, nproc=2)
    run_cli("mpi_stencil2d_gt", 16, 2, "--n-other", 32, "--jsonl", path, nproc=2)
    recs = [json.loads(ln) for ln in path.read_text().splitlines()]
    one_d = sorted((r for r in recs if r["program"] == "mpi_stencil_gt"), key=lambda r: r["rank"])
    assert [r["rank"] for r in one_d] == [0, 1] and all(r["world"] == 2 and r["n_local"] == 2048 for r in one_d)
    assert all(r["err_norm"] < 1e-3 and r["exchange_s"] >= 0 for r in one_d)
    two_d = [r for r in recs if r["program"] == "mpi_stencil2d_gt"]
    assert len(two_d) == 2
    for r in two_d:
        kinds = [(t["test"], t["dim"]) for t in r["tests"]]
        assert kinds.count(("deriv", 0)) >= 2 and kinds.count(("deriv", 1)) >= 2 and ("sum", 0) in kinds and ("sum", 1) in kinds
        assert all(t["err_norm"] < 1e-3 for t in r["tests"] if t["test"] == "deriv")


def test_rank_device_ratio_error_exit1():
    lines = run_cli("mpi_daxpy", nproc=3, env_extra={"GMT_FAKE_DEVICES": "2"}, expect_rc=1)
    match_all(lines, [r"ERROR: Number of ranks \(3\) not a multiple of number of GPUs \(2\)"])


def test_oversubscription_two_ranks_per_device():
    lines = run_cli("mpi_daxpy", nproc=4, env_extra={"GMT_FAKE_DEVICES": "2"})
    match_all(lines, [r"RANK\[1/4\] => DEVICE\[1/2\] mem=\d+", r"RANK\[2/4\] => DEVICE\[1/2\] mem=\d+",
                      r"RANK\[4/4\] => DEVICE\[2/2\] mem=\d+"])


def test_avg_and_run_command(tmp_path):
    from gpu_mpi_tests_b200.cli import avg, run
    f = tmp_path / "out-noum_none_1_2.txt"
    f.write_text("0/2 TIME gather : 0.500\n1/2 TIME gather : 0.250\n0/2 TIME kernel : 0.100\n")
    mean, n = avg.average(str(f), "gather")
    assert n == 2 and mean == pytest.approx(0.375)
    assert avg.average(str(f), "kernel")[0] == pytest.approx(0.1)
    cmd, outfile = run.build_command("um", "none", 1, 4, [])
    assert outfile == "out-um_none_1_4.txt" and "--managed" in cmd and "--nproc-per-node=4" in cmd
    cmd, _ = run.build_command("noum", "ncu", 1, 2, ["--barrier"])
    assert cmd[0] == "ncu" and "--profile-from-start" in cmd and "--managed" not in cmd and cmd[-1] == "--barrier"


def test_nvtx_ranges_of_daxpy_nvtx():
    from gpu_mpi_tests_b200.utils import nvtx
    nvtx.reset_trace()
    with nvtx.profiler_window():
        with nvtx.nvtx_range("mpiAllGather"):
            with nvtx.nvtx_range("x"):
                pass
            with nvtx.nvtx_range("y"):
                pass
    ev = nvtx.trace()
    assert ev[0][0] == "profiler_start" and ev[-1][0] == "profiler_stop"
    assert [(k, d, n) for (k, d, n) in ev if k == "push"] == [("push", 0, "mpiAllGather"), ("push", 1, "x"), ("push", 1, "y")]


@pytest.mark.gpu
def test_cli_gpu_fused_single_rank():
    """The GPU flavour of the programs on one B200 (fused kernels, world of one)."""
    lines = run_cli("mpi_stencil_gt", 4, cpu=False)
    match_all(lines, [r"n_global = 4194304", rf"0/1 exchange time {F8}", rf"0/1 \[0:0x[0-9a-f]{{8}}\] err_norm = {F8}"])
    lines = run_cli("mpi_stencil2d_gt", 64, 5, "--n-other", 4096, cpu=False)
    tests = [ln for ln in lines if ln.startswith("TEST")]
    assert len(tests) == 12
    for ln in tests[:8]:
        # one spacing (8/64) for both coordinates like the reference: the other axis reaches 512, and
        # err= is the rounding noise against y^2 ~ 2.6e5 (tests/test_cpu_units.py pins it to the oracle)
        assert float(re.search(r"err=(\S+)", ln).group(1)) < 1e-4
    lines = run_cli("mpi_daxpy_nvtx", "--n-per-node", 1 << 20, "--managed", cpu=False)
    match_all(lines, [r"0/1 SUM = 524288\.500000", r"0/1 ALLSUM = 524288\.500000",
                      r"CUDA PreferredLocation of 'd_x' is (CPU|DEVICE|INVALID) \(-?\d+\)"])
    lines = run_cli("mpi_daxpy", cpu=False)
    match_all(lines, [r"CUDA PreferredLocation of 'd_x' is UNMANAGED", r"CUDA PreferredLocation of 'x' is NOT CUDA",
                      r"CUDA PreferredLocation of 'm_x' is .+", r"0/1 SUM = 524800\.000000"])
    lines = run_cli("mpigatherinplace", 1 << 20, cpu=False)
    assert len(lines) == 1
