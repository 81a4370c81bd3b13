"""Multi-process tests: CPU/gloo (world 2 and 4, run anywhere) and multi-GPU (NCCL control
plane + peer-memory data plane, needs >= 2 GPUs)."""
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]
WORKER = str(ROOT / "tests" / "mp_worker.py")
_port = [29600 + (os.getpid() % 200)]


def run_workers(nproc: int, case: str, *args, env_extra=None, timeout=600):
    _port[0] += 1
    env = dict(os.environ)
    env.update(env_extra or {})
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_port[0]), WORKER, case, *map(str, args)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env, cwd=str(ROOT))
    assert res.returncode == 0, f"{case} x{nproc} failed:\n{res.stdout[-3000:]}\n{res.stderr[-3000:]}"
    return res.stdout


@pytest.mark.parametrize("nproc", [2, 4])
def test_baseline_exchange_gloo(nproc):
    out = run_workers(nproc, "baseline", env_extra={"GMT_FORCE_CPU": "1"})
    assert "baseline ok" in out


@pytest.mark.parametrize("nproc", [2, 4])
def test_user_compute_step_on_dist_arrays_gloo(nproc):
    """DistArray.exchange()/apply(fn) with arbitrary data and a user's own stencil (CPU/gloo flavour)."""
    out = run_workers(nproc, "darray", env_extra={"GMT_FORCE_CPU": "1"})
    assert "darray ok" in out


@pytest.mark.gpu
def test_user_compute_step_on_dist_arrays_across_gpus():
    """The same through the peer-memory exchange kernel on real GPUs."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    out = run_workers(n if n in (2, 4, 8) else 2, "darray")
    assert "darray ok" in out


@pytest.mark.gpu
def test_fused_kernels_across_gpus():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    out = run_workers(min(n, 8) if n in (2, 4, 8) else 2, "fused")
    assert "fused ok" in out


@pytest.mark.gpu
def test_baseline_exchange_nccl():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    out = run_workers(2, "baseline")
    assert "baseline ok" in out


@pytest.mark.gpu
def test_collectives_across_gpus():
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    out = run_workers(n if n in (2, 4, 8) else 2, "collectives")
    assert "collectives ok" in out
