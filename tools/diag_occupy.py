"""Diagnostic: does a kernel on another stream run while gmt_debug_occupy holds most of the GPU?"""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from gpu_mpi_tests_b200.ops._lib import lib
from gpu_mpi_tests_b200.parallel.collectives import PeerCollectives
from gpu_mpi_tests_b200.parallel.symm import VirtualSpace

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
sms = torch.cuda.get_device_properties(dev).multi_processor_count
F64 = torch.float64
n_out = 4096
pc = PeerCollectives(VirtualSpace(1, dev), stage_cap=n_out, spin_limit=1 << 26)
z = {0: torch.rand(n_out, 64, dtype=F64, device=dev)}
out = {0: torch.zeros(n_out, dtype=F64, device=dev)}
pc.axis_sum_allreduce(z, 0, out, n_out)
x = torch.zeros(1 << 20, device=dev)
torch.cuda.synchronize()

def trial(name, ctas, fn):
    hog, work = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    hog_done = torch.cuda.Event()
    t0 = time.perf_counter()
    with torch.cuda.stream(hog):
        lib.gmt_debug_occupy(ctas, 200_000_000, hog.cuda_stream)
    time.sleep(0.02)
    t1 = time.perf_counter()
    with torch.cuda.stream(work):
        fn()
    hog_done.record(hog)      # queued only after the work kernel: no event sits between the two kernels
    work.synchronize()
    t2 = time.perf_counter()
    early = not hog_done.query()
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    print(f"{name:40s} hog_ctas={ctas:4d} work took {1e3*(t2-t1):8.3f} ms, early={early}, hog total {1e3*(t3-t0):.1f} ms", flush=True)

import os
print("CUDA_DEVICE_MAX_CONNECTIONS =", os.environ.get("CUDA_DEVICE_MAX_CONNECTIONS"), flush=True)
for ctas in (sms // 2, sms, 2 * (sms - 4), 2 * sms - 16):
    trial("torch add_", ctas, lambda: x.add_(1.0))
    trial("axis_sum_allreduce axis0", ctas, lambda: pc.axis_sum_allreduce(z, 0, out, n_out))
    trial("barrier", ctas, lambda: pc.barrier())
pc.check_status()
