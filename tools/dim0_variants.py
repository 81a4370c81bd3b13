#!/usr/bin/env python
"""Tuning sweep of the fused dim-0 slab step (reference-default tile 1024 x 524288 per rank) under torchrun:
landing vs ghosts-from-inbox, chunks per fence, number of exporting CTAs.  Device-timed, max over ranks."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from gpu_mpi_tests_b200.models.stencil2d import Stencil2DWorkload
from gpu_mpi_tests_b200.parallel import runtime
from gpu_mpi_tests_b200.parallel.symm import make_space
from gpu_mpi_tests_b200.utils import timing

rt = runtime.init()
space = make_space(rt)
nx, ny = 1024, 524288
steps = int(os.environ.get("VARIANT_STEPS", "20"))
rows = []
for land in (True, False):
    wl = Stencil2DWorkload(rt, nx, ny, mode="dim0", impl="fused", with_err=False, space=space, land_ghosts=land)
    wl.setup()
    for group, ctas in ((0, 0), (1, 32), (4, 32), (8, 32), (4, 16), (4, 64), (4, 128), (2, 64), (1, 128), (15, 32)):
        os.environ.pop("GMT_PUSH_GROUP", None)
        os.environ.pop("GMT_PUSH_CTAS", None)
        if group:
            os.environ["GMT_PUSH_GROUP"] = str(group)
        if ctas:
            os.environ["GMT_PUSH_CTAS"] = str(ctas)
        ms = min(timing.time_steps(wl.step, steps, 3, rt.device)["ms_per_step"] for _ in range(2))
        wl.check()
        rows.append({"land_ghosts": land, "push_group": group or "default", "push_ctas": ctas or "default",
                     "ms_per_step": round(ms, 4), "n_gpus": rt.world_size})
        if rt.rank == 0:
            print(json.dumps(rows[-1]), flush=True)
    os.environ.pop("GMT_PUSH_GROUP", None)
    os.environ.pop("GMT_PUSH_CTAS", None)
    plain = min(timing.time_steps(wl.stencil_only, steps, 3, rt.device)["ms_per_step"] for _ in range(2))
    if rt.rank == 0:
        print(json.dumps({"plain_sweep_ms": round(plain, 4), "n_gpus": rt.world_size}), flush=True)
    del wl
    torch.cuda.empty_cache()
runtime.finalize()
