#!/usr/bin/env python
"""Tile-width sweep of the 2-D kernels ($GMT_TILE_JX / $GMT_TILE_JY), single GPU."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from gpu_mpi_tests_b200.ops import local
from gpu_mpi_tests_b200.parallel.halo import FusedStencil2D
from gpu_mpi_tests_b200.parallel.symm import VirtualSpace
from gpu_mpi_tests_b200.utils import timing

dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
space = VirtualSpace(1, dev)
res = {}


def t(fn):
    return timing.time_steps(fn, 20, 5, dev)["ms_per_step"]


for label, (nx, ny, gx, gy) in {"grid": (16384, 16384, 2, 2), "dim1": (524288, 1024, 0, 2), "dim0": (1024, 524288, 2, 0)}.items():
    p = FusedStencil2D(space, 1, 1, nx, ny, gx, gy, with_err=(label == "grid"))
    p.init_analytic()
    z, o = p.z[0], p.out[0]
    key = "GMT_TILE_JY" if gy else "GMT_TILE_JX"
    for w in (8, 16, 32, 64, 128, 256):
        os.environ[key] = str(w)
        fused = t(p.step)
        if gx and gy:
            plain = t(lambda: local.stencil2d_xy(z, p.sx, p.sy, out=o))
        elif gx:
            plain = t(lambda: local.stencil2d(z, 0, p.sx, out=o))
        else:
            plain = t(lambda: local.stencil2d(z, 1, p.sy, out=o))
        res[f"{label}_w{w}"] = {"fused_ms": fused, "plain_ms": plain, "fused_glups": nx * ny / fused / 1e6,
                                "plain_glups": nx * ny / plain / 1e6}
        print(label, w, res[f"{label}_w{w}"], flush=True)
    os.environ.pop(key, None)
    del p
    torch.cuda.empty_cache()
Path("gpurun_out").mkdir(exist_ok=True)
Path("gpurun_out/tile_sweep.json").write_text(json.dumps(res, indent=1))
