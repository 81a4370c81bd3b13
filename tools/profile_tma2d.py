#!/usr/bin/env python
"""ncu target: the dim-1 TMA/cluster stencil variant and the LSU variant at the reference tile."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch
from gpu_mpi_tests_b200.ops import local
dev = torch.device("cuda", 0)
nx, ny = 524288, 1024
z = torch.rand(ny + 4, nx, dtype=torch.float64, device=dev)
o = torch.empty(ny, nx, dtype=torch.float64, device=dev)
for v in ("tma", "vector"):
    for _ in range(3):
        local.stencil2d(z, 1, 1.0, out=o, variant=v)
torch.cuda.synchronize()
print("done")
