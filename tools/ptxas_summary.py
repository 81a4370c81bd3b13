#!/usr/bin/env python
"""Summarise `nvcc -Xptxas -v` output saved by the build: registers, stack, spills per kernel."""
import glob
import re
import subprocess
from pathlib import Path

obj = Path(__file__).resolve().parents[1] / "gpu_mpi_tests_b200" / "_native" / "obj"
pat = re.compile(
    r"Compiling entry function '(\S+)' for 'sm_100a'\n.*?Function properties for \S+\n\s+(\d+) bytes stack frame, "
    r"(\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(.*)")
for f in sorted(glob.glob(str(obj / "*.ptxas.txt"))):
    txt = open(f).read()
    for m in pat.finditer(txt):
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        print(f"{Path(f).name[:-10]:12s} regs={m.group(5):>3s} stack={m.group(2)} spill={m.group(3)}/{m.group(4)} "
              f"{name[:100]} |{m.group(6)[:70]}")
