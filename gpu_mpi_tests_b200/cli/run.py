"""Launcher: the run.sh of the reference (summit/run.sh:1-32, jlse/run.sh:1-34).

    python -m gpu_mpi_tests_b200.cli.run um|noum ncu|nsys|none nodes ppn [-- extra program args]

Picks the managed / unmanaged flavour of mpi_daxpy_nvtx, optionally wraps every rank in a
profiler that honours the cudaProfilerStart/Stop window (`ncu --profile-from-start off`, or
nsys `-c cudaProfilerApi` where nsys exists), launches nodes*ppn ranks with torchrun on this
node and redirects stdout to out-<um>_<prof>_<nodes>_<ppn>.txt (feed those to cli.avg)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys


def build_command(um: str, prof: str, nodes: int, ppn: int, extra: list, port: int = 29533) -> tuple:
    if um not in ("um", "noum"):
        raise SystemExit("Usage: run um|noum nsys|ncu|none nodes ppn")
    tag = f"{um}_{prof}_{nodes}_{ppn}"
    prog = ["-m", "gpu_mpi_tests_b200.cli.mpi_daxpy_nvtx"] + (["--managed"] if um == "um" else []) + list(extra)
    nproc = nodes * ppn
    launcher = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
                "--master-addr", "127.0.0.1", "--master-port", str(port)]
    if prof == "none":
        cmd = launcher + prog
    elif prof == "ncu":
        os.makedirs("profile", exist_ok=True)
        cmd = ["ncu", "--target-processes", "all", "--profile-from-start", "off", "--clock-control", "none",
               "-o", f"profile/{tag}"] + launcher + prog
    elif prof == "nsys":
        if not shutil.which("nsys"):
            raise SystemExit("nsys is not installed on this image; use ncu or none")
        os.makedirs("profile", exist_ok=True)
        cmd = ["nsys", "profile", "--kill=none", "-c", "cudaProfilerApi", "-o", f"profile/{tag}.%q{{RANK}}"] + launcher + prog
    else:
        raise SystemExit("profiler must be one of nsys, ncu, none")
    return cmd, f"out-{tag}.txt"


def main(argv=None) -> int:
    argv = list(sys.argv[1:] if argv is None else argv)
    extra = []
    if "--" in argv:
        i = argv.index("--")
        argv, extra = argv[:i], argv[i + 1:]
    if len(argv) != 4:
        print("Usage: run um|noum nsys|ncu|none nodes ppn [-- program args]")
        return 1
    um, prof, nodes, ppn = argv[0], argv[1], int(argv[2]), int(argv[3])
    cmd, outfile = build_command(um, prof, nodes, ppn, extra)
    with open(outfile, "w") as f:
        rc = subprocess.run(cmd, stdout=f, stderr=subprocess.STDOUT).returncode
    print(f"{' '.join(cmd)}\n-> {outfile} (exit {rc})")
    return rc


if __name__ == "__main__":
    sys.exit(main())
