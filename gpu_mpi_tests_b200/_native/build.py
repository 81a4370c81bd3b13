"""Build libgmt_b200.so in-tree with nvcc for sm_100a.

The reference builds with a bare Makefile without any -arch flag (Makefile:4-20) and a CMake
project that fetches gtensor (CMakeLists.txt:9-13).  Here there is one shared library with a C
ABI (include/gmt_b200.h); nvcc cross-compiles it without a GPU, one translation unit per
process in parallel, relinking only what changed.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
CSRC = ROOT / "csrc"
OUT_DIR = Path(__file__).resolve().parent
LIB = OUT_DIR / "libgmt_b200.so"
OBJ_DIR = OUT_DIR / "obj"

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
    "--expt-relaxed-constexpr",
]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _headers_mtime() -> float:
    hs = list(CSRC.glob("*.h")) + list(CSRC.glob("*.cuh")) + list((ROOT / "include").glob("*.h"))
    return max(h.stat().st_mtime for h in hs)


def _compile(nvcc: str, src: Path, obj: Path, verbose: bool) -> str:
    extra = os.environ.get("GMT_NVCC_EXTRA", "").split()
    cmd = [nvcc, *NVCC_FLAGS, *extra, "-I", str(ROOT / "include"), "-c", str(src), "-o", str(obj)]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src.name}:\n{res.stdout}\n{res.stderr}")
    (obj.with_suffix(".ptxas.txt")).write_text(res.stderr)
    if verbose:
        sys.stderr.write(f"[build] {src.name} ok\n")
    return res.stderr


def build(force: bool = False, verbose: bool = False) -> Path:
    """Compile every csrc/*.cu for sm_100a and link libgmt_b200.so; returns its path."""
    nvcc = find_nvcc()
    OBJ_DIR.mkdir(parents=True, exist_ok=True)
    hdr_m = _headers_mtime()
    jobs = []
    objs = []
    for src in sources():
        obj = OBJ_DIR / (src.stem + ".o")
        objs.append(obj)
        stale = force or not obj.exists() or obj.stat().st_mtime < max(src.stat().st_mtime, hdr_m)
        if stale:
            jobs.append((src, obj))
    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            list(ex.map(lambda so: _compile(nvcc, so[0], so[1], verbose), jobs))
    if jobs or not LIB.exists():
        cmd = [nvcc, "-shared", "-o", str(LIB), *map(str, objs), "-lcudart", "-ldl"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
        if verbose:
            sys.stderr.write(f"[build] linked {LIB}\n")
    build_programs(nvcc, verbose)
    return LIB


NATIVE_SRC = ROOT / "native"
BIN_DIR = OUT_DIR / "bin"


def build_programs(nvcc: str | None = None, verbose: bool = False) -> list[Path]:
    """Native programs (native/*.cpp): plain C++ against the C ABI, linked to the in-tree library
    through an $ORIGIN-relative rpath so the binaries work wherever the tree is copied."""
    nvcc = nvcc or find_nvcc()
    BIN_DIR.mkdir(parents=True, exist_ok=True)
    built = []
    for src in sorted(NATIVE_SRC.glob("*.cpp")):
        exe = BIN_DIR / src.stem
        newest = max(src.stat().st_mtime, LIB.stat().st_mtime,
                     *(h.stat().st_mtime for h in (ROOT / "include").glob("*.h*")))
        if not exe.exists() or exe.stat().st_mtime < newest:
            cmd = [nvcc, "-O2", "-std=c++17", "-I", str(ROOT / "include"), str(src), "-o", str(exe),
                   "-L", str(OUT_DIR), "-lgmt_b200", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/..", "-lcudart"]
            res = subprocess.run(cmd, capture_output=True, text=True)
            if res.returncode != 0:
                raise RuntimeError(f"building {src.name} failed:\n{res.stdout}\n{res.stderr}")
            if verbose:
                sys.stderr.write(f"[build] {exe}\n")
        built.append(exe)
    return built


def ensure_built() -> Path:
    """Build the library and the programs if the tree has none (a fresh checkout: build outputs are
    git-ignored).  Safe to call from every rank of a multi-process launch: an exclusive file lock
    lets one process build while the others wait and then find the result."""
    import fcntl
    if LIB.exists():
        return LIB
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    with open(OUT_DIR / ".build.lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if not LIB.exists():
                build()
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print(LIB)
