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
    if LIB.exists() and not force and MANIFEST.exists() and is_stale():
        # contents changed: trust hashes over mtimes for the objects whose source differs
        import json
        try:
            old = json.loads(MANIFEST.read_text())
        except ValueError:
            old = {}
        new = _fingerprint()
        changed = {k for k in new if old.get(k) != new[k]}
        if any(not k.endswith(".cu") for k in changed):
            force = True           # a header or the flags changed: everything is stale
        else:
            for k in changed:
                (OBJ_DIR / (Path(k).stem + ".o")).unlink(missing_ok=True)
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
        # link under a temporary name and rename: a rank that looks for the library while the link
        # is running must see the old file or the new one, never a partial one
        tmp = LIB.with_name(f".{LIB.name}.{os.getpid()}.tmp")
        cmd = [nvcc, "-shared", "-o", str(tmp), *map(str, objs), "-lcudart", "-ldl"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            tmp.unlink(missing_ok=True)
            raise RuntimeError(f"link failed:\n{res.stdout}\n{res.stderr}")
        os.replace(tmp, LIB)
        if verbose:
            sys.stderr.write(f"[build] linked {LIB}\n")
    import json
    MANIFEST.write_text(json.dumps(_fingerprint(), indent=0))
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


MANIFEST = OUT_DIR / "build_manifest.json"


def _fingerprint() -> dict:
    """Content hashes of everything the library is built from (mtimes do not survive every copy of
    the tree, contents do)."""
    import hashlib
    files = sources() + sorted(CSRC.glob("*.h")) + sorted(CSRC.glob("*.cuh")) + sorted((ROOT / "include").glob("*.h"))
    out = {str(f.relative_to(ROOT)): hashlib.sha256(f.read_bytes()).hexdigest() for f in files}
    out["flags"] = " ".join(NVCC_FLAGS + os.environ.get("GMT_NVCC_EXTRA", "").split())
    return out


def is_stale() -> bool:
    """True when the library is missing or was built from other sources than the tree holds now."""
    import json
    if not LIB.exists():
        return True
    try:
        return json.loads(MANIFEST.read_text()) != _fingerprint()
    except (OSError, ValueError):
        return True


def ensure_built() -> Path:
    """Bring the library and the programs up to date with the sources (a fresh checkout has none:
    build outputs are git-ignored; after an edit of csrc/ or include/ the stale objects are
    recompiled).  Safe to call from every rank of a multi-process launch: the exclusive file lock is
    taken FIRST, so one process builds while the others wait and then find nothing left to do.
    Without a compiler an existing library is used as it is."""
    import fcntl
    OUT_DIR.mkdir(parents=True, exist_ok=True)
    with open(OUT_DIR / ".build.lock", "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            if is_stale():
                try:
                    build()
                except RuntimeError:
                    if not LIB.exists():
                        raise
                    sys.stderr.write("[build] WARNING: sources are newer than libgmt_b200.so and rebuilding failed; "
                                     "using the existing library\n")
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print(LIB)
