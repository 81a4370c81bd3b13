#!/usr/bin/env python
"""First-contact GPU probe: environment facts, kernel numerics against torch fp64, and
CUDA-event timings of every single-GPU kernel at the BASELINE shapes.  Writes
gpurun_out/probe.json.  (Development tool; the product benchmarks live in bench.py and
benchmarks/.)"""
import json
import shutil
import subprocess
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import torch

from gpu_mpi_tests_b200.ops._lib import lib, Halo1dRank, Halo2dRank

OUT = Path("gpurun_out")
OUT.mkdir(exist_ok=True)
res = {"env": {}, "numerics": {}, "timing": {}}


def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=60).stdout.strip()
    except Exception as e:  # noqa
        return f"ERR {e}"


res["env"]["mpi"] = {k: shutil.which(k) for k in ("mpicxx", "mpirun", "mpifort", "mpicc")}
res["env"]["nvidia_smi"] = sh("nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm,memory.total --format=csv")
res["env"]["topo"] = sh("nvidia-smi topo -m | head -20")
res["env"]["ndev"] = torch.cuda.device_count()

dev = torch.device("cuda:0")
torch.cuda.set_device(0)
st = lambda: torch.cuda.current_stream().cuda_stream
f64 = torch.float64
scratch = torch.zeros(lib.gmt_reduce_scratch_bytes(), dtype=torch.uint8, device=dev)
C0, C1, C3, C4 = 1 / 12, -2 / 3, 2 / 3, -1 / 12


def ref1d(y, scale):
    return (C0 * y[:-4] + C1 * y[1:-3] + C3 * y[3:-1] + C4 * y[4:]) * scale


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        b.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return {"min_ms": ts[0], "med_ms": ts[len(ts) // 2]}


def relerr(a, b):
    return float((a - b).abs().max() / (b.abs().max() + 1e-300))


# ---------------------------------------------------------------- numerics (small)
g = torch.Generator(device="cuda").manual_seed(0)
for n in (1024, 1000, 4096 * 3 + 8, 1 << 20, (1 << 20) + 6):
    x = torch.rand(n, dtype=f64, device=dev, generator=g)
    y = torch.rand(n, dtype=f64, device=dev, generator=g)
    yr = 2.0 * x + y
    y1 = y.clone()
    lib.gmt_daxpy_f64(n, 2.0, x.data_ptr(), y1.data_ptr(), st())
    res["numerics"][f"daxpy_{n}"] = relerr(y1, yr)
    y2 = y.clone()
    out = torch.zeros(1, dtype=f64, device=dev)
    lib.gmt_daxpy_sum_f64(n, 2.0, x.data_ptr(), y2.data_ptr(), out.data_ptr(), scratch.data_ptr(), st())
    res["numerics"][f"daxpy_sum_{n}"] = [relerr(y2, yr), abs(out.item() - yr.sum().item()) / abs(yr.sum().item())]
    lib.gmt_sum_f64(n, x.data_ptr(), out.data_ptr(), scratch.data_ptr(), st())
    res["numerics"][f"sum_{n}"] = abs(out.item() - x.sum().item()) / x.sum().item()
    lib.gmt_sumsq_diff_f64(n, x.data_ptr(), y.data_ptr(), out.data_ptr(), scratch.data_ptr(), st())
    r = ((x - y) ** 2).sum().item()
    res["numerics"][f"sumsq_{n}"] = abs(out.item() - r) / r
    yin = torch.rand(n + 4, dtype=f64, device=dev, generator=g)
    rr = ref1d(yin, 3.5)
    for v in (0, 1, 2):
        o = torch.zeros(n, dtype=f64, device=dev)
        lib.gmt_stencil1d_f64(n, yin.data_ptr(), o.data_ptr(), 3.5, v, st())
        res["numerics"][f"sten1d_v{v}_{n}"] = relerr(o, rr)

for (nx, ny) in ((8, 8), (1024, 64), (2048, 130), (12, 7), (4096, 257)):
    # dim 0: z (nx+4, ny) column-major == torch (ny, nx+4)
    z = torch.rand(ny, nx + 4, dtype=f64, device=dev, generator=g)
    rr = (C0 * z[:, :-4] + C1 * z[:, 1:-3] + C3 * z[:, 3:-1] + C4 * z[:, 4:]) * 2.5
    for v in (0, 1):
        o = torch.zeros(ny, nx, dtype=f64, device=dev)
        lib.gmt_stencil2d_f64(0, nx, ny, z.data_ptr(), nx + 4, o.data_ptr(), nx, 2.5, v, st())
        res["numerics"][f"sten2d_d0_v{v}_{nx}x{ny}"] = relerr(o, rr)
    z = torch.rand(ny + 4, nx, dtype=f64, device=dev, generator=g)
    rr = (C0 * z[:-4] + C1 * z[1:-3] + C3 * z[3:-1] + C4 * z[4:]) * 2.5
    for v in (0, 1):
        o = torch.zeros(ny, nx, dtype=f64, device=dev)
        lib.gmt_stencil2d_f64(1, nx, ny, z.data_ptr(), nx, o.data_ptr(), nx, 2.5, v, st())
        res["numerics"][f"sten2d_d1_v{v}_{nx}x{ny}"] = relerr(o, rr)
    z = torch.rand(ny + 4, nx + 4, dtype=f64, device=dev, generator=g)
    zc = z[2:-2]
    rx = (C0 * zc[:, :-4] + C1 * zc[:, 1:-3] + C3 * zc[:, 3:-1] + C4 * zc[:, 4:]) * 2.5
    zr = z[:, 2:-2]
    ry = (C0 * zr[:-4] + C1 * zr[1:-3] + C3 * zr[3:-1] + C4 * zr[4:]) * 1.5
    o = torch.zeros(ny, nx, dtype=f64, device=dev)
    lib.gmt_stencil2d_xy_f64(nx, ny, z.data_ptr(), nx + 4, o.data_ptr(), nx, 2.5, 1.5, st())
    res["numerics"][f"sten2d_xy_{nx}x{ny}"] = relerr(o, rx + ry)
torch.cuda.synchronize()
print(json.dumps(res["numerics"], indent=1))
bad = {k: v for k, v in res["numerics"].items() if (max(v) if isinstance(v, list) else v) > 1e-12}
res["numerics_bad"] = bad
print("BAD:", bad)


# ---------------------------------------------------------------- fused, virtual ranks on one GPU
def dev_bytes(structs):
    raw = b"".join(bytes(s) for s in structs)
    t = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
    return t


def fused1d_emulated(R, n, iters=3, with_err=True):
    nglob = R * n
    dx = 8.0 / nglob
    scale = nglob / 8.0
    keep = []
    zs, outs = [], []
    inbox = torch.zeros(R, 2, 2, 4, dtype=f64, device=dev)  # [rank][side][parity][2 doubles, padded to 4]
    flags = torch.zeros(R, 2, 4, dtype=torch.int64, device=dev)
    epoch = torch.zeros(R, 8, dtype=torch.int64, device=dev)
    done = torch.zeros(R, 8, dtype=torch.int32, device=dev)
    status = torch.zeros(R, 8, dtype=torch.int32, device=dev)
    errs = torch.zeros(R, dtype=f64, device=dev)
    scr = torch.zeros(R, lib.gmt_reduce_scratch_bytes(), dtype=torch.uint8, device=dev)
    structs = []
    for r in range(R):
        z = torch.full((n + 4,), 1e30, dtype=f64, device=dev)
        lib.gmt_init_stencil1d_f64(n, 2, r * n * dx, dx, z.data_ptr(), 0, int(r == 0), int(r == R - 1), st())
        o = torch.zeros(n, dtype=f64, device=dev)
        zs.append(z)
        outs.append(o)
    for r in range(R):
        s = Halo1dRank()
        s.z, s.out, s.n, s.scale = zs[r].data_ptr(), outs[r].data_ptr(), n, scale
        s.has[0], s.has[1] = int(r > 0), int(r < R - 1)
        for side in range(2):
            nb = r - 1 if side == 0 else r + 1
            for p in range(2):
                s.my_inbox[side][p] = inbox[r, side, p].data_ptr()
                s.peer_inbox[side][p] = inbox[nb, 1 - side, p].data_ptr() if 0 <= nb < R else 0
            s.my_flag[side] = flags[r, side].data_ptr()
            s.peer_flag[side] = flags[nb, 1 - side].data_ptr() if 0 <= nb < R else 0
        s.epoch, s.done, s.status = epoch[r].data_ptr(), done[r].data_ptr(), status[r].data_ptr()
        s.err_scratch, s.err_sumsq = scr[r].data_ptr(), errs[r].data_ptr()
        s.x0, s.dx, s.spin_limit = r * n * dx, dx, 1 << 22
        structs.append(s)
    rd = dev_bytes(structs)
    for _ in range(iters):
        lib.gmt_fused_stencil1d_f64(rd.data_ptr(), R, n, int(with_err), st())
    torch.cuda.synchronize()
    # oracle
    xs = (torch.arange(nglob, dtype=f64, device=dev)) * dx
    worst = 0.0
    for r in range(R):
        act = 3 * xs[r * n:(r + 1) * n] ** 2
        worst = max(worst, float((outs[r] - act).abs().max()))
    return {"max_abs_err": worst, "status": status[:, 0].tolist(), "epoch": epoch[:, 0].tolist(),
            "err_norm_dev": [float(e) ** 0.5 for e in errs.tolist()]}


def fused2d_emulated(PX, PY, nx, ny, gx, gy, iters=3, with_err=True, time_it=False):
    """PX x PY virtual ranks; rank = px + py*PX; x split over px, y over py."""
    R = PX * PY
    NX, NY = nx * PX, ny * PY
    dx, dy = 8.0 / NX, 8.0 / NY
    sx, sy = (NX / 8.0 if gx else 0.0), (NY / 8.0 if gy else 0.0)
    ldz = nx + 2 * gx
    zs, outs = [], []
    nfx = int(lib.gmt_halo2d_flag_count(0, nx, ny))
    nfy = int(lib.gmt_halo2d_flag_count(2, nx, ny))
    nf = max(nfx, nfy)
    inbox_x = torch.zeros(R, 2, 2, 2 * ny, dtype=f64, device=dev)
    inbox_y = torch.zeros(R, 2, 2, 2 * nx, dtype=f64, device=dev)
    flags = torch.zeros(R, 4, nf, dtype=torch.int64, device=dev)
    epoch = torch.zeros(R, 8, dtype=torch.int64, device=dev)
    done = torch.zeros(R, 8, dtype=torch.int32, device=dev)
    status = torch.zeros(R, 8, dtype=torch.int32, device=dev)
    errs = torch.zeros(R, dtype=f64, device=dev)
    scr = torch.zeros(R, lib.gmt_reduce_scratch_bytes(), dtype=torch.uint8, device=dev)
    structs = []
    for r in range(R):
        px, py = r % PX, r // PX
        z = torch.empty(ny + 2 * gy, ldz, dtype=f64, device=dev)
        mask = (1 if px == 0 else 0) | (2 if px == PX - 1 else 0) | (4 if py == 0 else 0) | (8 if py == PY - 1 else 0)
        lib.gmt_init_stencil2d_f64(nx, ny, gx, gy, px * nx * dx, py * ny * dy, dx, dy, z.data_ptr(), ldz, 0, nx, 0,
                                   sx, sy, mask, 1e30, st())
        zs.append(z)
        outs.append(torch.zeros(ny, nx, dtype=f64, device=dev))
    nbr = lambda px, py: (px + py * PX) if (0 <= px < PX and 0 <= py < PY) else -1
    for r in range(R):
        px, py = r % PX, r // PX
        s = Halo2dRank()
        s.z, s.out = zs[r].data_ptr(), outs[r].data_ptr()
        s.nx, s.ny, s.ldz, s.ldo, s.gx, s.gy, s.sx, s.sy = nx, ny, ldz, nx, gx, gy, sx, sy
        nbs = [nbr(px - 1, py) if gx else -1, nbr(px + 1, py) if gx else -1,
               nbr(px, py - 1) if gy else -1, nbr(px, py + 1) if gy else -1]
        for side in range(4):
            nb = nbs[side]
            s.has[side] = int(nb >= 0)
            opp = side ^ 1
            for p in range(2):
                box = inbox_x if side < 2 else inbox_y
                s.my_inbox[side][p] = box[r, side & 1, p].data_ptr()
                s.peer_inbox[side][p] = box[nb, opp & 1, p].data_ptr() if nb >= 0 else 0
            s.my_flag[side] = flags[r, side].data_ptr()
            s.peer_flag[side] = flags[nb, opp].data_ptr() if nb >= 0 else 0
        s.epoch, s.done, s.status = epoch[r].data_ptr(), done[r].data_ptr(), status[r].data_ptr()
        s.err_scratch, s.err_sumsq = scr[r].data_ptr(), errs[r].data_ptr()
        s.x0, s.y0, s.dx, s.dy, s.spin_limit = px * nx * dx, py * ny * dy, dx, dy, 1 << 22
        structs.append(s)
    rd = dev_bytes(structs)
    fn = lambda: lib.gmt_fused_stencil2d_f64(rd.data_ptr(), R, nx, ny, gx, gy, int(with_err), st())
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    worst = 0.0
    for r in range(R):
        px, py = r % PX, r // PX
        xs = (px * nx + torch.arange(nx, dtype=f64, device=dev)) * dx
        ys = (py * ny + torch.arange(ny, dtype=f64, device=dev)) * dy
        act = (3 * xs[None, :] ** 2 if gx else 0.0) + (2 * ys[:, None] if gy else 0.0) + torch.zeros(ny, nx, dtype=f64, device=dev)
        worst = max(worst, float((outs[r] - act).abs().max()))
    out = {"max_abs_err": worst, "status": status[:, 0].tolist(), "epoch": epoch[:, 0].tolist(),
           "err_norm_dev": [float(e) ** 0.5 for e in errs.tolist()]}
    if time_it:
        out["time"] = timeit(fn)
    return out


res["fused"] = {}
try:
    res["fused"]["1d_R1_n4096"] = fused1d_emulated(1, 4096)
    res["fused"]["1d_R4_n8192"] = fused1d_emulated(4, 8192)
    res["fused"]["1d_R3_n1000"] = fused1d_emulated(3, 1000)
    res["fused"]["1d_R8_n65536"] = fused1d_emulated(8, 65536, iters=7)
    res["fused"]["2d_d0_R4"] = fused2d_emulated(4, 1, 1024, 256, 2, 0)
    res["fused"]["2d_d0_R2_wide"] = fused2d_emulated(2, 1, 4096, 100, 2, 0, iters=4)
    res["fused"]["2d_d1_R4"] = fused2d_emulated(1, 4, 2048, 256, 0, 2)
    res["fused"]["2d_d1_R3_odd"] = fused2d_emulated(1, 3, 1024, 130, 0, 2, iters=5)
    res["fused"]["2d_grid_2x2"] = fused2d_emulated(2, 2, 2048, 256, 2, 2)
    res["fused"]["2d_grid_4x2"] = fused2d_emulated(4, 2, 1024, 384, 2, 2, iters=6)
    res["fused"]["2d_grid_1x1"] = fused2d_emulated(1, 1, 2048, 512, 2, 2)
except Exception as e:  # noqa
    res["fused"]["exception"] = repr(e)
print(json.dumps(res["fused"], indent=1))

# ---------------------------------------------------------------- timings at BASELINE shapes
if "--no-timing" not in sys.argv:
    T = res["timing"]
    n = 48 << 20
    x = torch.rand(n, dtype=f64, device=dev)
    y = torch.rand(n, dtype=f64, device=dev)
    out = torch.zeros(1, dtype=f64, device=dev)
    T["daxpy_48Mi"] = timeit(lambda: lib.gmt_daxpy_f64(n, 2.0, x.data_ptr(), y.data_ptr(), st()))
    T["daxpy_48Mi"]["GBs"] = 24 * n / T["daxpy_48Mi"]["min_ms"] / 1e6
    T["daxpy_sum_48Mi"] = timeit(lambda: lib.gmt_daxpy_sum_f64(n, 2.0, x.data_ptr(), y.data_ptr(), out.data_ptr(),
                                                               scratch.data_ptr(), st()))
    T["daxpy_sum_48Mi"]["GBs"] = 24 * n / T["daxpy_sum_48Mi"]["min_ms"] / 1e6
    T["torch_axpy_48Mi"] = timeit(lambda: y.add_(x, alpha=2.0))
    T["torch_axpy_48Mi"]["GBs"] = 24 * n / T["torch_axpy_48Mi"]["min_ms"] / 1e6
    T["sum_48Mi"] = timeit(lambda: lib.gmt_sum_f64(n, x.data_ptr(), out.data_ptr(), scratch.data_ptr(), st()))
    T["sum_48Mi"]["GBs"] = 8 * n / T["sum_48Mi"]["min_ms"] / 1e6
    T["copy_48Mi_torch"] = timeit(lambda: y.copy_(x))
    T["copy_48Mi_torch"]["GBs"] = 16 * n / T["copy_48Mi_torch"]["min_ms"] / 1e6
    n2 = 1 << 20
    x2 = x[:n2]; y2 = y[:n2]
    T["daxpy_sum_1Mi"] = timeit(lambda: lib.gmt_daxpy_sum_f64(n2, 2.0, x2.data_ptr(), y2.data_ptr(), out.data_ptr(),
                                                              scratch.data_ptr(), st()), iters=50)
    del x, y
    n = 1 << 27
    yin = torch.rand(n + 4, dtype=f64, device=dev)
    o = torch.empty(n, dtype=f64, device=dev)
    for v in (0, 1, 2):
        T[f"sten1d_v{v}_2^27"] = timeit(lambda: lib.gmt_stencil1d_f64(n, yin.data_ptr(), o.data_ptr(), 1.0, v, st()))
        T[f"sten1d_v{v}_2^27"]["GLUPs"] = n / T[f"sten1d_v{v}_2^27"]["min_ms"] / 1e6
    del yin, o
    for (nx, ny, tag) in ((1024, 524288, "ref_d0"), (16384, 16384, "sq")):
        z = torch.rand(ny, nx + 4, dtype=f64, device=dev)
        o = torch.empty(ny, nx, dtype=f64, device=dev)
        for v in (0, 1):
            k = f"sten2d_d0_v{v}_{tag}"
            T[k] = timeit(lambda: lib.gmt_stencil2d_f64(0, nx, ny, z.data_ptr(), nx + 4, o.data_ptr(), nx, 1.0, v, st()))
            T[k]["GLUPs"] = nx * ny / T[k]["min_ms"] / 1e6
        del z, o
    for (nx, ny, tag) in ((524288, 1024, "ref_d1"), (16384, 16384, "sq")):
        z = torch.rand(ny + 4, nx, dtype=f64, device=dev)
        o = torch.empty(ny, nx, dtype=f64, device=dev)
        for v in (0, 1):
            k = f"sten2d_d1_v{v}_{tag}"
            T[k] = timeit(lambda: lib.gmt_stencil2d_f64(1, nx, ny, z.data_ptr(), nx, o.data_ptr(), nx, 1.0, v, st()))
            T[k]["GLUPs"] = nx * ny / T[k]["min_ms"] / 1e6
        del z, o
    nx = ny = 16384
    z = torch.rand(ny + 4, nx + 4, dtype=f64, device=dev)
    o = torch.empty(ny, nx, dtype=f64, device=dev)
    T["sten2d_xy_sq"] = timeit(lambda: lib.gmt_stencil2d_xy_f64(nx, ny, z.data_ptr(), nx + 4, o.data_ptr(), nx, 1.0, 1.0, st()))
    T["sten2d_xy_sq"]["GLUPs"] = nx * ny / T["sten2d_xy_sq"]["min_ms"] / 1e6
    del z, o
    # fused kernels, world of one and emulated 2x2 (timing only indicative for R>1)
    for tag, args in (("fused_grid_1x1_16384", (1, 1, 16384, 16384, 2, 2)), ("fused_d0_1_ref", (1, 1, 1024, 524288, 2, 0)),
                      ("fused_d1_1_ref", (1, 1, 524288, 1024, 0, 2))):
        for we in (False, True):
            r = fused2d_emulated(*args, iters=1, with_err=we, time_it=True)
            r["GLUPs"] = args[2] * args[3] / r["time"]["min_ms"] / 1e6
            T[f"{tag}_err{int(we)}"] = r
    print(json.dumps(T, indent=1))

(OUT / "probe.json").write_text(json.dumps(res, indent=1))
print("probe done")

# ---- TMA / cluster variants at benchmark shapes (appended probe)
if "--tma" in sys.argv:
    from gpu_mpi_tests_b200.ops import local as _local
    T2 = {}
    n = 1 << 27
    yin = torch.rand(n + 4, dtype=f64, device=dev)
    o = torch.empty(n, dtype=f64, device=dev)
    for v in ("vector", "tma"):
        r = timeit(lambda: _local.stencil1d(yin, 1.0, out=o, variant=v))
        T2[f"sten1d_{v}"] = {"ms": r["min_ms"], "GLUPs": n / r["min_ms"] / 1e6}
    del yin, o
    for (nx, ny) in ((524288, 1024), (16384, 16384)):
        z = torch.rand(ny + 4, nx, dtype=f64, device=dev)
        o = torch.empty(ny, nx, dtype=f64, device=dev)
        for v in ("vector", "tma"):
            r = timeit(lambda: _local.stencil2d(z, 1, 1.0, out=o, variant=v))
            T2[f"sten2d_d1_{v}_{nx}x{ny}"] = {"ms": r["min_ms"], "GLUPs": nx * ny / r["min_ms"] / 1e6}
        del z, o
    print(json.dumps(T2, indent=1))
    (OUT / "probe_tma.json").write_text(json.dumps(T2, indent=1))
