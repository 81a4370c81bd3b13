#!/usr/bin/env python
"""Markdown rows of BASELINE.md section 4a from the committed records (profiles/bench_n*.json, sweep_n*.jsonl)."""
import json
from pathlib import Path

P = Path(__file__).resolve().parents[1] / "profiles"
HBM = 6555.2


def bench(n):
    return json.loads((P / f"bench_n{n}.json").read_text().strip().splitlines()[-1])


def sweep(n):
    return [json.loads(l) for l in (P / f"sweep_n{n}.jsonl").read_text().splitlines() if l.strip()]


print("| N | fused GLUP/s (ms/step) | library arm GLUP/s (ms) | `vs_baseline` | e2e pipelined GLUP/s (ms) | e2e synchronous (ms) | library e2e | e2e ratio | weak-scaling eff. | ghost mismatches |")
print("|---|---|---|---|---|---|---|---|---|---|")
t1 = bench(1)["ms_per_step"]
for n in (1, 2, 4, 8):
    r = bench(n)
    b = r["baseline_arm"]
    print(f"| {n} | **{r['value']:.0f}** ({r['ms_per_step']:.4f}) | {b['value']:.0f} ({b['ms_per_step']:.4f}) | {r['vs_baseline']:.3f} | "
          f"**{r['e2e']['value']:.0f}** ({r['e2e']['ms_per_step']:.4f}) | {r['e2e']['sync_value']:.0f} ({r['e2e']['sync_ms_per_step']:.4f}) | "
          f"{b['e2e']:.0f} ({b['e2e_ms_per_step']:.3f}) | {b['e2e_ratio']:.2f} | {100 * t1 / r['ms_per_step']:.1f}% | {r['ghost_mismatch']} |")
print()
print("| Row (fused / library) | N=1 | N=2 | N=4 | N=8 |")
print("|---|---|---|---|---|")


def row(label, fn):
    print(f"| {label} | " + " | ".join(fn(bench(n)["extra"]) for n in (1, 2, 4, 8)) + " |")


row("1-D stencil 2^27/GPU, GLUP/s", lambda e: f"**{e['stencil1d_2p27']['fused']['glups']:.0f}** / {e['stencil1d_2p27']['lib']['glups']:.0f}")
row("dim-0 slab 1024x524288, ms/step", lambda e: f"**{e['stencil2d_ref_dim0']['fused']['ms_per_step']:.4f}** / {e['stencil2d_ref_dim0']['lib']['ms_per_step']:.4f}")
row("dim-0 slab, ghosts from the inbox, ms/step", lambda e: f"**{e['stencil2d_ref_dim0_inbox']['ms_per_step']:.4f}**")
row("dim-1 slab 524288x1024, ms/step", lambda e: f"**{e['stencil2d_ref_dim1']['fused']['ms_per_step']:.4f}** / {e['stencil2d_ref_dim1']['lib']['ms_per_step']:.4f}")
row("16384^2 periodic grid (4 live faces on every rank), ms/step", lambda e: f"**{e['stencil2d_periodic4']['ms_per_step']:.4f}**")
for key, label in (("1d_16B", "exchange alone, 1-D 16 B, us"), ("grid_faces_4x256KiB", "exchange alone, four 256 KiB faces, us"),
                   ("dim0_8MiB_strided", "exchange alone, 8 MiB strided (dim 0), us"),
                   ("dim1_8MiB_contiguous", "exchange alone, 8 MiB contiguous (dim 1), us")):
    row(label, lambda e, k=key: f"**{e['halo_exchange_us'][k]['fused']:.1f}** / {e['halo_exchange_us'][k]['lib']:.1f}")
row("daxpy + global sum n = 2^20/GPU, us", lambda e: f"**{e['daxpy_sum']['n_2p20']['fused']['us']:.2f}** / {e['daxpy_sum']['n_2p20']['lib']['us']:.2f}")
row("daxpy + global sum n = 48 Mi/N, us (GB/s per GPU, % of measured copy)",
    lambda e: (lambda f, l: f"**{f['us']:.1f}** ({f['gbs_per_gpu']:.0f}, {100 * f['gbs_per_gpu'] / HBM:.0f}%) / {l['us']:.1f}")(
        e['daxpy_sum']['n_48Mi_over_N']['fused'], e['daxpy_sum']['n_48Mi_over_N']['lib']))
row("8 KiB in-place all-reduce, us", lambda e: f"**{e['allreduce_8KiB_us']['fused']:.2f}** / " + (f"{e['allreduce_8KiB_us']['lib']:.2f}" if e['allreduce_8KiB_us']['lib'] else "-"))
row("barrier, us", lambda e: f"**{e['barrier_us']['fused']:.2f}** / " + (f"{e['barrier_us']['lib_allreduce_4B']:.2f}" if e['barrier_us']['lib_allreduce_4B'] else "-"))
row("in-place allgather 64 KiB slices, us", lambda e: f"**{e['allgather']['64KiB']['fused_us']:.1f}** / " + (f"{e['allgather']['64KiB']['lib_us']:.1f}" if e['allgather']['64KiB']['lib_us'] else "-"))
row("in-place allgather 256 MiB slices, ingress GB/s per GPU", lambda e: (f"**{e['allgather']['256MiB']['fused_ingress_gbs_per_gpu']:.0f}** / {e['allgather']['256MiB']['lib_ingress_gbs_per_gpu']:.0f}" if e['allgather']['256MiB']['lib_us'] else "-"))
print()
print("| in-place allgather, slice per rank | " + " | ".join(f"N={n}: P2P / LL / NCCL (us; ingress GB/s of the best)" for n in (2, 4, 8)) + " |")
print("|---|---|---|---|")
sw = {n: sweep(n) for n in (2, 4, 8)}
sizes = sorted({r["slice_bytes"] for r in sw[8] if r["config"] == "allgather_inplace"})
for s in sizes:
    cells = []
    for n in (2, 4, 8):
        g = {r["impl"]: r for r in sw[n] if r["config"] == "allgather_inplace" and r["slice_bytes"] == s}
        if not g:
            cells.append("-")
            continue
        p2p, ll, nc = g.get("fused_p2p"), g.get("fused_ll"), g.get("nccl_allgather")
        best = min(x["us"] for x in (p2p, ll) if x)
        cells.append(f"{p2p['us']:.1f} / " + (f"{ll['us']:.1f}" if ll else "-") + f" / {nc['us']:.1f}; {(n - 1) * s / best / 1e3:.0f}")
    lab = f"{s >> 30} GiB" if s >= 1 << 30 else (f"{s >> 20} MiB" if s >= 1 << 20 else f"{s >> 10} KiB")
    print(f"| {lab} | " + " | ".join(cells) + " |")
