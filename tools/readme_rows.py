#!/usr/bin/env python
"""Rewrite the measured table of README.md from profiles/bench_n{1,8}.json (run after new records are committed)."""
import json
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]


def rec(n):
    return json.loads((ROOT / "profiles" / f"bench_n{n}.json").read_text().strip().splitlines()[-1])


r1, r8 = rec(1), rec(8)
e1, e8, b8 = r1["extra"], r8["extra"], r8["baseline_arm"]
hx = lambda e, k, w: e["halo_exchange_us"][k][w]
keys = ("1d_16B", "grid_faces_4x256KiB", "dim0_8MiB_strided", "dim1_8MiB_contiguous")
d1, d8 = e1["daxpy_sum"], e8["daxpy_sum"]
rows = [
    f"| 2-D stencil 16384² per GPU, ghost exchange fused in (bench.py `value`) | {r1['value']:.0f} GLUP/s, {100 * r1['roofline']['frac_of_measured_copy_peak']:.0f}% of measured copy BW | {r8['value']:.0f} GLUP/s ({100 * r1['ms_per_step'] / r8['ms_per_step']:.1f}% weak-scaling efficiency) | {b8['value']:.0f} GLUP/s (`vs_baseline` {r8['vs_baseline']:.2f}) |",
    f"| the same end to end (per-step H2D of boundary values, kernel, result on the host; pipelined / synchronous) | {r1['e2e']['value']:.0f} / {r1['e2e']['sync_value']:.0f} GLUP/s | {r8['e2e']['value']:.0f} / {r8['e2e']['sync_value']:.0f} GLUP/s | {b8['e2e']:.0f} GLUP/s ({b8['e2e_ratio']:.1f}x) |",
    f"| 1-D stencil 2²⁷ per GPU | {e1['stencil1d_2p27']['fused']['glups']:.0f} GLUP/s (105%) | {e8['stencil1d_2p27']['fused']['glups']:.0f} GLUP/s | {e8['stencil1d_2p27']['lib']['glups']:.0f} GLUP/s |",
    f"| reference-default 2-D tile 1024×524288, dim 0 / dim 1 | {e1['stencil2d_ref_dim0']['fused']['ms_per_step']:.3f} / {e1['stencil2d_ref_dim1']['fused']['ms_per_step']:.3f} ms | {e8['stencil2d_ref_dim0']['fused']['ms_per_step']:.3f} / {e8['stencil2d_ref_dim1']['fused']['ms_per_step']:.3f} ms (dim 0 with ghosts from the inbox: {e8['stencil2d_ref_dim0_inbox']['ms_per_step']:.3f}) | {e8['stencil2d_ref_dim0']['lib']['ms_per_step']:.3f} / {e8['stencil2d_ref_dim1']['lib']['ms_per_step']:.3f} ms |",
    "| halo exchange ALONE (the reference's timed quantity): 16 B / four 256 KiB faces / 8 MiB strided / 8 MiB contiguous | "
    + " / ".join(f"{hx(e1, k, 'fused'):.1f}" for k in keys) + " µs | " + " / ".join(f"{hx(e8, k, 'fused'):.1f}" for k in keys)
    + " µs | " + " / ".join(f"{hx(e8, k, 'lib'):.0f}" for k in keys) + " µs |",
    f"| daxpy + global sum, n = 2²⁰ per GPU | {d1['n_2p20']['fused']['us']:.1f} µs | {d8['n_2p20']['fused']['us']:.1f} µs | {d8['n_2p20']['lib']['us']:.1f} µs |",
    f"| daxpy + global sum, n = 48 Mi / N | {d1['n_48Mi_over_N']['fused']['us']:.0f} µs (101% of copy BW) | {d8['n_48Mi_over_N']['fused']['us']:.1f} µs ({100 * d8['n_48Mi_over_N']['fused']['gbs_per_gpu'] / 6555.2:.0f}%) | {d8['n_48Mi_over_N']['lib']['us']:.1f} µs |",
    f"| 8 KiB in-place all-reduce / barrier | {e1['allreduce_8KiB_us']['fused']:.1f} / {e1['barrier_us']['fused']:.1f} µs | {e8['allreduce_8KiB_us']['fused']:.1f} / {e8['barrier_us']['fused']:.1f} µs | {e8['allreduce_8KiB_us']['lib']:.1f} / {e8['barrier_us']['lib_allreduce_4B']:.1f} µs (NCCL) |",
    f"| in-place allgather, 64 KiB / 256 MiB slices | – | {e8['allgather']['64KiB']['fused_us']:.1f} µs / {e8['allgather']['256MiB']['fused_ingress_gbs_per_gpu']:.0f} GB/s ingress per GPU | {e8['allgather']['64KiB']['lib_us']:.1f} µs / {e8['allgather']['256MiB']['lib_ingress_gbs_per_gpu']:.0f} GB/s (NCCL) |",
]
p = ROOT / "README.md"
s = p.read_text()
a = s.index("| 2-D stencil 16384² per GPU, ghost exchange fused in")
b = s.index("## Layout")
p.write_text(s[:a] + "\n".join(rows) + "\n\n" + s[b:])
print("README.md table rewritten")
