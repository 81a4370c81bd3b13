#!/usr/bin/env python
"""Per-kernel SASS hot spots from an .ncu-rep (`--page source --csv`): where the warp-stall
samples land (by opcode and the top instructions) and which memory/synchronisation opcodes the
kernel contains.  Writes profiles/source_hotspots.md.

    python tools/ncu_hotspots.py [gpurun_out/prof_all.ncu-rep]
    python tools/ncu_hotspots.py gpurun_out/prof_round2.ncu-rep      (round-2 kernels -> source_hotspots_round2.md)"""
import collections
import csv
import io
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
REP = Path(sys.argv[1]) if len(sys.argv) > 1 else ROOT / "gpurun_out" / "prof_all.ncu-rep"
# (kernel regex, how many earlier launches of it to skip so that a warm one is read)
KERNELS = [("fused2d_kernel", 3),   # 4th launch = 2x2 virtual ranks in one cooperative launch (real exchange)
           ("fused_dim0_ws_kernel", 1), ("fused1d_kernel", 1),
           ("stencil2d_d1_cluster_kernel", 1), ("stencil1d_tma_kernel", 1), ("daxpy_sum_vec4_kernel", 1)]
OUT_NAME = "source_hotspots.md"
if "round2" in REP.name:
    # exchange alone (2 virtual ranks, 4 MiB strided faces), dim-0 step with ghosts from the inbox, fused 1-D on the
    # TMA path, daxpy + global sum (n = 48 Mi: third launch), flag-based and LL allgather, two-shot all-reduce
    KERNELS = [("halo_exchange2d_kernel", 1), ("fused_dim0_ws_kernel", 1), ("fused1d_tma_kernel", 1),
               ("daxpy_allreduce_kernel", 3), ("gather_kernel", 1), ("gather_ll_kernel", 1), ("axis_sum_allreduce_kernel", 1)]
    OUT_NAME = "source_hotspots_round2.md"
MEM = ("LD", "ST", "ATOM", "RED", "MEMBAR", "ERRBAR", "UBLKCP", "UTMA", "SYNCS", "UCGABAR", "NANOSLEEP")


def opcode(sass: str) -> str:
    parts = sass.split()
    if not parts:
        return ""
    return parts[1] if parts[0].startswith("@") and len(parts) > 1 else parts[0]


def main():
    out = ["# SASS hot spots (ncu `--page source`, warp-stall samples)", "",
           f"Source: `{REP.name}` (captured with `--set full --import-source on`, library built with `-lineinfo`).  "
           "For HBM-bound streaming kernels the samples are expected to pile up on the first arithmetic instruction "
           "that consumes a global load (`long_scoreboard`).", ""]
    for regex, skip in KERNELS:
        res = subprocess.run(["ncu", "-i", str(REP), "--page", "source", "--csv", "--kernel-name", f"regex:{regex}",
                              "--launch-skip", str(skip), "--launch-count", "1"], capture_output=True, text=True)
        rows = list(csv.reader(io.StringIO(res.stdout)))
        if len(rows) < 3:
            continue
        name = rows[0][1] if len(rows[0]) > 1 else regex
        hdr = rows[1]
        si = hdr.index("Warp Stall Sampling (All Samples)")
        seen, data = set(), []
        for r in rows[2:]:
            if len(r) <= si or not r[si].isdigit() or r[0] in seen:
                continue
            seen.add(r[0])
            data.append((int(r[si]), r[1].strip()))
        total = sum(n for n, _s in data) or 1
        by_op = collections.Counter()
        mem = collections.Counter()
        for n, s in data:
            op = opcode(s)
            by_op[op] += n
            if any(k in op for k in MEM):
                mem[op] += 1
        out.append(f"## `{name.split('(const')[0].split('(long')[0].replace('gmt::', '')}`")
        out.append("")
        out.append(f"{len(data)} SASS instructions, {total} stall samples.")
        out.append("")
        out.append("Samples by opcode: " + ", ".join(f"`{op}` {100 * n / total:.1f}%" for op, n in by_op.most_common(6)))
        out.append("")
        out.append("Top instructions: " + "; ".join(f"`{s[:60]}` {100 * n / total:.1f}%"
                                                    for n, s in sorted(data, reverse=True)[:4]))
        out.append("")
        out.append("Memory / synchronisation opcodes present (static count): " +
                   ", ".join(f"`{op}` x{n}" for op, n in sorted(mem.items())))
        out.append("")
    (ROOT / "profiles" / OUT_NAME).write_text("\n".join(out) + "\n")
    print(ROOT / "profiles" / OUT_NAME)


if __name__ == "__main__":
    main()
