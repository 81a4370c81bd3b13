#!/usr/bin/env python
"""Turn the .ncu-rep captures in gpurun_out/ into committed text: profiles/<name>.raw.csv (the
`--page raw --csv` export) and profiles/ncu_summary.md (the handful of metrics the roofline
argument rests on)."""
import csv
import io
import json
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
OUT = ROOT / "profiles"
WANT = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM % of peak"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "regs/thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
]


def main():
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    hbm = float(peaks.get("hbm_gbs", 6650.0))
    lines = ["# ncu captures (B200, `--set full --clock-control none --import-source on`)", "",
             f"Roofline denominator: measured copy bandwidth {hbm:.0f} GB/s (MEASURED_PEAKS.json).  Durations below are "
             "under the profiler (cold caches, serialised) -- the benchmark numbers in BASELINE.md are CUDA-event "
             "timings without it.  Raw exports: `profiles/*.raw.csv`.", ""]
    for rep in sorted((ROOT / "gpurun_out").glob("*.ncu-rep")):
        res = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True)
        if res.returncode != 0:
            continue
        (OUT / (rep.stem + ".raw.csv")).write_text(res.stdout)
        rows = list(csv.reader(io.StringIO(res.stdout)))
        hdr, units = rows[0], rows[1]
        lines.append(f"## {rep.name}")
        lines.append("")
        lines.append("| kernel | " + " | ".join(label for _m, label in WANT) + " | algorithmic GB/s | top warp stalls (warps per issue) |")
        lines.append("|---|" + "---|" * (len(WANT) + 2))
        # one row per kernel (and grid size): the LAST capture, i.e. a warm launch
        last = {}
        for r in rows[2:]:
            key = (r[hdr.index("Kernel Name")].split("(")[0].strip(), r[hdr.index("Grid Size")]
                   if "Grid Size" in hdr else "")
            last[key] = r
        stall_cols = [(i, h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")])
                      for i, h in enumerate(hdr)
                      if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")
                      and "not_issued" not in h]
        for (name, grid_dims), r in last.items():
            if grid_dims.replace(" ", "").split(",")[1:2] not in ([], ["1"]):
                name += f" grid {grid_dims}: {grid_dims.split(',')[1].strip()} virtual ranks, real exchange"
            cells = []
            vals = {}
            for metric, _label in WANT:
                if metric in hdr:
                    i = hdr.index(metric)
                    vals[metric] = (r[i], units[i])
                    cells.append(f"{r[i]} {units[i]}".strip())
                else:
                    cells.append("-")
            algo = "-"
            try:
                rd, ru = vals["dram__bytes_read.sum"]
                wr, wu = vals["dram__bytes_write.sum"]
                dur, du = vals["gpu__time_duration.sum"]
                scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
                tsc = {"us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}
                total = float(rd) * scale[ru] + float(wr) * scale[wu]
                algo = f"{total / (float(dur) * tsc[du]) / 1e9:.0f} (DRAM bytes/time; {total / (float(dur) * tsc[du]) / 1e9 / hbm:.2f} of measured)"
            except Exception:  # noqa: BLE001
                pass
            stalls = []
            for i, reason in stall_cols:
                try:
                    stalls.append((float(r[i]), reason))
                except ValueError:
                    pass
            top = ", ".join(f"{reason} {v:.1f}" for v, reason in sorted(stalls, reverse=True)[:3]) or "-"
            lines.append(f"| `{name}` | " + " | ".join(cells) + f" | {algo} | {top} |")
        lines.append("")
    (OUT / "ncu_summary.md").write_text("\n".join(lines) + "\n")
    print(OUT / "ncu_summary.md")


if __name__ == "__main__":
    main()
