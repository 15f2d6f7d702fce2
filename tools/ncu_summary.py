"""Summarise ncu --set full reports (gpurun_out/*.ncu-rep) into a markdown table + profiles/ncu_traffic.json.
python tools/ncu_summary.py gpurun_out/r01_full_*.ncu-rep > profiles/r01_ncu_full_C2.md"""
import csv
import io
import json
import subprocess
import sys
from pathlib import Path

KEYS = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "dram rd"), ("dram__bytes_write.sum", "dram wr"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma pipe %"),
        ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "alu pipe %"),
        ("sm__inst_issued.avg.pct_of_peak_sustained_active", "issue %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp insts")]
UNIT = {"Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "Gbyte": 1e9}
traffic = {}
counters = {}
print("| kernel | " + " | ".join(k for _, k in KEYS) + " | top stalls (pc samples) |")
print("|---|" + "---|" * (len(KEYS) + 1))
for rep in sys.argv[1:]:
    raw = Path(rep).read_text() if rep.endswith(".csv") else \
        subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    seen = {}
    for r in rows[2:]:                 # one line per distinct kernel of the report; the last launch of each wins
        if len(r) == len(hdr) and "at::" not in r[idx["Kernel Name"]]:      # the path's kernels, not torch's
            seen[r[idx["Kernel Name"]] + "|" + r[idx["Grid Size"]]] = r
    if not seen:
        seen = {"last": rows[-1]}
    for r in seen.values():
        name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "")
        cells = []
        for key, _ in KEYS:
            cells.append(f"{r[idx[key]]} {units[idx[key]]}".strip() if key in idx else "-")
        stalls = sorted(((float(r[i]), h.replace("smsp__pcsamp_warps_issue_stalled_", "")) for h, i in idx.items()
                         if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued") and r[i]), reverse=True)[:4]
        print(f"| `{name}` grid {r[idx['Grid Size']]} | " + " | ".join(cells) + " | " + ", ".join(f"{n} {int(v)}" for v, n in stalls) + " |")

        def b(key):
            return float(r[idx[key]]) * UNIT.get(units[idx[key]], 1.0)
        short = next((k for k in ("raster_bwd", "raster_fwd", "bin_small", "scan_kernel", "onesweep", "decode_project", "project_fwd",
                                  "project_bwd", "tile_sort_pack", "tile_count_scatter", "tile_scan", "repack", "adan") if k in name), name)
        if short == "raster_fwd" and "true, false, true" in r[idx["Kernel Name"]].replace("1, 0, 1", "true, false, true"):
            short = "raster_fwd_loss"
        traffic[short + ("@c5" if Path(rep).stem.endswith("c5") else "")] = {
            "dram_bytes": b("dram__bytes_read.sum") + b("dram__bytes_write.sum"), "us": float(r[idx["gpu__time_duration.sum"]]),
            "report": Path(rep).name}

        def pct(key):
            return float(r[idx[key]]) if key in idx and r[idx[key]] not in ("", "n/a") else None
        counters[short + ("@c5" if Path(rep).stem.endswith("c5") else "")] = {
            "duration_us": float(r[idx["gpu__time_duration.sum"]]),
            "fma_pipe_pct": pct("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
            "alu_pipe_pct": pct("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active"),
            "issue_active_pct": pct("sm__inst_issued.avg.pct_of_peak_sustained_active"),
            "achieved_occupancy_pct": pct("sm__warps_active.avg.pct_of_peak_sustained_active"),
            "dram_throughput_pct": pct("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"),
            "dram_bytes": b("dram__bytes_read.sum") + b("dram__bytes_write.sum"),
            "warp_instructions": float(r[idx["smsp__inst_executed.sum"]]) if "smsp__inst_executed.sum" in idx else None,
            "registers_per_thread": float(r[idx["launch__registers_per_thread"]]) if "launch__registers_per_thread" in idx else None,
            "grid": r[idx["Grid Size"]], "top_stalls": [n for _, n in stalls],
            "report": Path(rep).name, "how": "ncu --set full --clock-control none, one launch, L2 flushed before it (tools/prof_raster.py)"}
Path("profiles/ncu_traffic.json").write_text(json.dumps(traffic, indent=1) + "\n")
Path("profiles/ncu_counters.json").write_text(json.dumps(counters, indent=1) + "\n")
