#!/usr/bin/env python3
"""Summarise `ncu --set full` reports (gpurun_out/prof_<tag>_<name>.ncu-rep) into profiles/<round>/:
the raw metric page of every captured launch as CSV, and one markdown table with the figures DESIGN.md quotes
(duration, DRAM bytes, pipe utilisation, occupancy, issue rate, top stall reasons).  Runs where ncu is installed
(the build container reads the reports the GPU box wrote)."""
import csv, io, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r02"
outdir = os.path.join(ROOT, "profiles", sys.argv[2] if len(sys.argv) > 2 else tag)
os.makedirs(outdir, exist_ok=True)
KEYS = [("gpu__time_duration.sum", "duration"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM written"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM % of peak"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe %"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe %"),
        ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe %"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe %"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads per warp instruction"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
        ("launch__registers_per_thread", "registers"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
        ("launch__shared_mem_per_block_dynamic", "dyn smem / CTA"), ("smsp__inst_executed.sum", "warp instructions"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %")]
lines = [f"# ncu --set full, {tag} (B200, --clock-control none; cold-cache, serialised launches)\n"]
for name in ("sigma", "var", "chain", "tree", "slash_mid"):
    rep = os.path.join(ROOT, "gpurun_out", f"prof_{tag}_{name}.ncu-rep")
    kept = os.path.join(outdir, f"prof_{name}_raw.csv")
    if os.path.exists(rep):
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        open(kept, "w").write(raw)
    elif os.path.exists(kept):                  # the report is scratch (gpurun_out/); its raw page was kept
        raw = open(kept).read()
    else:
        continue
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kn = hdr.index("Kernel Name")
    lines.append(f"\n## {name}: " + "; ".join(sorted({r[kn].split('(')[0].replace('void <unnamed>::', '') for r in data})) + "\n")
    lines.append("| metric | " + " | ".join(f"launch {i + 1}" for i in range(len(data))) + " |")
    lines.append("|---|" + "---|" * len(data))
    for key, label in KEYS:
        if key in hdr:
            i = hdr.index(key)
            lines.append(f"| {label} ({units[i]}) | " + " | ".join(r[i] for r in data) + " |")
    stalls = [h for h in hdr if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h]
    top = sorted(stalls, key=lambda h: -max(float(r[hdr.index(h)] or 0) for r in data))[:6]
    for h in top:
        i = hdr.index(h)
        lines.append(f"| stall {h[len('smsp__average_warps_issue_stalled_'):-len('_per_issue_active.ratio')]} (per issue) | " +
                     " | ".join(f"{float(r[i]):.2f}" for r in data) + " |")
open(os.path.join(outdir, "ncu_full_summary.md"), "w").write("\n".join(lines) + "\n")
print("\n".join(lines))
