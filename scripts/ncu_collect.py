#!/usr/bin/env python
"""Turn one `ncu --set full` report of a steady-state k_wf_trace / k_wf_shade launch into the committed summary that
bench.py parses for its roofline object: profiles/ncu_<kernel>_<workload>.txt.

usage: ncu_collect.py <report.ncu-rep> <trace|shade> <workload> <rays_in_launch|0> "<command that was profiled>" [output name]

Lines `name value unit` carry the raw counters (bench.py: ncu_summary); `#` lines carry provenance — the commit the
kernels were built from and the rays the profiled launch traced (from the wave log of the same command, so approximate:
work claiming is dynamic).  The per-source-line table of scripts/ncu_lines.py follows."""
import csv, pathlib, subprocess, sys

rep, kernel, workload, rays, cmd = sys.argv[1], sys.argv[2], sys.argv[3], float(sys.argv[4]), sys.argv[5]
root = pathlib.Path(__file__).resolve().parents[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, u, v = rows[0], rows[1], rows[-1]
keep = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
        "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "smsp__average_warp_latency_per_inst_issued.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio")
import os
commit = subprocess.run(["git", "-C", str(root), "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip() \
    or os.environ.get("MRT_COMMIT", "").strip() or ((root / "profiles" / "HEAD_COMMIT").read_text().strip() if (root / "profiles" / "HEAD_COMMIT").exists() else "unknown")
dirty = subprocess.run(["git", "-C", str(root), "status", "--porcelain", "--", "metal-renderer_b200/csrc"], capture_output=True, text=True).stdout.strip()
name = v[h.index("Kernel Name")] if "Kernel Name" in h else "?"
out = [f"# ncu --set full --clock-control none --import-source on, one steady-state launch of {name}",
       f"# command: {cmd}",
       f"# kernels built from commit {commit}{' + uncommitted changes in csrc/' if dirty else ''}",
       f"# rays_in_launch {rays:.0f}   (wave log of the same command without ncu; dynamic work claiming makes this approximate)"]
SCALE = {"Kbyte": (1e3, "byte"), "Mbyte": (1e6, "byte"), "Gbyte": (1e9, "byte"), "Tbyte": (1e12, "byte"),
         "us": (1e3, "ns"), "ms": (1e6, "ns"), "s": (1e9, "ns"), "usecond": (1e3, "ns"), "msecond": (1e6, "ns"), "second": (1e9, "ns"), "nsecond": (1, "ns")}
for w in keep:                      # values normalised to bytes and nanoseconds (ncu picks a unit per value)
    if w not in h:
        continue
    val, unit = v[h.index(w)].replace(",", ""), u[h.index(w)]
    if unit in SCALE:
        val, unit = "%.6g" % (float(val) * SCALE[unit][0]), SCALE[unit][1]
    out.append("%-90s %s %s" % (w, val, unit))
lines = subprocess.run([sys.executable, str(root / "scripts/ncu_lines.py"), rep, "24"], capture_output=True, text=True).stdout
dst = root / "profiles" / (sys.argv[6] if len(sys.argv) > 6 else f"ncu_{kernel}_{workload}.txt")
dst.write_text("\n".join(out) + "\n\n# per-source-line summary (scripts/ncu_lines.py): share of warp instructions, lanes, share of stall samples\n"
               + "\n".join("# " + l for l in lines.splitlines()) + "\n")
print("\n".join(out))
print(dst)
