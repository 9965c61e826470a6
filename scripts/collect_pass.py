#!/usr/bin/env python
"""Copy the evidence of one GPU pass from gpurun_out/ (scratch) into profiles/ (tracked): bench JSON lines, the ncu launch
list of the bench command with a per-kernel summary, and the --set full summary of the steady-state trace launch.
usage: collect_pass.py <tag>      (e.g. x for scripts/gpu_r1_x.sh)"""
import collections, csv, json, pathlib, re, shutil, subprocess, sys

tag = sys.argv[1]
root = pathlib.Path(__file__).resolve().parents[1]
src, dst = root / "gpurun_out", root / "profiles"
names = {f"bench_r1_{tag}": "bench", f"bench_r1_{tag}_ref": "bench_reference", f"bench_r1_{tag}_cfg2": "bench_config2",
         f"bench_r1_{tag}_cfg4": "bench_config4_16frames", f"bench_r1_{tag}_cfg5": "bench_config5_64frames"}
share = None
for a, b in names.items():
    p = src / f"{a}.json"
    if p.exists():
        shutil.copy(p, dst / f"r1_{tag}_{b}.json")
        d = json.loads(p.read_text().strip().splitlines()[-1])
        r = d.get("roofline", {})
        if b == "bench": share = r.get("kernel_share_of_step")
        print(f"{b:24s} value {d['value']:9.1f}  e2e {d['e2e']['value']:9.1f}  ms/step {d['ms_per_step']:7.1f}  trace share {r.get('kernel_share_of_step')}"
              f"  launch_ms {r.get('launch_ms')}  cpu {d.get('cpu_baseline', {}).get('value')}")
lst = src / f"launches_bench_{tag}.csv"
if lst.exists():
    rows = [r for r in csv.reader(lst.open()) if len(r) > 14 and r[0].isdigit()]
    agg = collections.OrderedDict()
    for r in rows:
        a = agg.setdefault(re.sub(r"\(.*", "", r[4]), [0, 0.0]); a[0] += 1; a[1] += float(r[14])
    tot = sum(a[1] for a in agg.values())
    out = [f"ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 400: python bench.py --steps 1 --warmup 3 --frames 8 --no-cpu-baseline (scripts/gpu_r1_{tag}.sh)",
           f"launches 400..799 of the run (cold-cache, serialised: compare SHARES with bench.py's live kernel_share_of_step = {share}); full list profiles/r1_{tag}_launches_bench.csv"]
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        out.append("%-52s n=%4d sum=%9.3f ms share=%5.1f%%  mean=%8.1f us" % (n, a[0], a[1] / 1e6, 100 * a[1] / tot, a[1] / a[0] / 1e3))
    (dst / f"r1_{tag}_launches_bench_summary.txt").write_text("\n".join(out) + "\n")
    (dst / f"r1_{tag}_launches_bench.csv").write_text("".join(l for l in lst.open() if not l.startswith("==")))
    print("\n".join(out[2:]))
rep = src / f"prof_wf_{tag}_trace.ncu-rep"
if rep.exists():
    raw = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines())); h, u, v = rows[0], rows[1], rows[-1]
    keep = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
            "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "launch__grid_size", "launch__block_size", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "smsp__inst_executed.sum",
            "smsp__average_warp_latency_per_inst_issued.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
            "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio")
    out = [f"ncu --set full --clock-control none --import-source on -k regex:k_wf_trace -s 30 -c 1  (scripts/gpu_r1_{tag}.sh), steady-state wave",
           f"mrt_render --scene spheres-1m --size 1920x1080 --seed 3 --frames 8 --pipeline wavefront; source report gpurun_out/{rep.name}"]
    out += ["%-90s %s %s" % (w, v[h.index(w)], u[h.index(w)]) for w in keep if w in h]
    lines = subprocess.run([sys.executable, str(root / "scripts/ncu_lines.py"), str(rep), "20"], capture_output=True, text=True).stdout
    (dst / f"r1_{tag}_wf_trace_steady_ncu_summary.txt").write_text("\n".join(out) + "\n\n" + lines)
    print("\n".join(out[2:12]))
