#!/usr/bin/env python
"""Per-source-line summary of an ncu report: share of warp instructions, average active threads, share of stall samples.
usage: ncu_lines.py report.ncu-rep [top]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]; top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur = None; hdr = None
agg = collections.defaultdict(lambda: [0, 0, 0, ""])
for r in rows:
    if not r: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if r[0] == "Function Name" or hdr is None: continue
    try: line = int(r[0])
    except ValueError: continue
    ie, te, sm = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("# Samples")
    try:
        a = agg[(cur, line)]; a[0] += int(r[ie] or 0); a[1] += int(r[te] or 0); a[2] += int(r[sm] or 0); a[3] = r[1][:100]
    except (ValueError, IndexError): pass
tot = sum(a[0] for a in agg.values()) or 1; ts = sum(a[2] for a in agg.values()) or 1
print("total warp inst", tot, "samples", ts, "avg threads %.2f" % (sum(a[1] for a in agg.values()) / tot))
pf = collections.defaultdict(lambda: [0, 0, 0])
for (f, l), a in agg.items():
    pf[f][0] += a[0]; pf[f][1] += a[1]; pf[f][2] += a[2]
for f, a in sorted(pf.items(), key=lambda kv: -kv[1][0]):
    print(f"{f:32s} {100*a[0]/tot:5.1f}% inst  avg thr {a[1]/max(a[0],1):5.1f}  {100*a[2]/ts:5.1f}% samples")
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][2])[:top]:
    print(f"{f}:{l:<4d} {100*a[0]/tot:5.1f}%i thr {a[1]/max(a[0],1):5.1f} {100*a[2]/ts:5.1f}%s  {a[3]}")
