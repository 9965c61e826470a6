#!/usr/bin/env python
"""SASS of the production trace kernel from the built library: full listing + opcode histogram, so that the instruction
budget DESIGN.md quotes (node step / triangle step) can be checked.  usage: sass_excerpt.py [out prefix]"""
import collections, pathlib, re, subprocess, sys

root = pathlib.Path(__file__).resolve().parents[1]
lib = root / "metal-renderer_b200" / "lib" / "libmrt_b200.so"
prefix = sys.argv[1] if len(sys.argv) > 1 else str(root / "profiles" / "sass_k_wf_trace")
sass = subprocess.run(["cuobjdump", "-sass", str(lib)], capture_output=True, text=True).stdout
blocks = re.split(r"\n\s*Function : ", sass)
commit = subprocess.run(["git", "-C", str(root), "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
for want, tag in (("k_wf_traceILi8ELi1ELb0E", "trace"), ("k_wf_shadeILb0ELi4ELb0E", "shade")):
    body = next(b for b in blocks if want in b.split("\n", 1)[0])
    lines = [l for l in body.splitlines() if re.match(r"\s+/\*[0-9a-f]{4}\*/", l)]
    ops = collections.Counter()
    for l in lines:
        m = re.match(r"\s+/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m:
            ops[m.group(1)] += 1
    out = [f"# cuobjdump -sass of {body.splitlines()[0].strip()} ({lib.name}, sm_100a), built from commit {commit}",
           f"# {len(lines)} instructions; opcode histogram (static counts, whole kernel):"]
    out += [f"{n:6d}  {op}" for op, n in ops.most_common()]
    base = collections.Counter()
    for op, n in ops.items():
        base[op.split(".")[0]] += n
    out += ["", "# by base opcode:"] + [f"{n:6d}  {op}" for op, n in base.most_common()]
    name = prefix.replace("trace", tag)
    pathlib.Path(name + "_opcodes.txt").write_text("\n".join(out) + "\n")
    pathlib.Path(name + ".sass").write_text(out[0] + "\n" + "\n".join(lines) + "\n")
    print(name, len(lines), dict(base.most_common(12)))

# ---- static instruction counts per source file / line of the trace kernel (nvdisasm -g on the embedded cubin) ----------
import tempfile, os
with tempfile.TemporaryDirectory() as tmp:
    subprocess.run(["cuobjdump", "-xelf", "all", str(lib)], cwd=tmp, capture_output=True)
    cubin = os.path.join(tmp, "wavefront.sm_100a.cubin")
    dis = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
m0 = re.search(r"^\.text\._ZN3mrt10k_wf_traceILi8ELi1ELb0E[^\n]*:\n", dis, re.M)
body = dis[m0.end():] if m0 else ""
body = body.split("//--------------------- ", 1)[0]
per_line, per_file, cur = collections.Counter(), collections.Counter(), ("?", 0)
for l in body.splitlines():
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    if re.match(r"\s+/\*[0-9a-f]{4}\*/", l):
        per_line[cur] += 1; per_file[cur[0]] += 1
out = [f"# static SASS instruction counts of k_wf_trace<8,1,false> by source position (nvdisasm -g), commit {commit}",
       f"total {sum(per_file.values())}"]
out += [f"{n:6d}  {f}" for f, n in per_file.most_common()]
out += ["", "# inlined helpers are attributed to their own lines, so a FILE is the unit to read:",
        f"bvh8.h = the 8-box test of a node step with its plane decode (mrt_plane / mrt_fma / min-max helpers) and the leaf expansion: {per_file['bvh8.h']}",
        f"device_trace.cuh + device_math.cuh = the ray/triangle test of a triangle step: {per_file['device_trace.cuh']} + {per_file['device_math.cuh']}",
        f"device_trace8.cuh + intrinsics = refill, vote, pop / push, node fetch, result store: {per_file['device_trace8.cuh']} + {per_file['sm_30_intrinsics.hpp'] + per_file['sm_32_intrinsics.hpp'] + per_file['device_atomic_functions.hpp']}",
        f"wavefront.cu = queue entry -> ray (WfIO::load), hit / visibility store, batch control of block 0: {per_file['wavefront.cu']}",
        "", "# per line, 25 largest:"]
out += [f"{n:6d}  {f}:{ln}" for (f, ln), n in per_line.most_common(25)]
pathlib.Path(prefix + "_by_source.txt").write_text("\n".join(out) + "\n")
print("\n".join(out[:14]))
