"""Aggregates an ncu report's per-SASS-instruction samples by CUDA source line.

usage: ncu_lines.py <report.ncu-rep> <object.o> <source.cu> [top]
Needs the object compiled with -lineinfo; uses `ncu --page source --csv`, `cuobjdump -xelf`
and `nvdisasm -g` (all run here, no GPU)."""
import collections
import csv
import io
import os
import re
import subprocess
import sys
import tempfile

rep, obj, src = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
kernel = sys.argv[5] if len(sys.argv) > 5 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hi = next(i for i, r in enumerate(rows) if "Source" in r and "# Samples" in r)
h = rows[hi]
sa, ie = h.index("# Samples"), h.index("Instructions Executed")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=tmp, capture_output=True)
cubin = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout
# per function: offset -> (file, line)
funcs, cur, fn = {}, None, None
for ln in sass.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        fn = m.group(1)
        funcs[fn] = {}
        continue
    m = re.search(r'//## File "(.*)", line (\d+)', ln)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]+)\*/\s+\S", ln)
    if m and fn:
        funcs[fn][int(m.group(1), 16)] = cur
kname = rows[0][1] if rows[0] and rows[0][0] == "Kernel Name" else ""
# pick the function with as many instructions as the report rows
body = [r for r in rows[hi + 1:] if len(r) == len(h)]
cands = [f for f, d in funcs.items() if len(d) == len(body)]
if kernel:
    cands = [f for f in funcs if kernel in f] or cands
f = cands[0] if cands else max(funcs, key=lambda k: len(funcs[k]))
offs = sorted(funcs[f])
agg = collections.defaultdict(lambda: [0, 0])
base = int(body[0][0], 16)
for r in body:
    key = funcs[f].get(int(r[0], 16) - base)
    agg[key][0] += int(r[sa])
    agg[key][1] += int(r[ie])
tot = sum(v[0] for v in agg.values()) or 1
toti = sum(v[1] for v in agg.values()) or 1
text = open(src).read().split("\n")
print(f"kernel {kname}  function {f}  samples {tot}  warp-instructions {toti}")
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    s = text[k[1] - 1].strip()[:95] if k and k[0] == os.path.basename(src) else ""
    print(f"{v[0] / tot * 100:5.1f}% smp {v[1] / toti * 100:5.1f}% inst  {k}  {s}")
