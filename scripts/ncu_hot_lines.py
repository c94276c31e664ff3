#!/usr/bin/env python
"""Aggregates an ncu SASS source page by CUDA source line (needs -lineinfo):
   python scripts/ncu_hot_lines.py <rep> <object.o> <kernel-name-substring> [top]"""
import csv, re, subprocess, sys, tempfile, os, glob, collections
rep, obj, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if len(r) > 5 and r[0] == "Address"][0]
h = rows[hi]
si, ie, at = h.index("# Samples"), h.index("Instructions Executed"), h.index("Avg. Threads Executed")
ins = [r for r in rows[hi + 1:] if len(r) > at]
base = int(ins[0][0], 16)
td = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=td, check=True, capture_output=True)
cub = glob.glob(os.path.join(td, "*.cubin"))[0]
dis = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout
# walk the disassembly of the wanted function: remember the current line, map offset -> (file,line)
cur = None; infn = False; off2line = {}
for ln in dis.splitlines():
    m = re.match(r"\s*//## File \"(.*)\", line (\d+)", ln)
    if ".text." in ln and ln.strip().startswith(".section"):
        infn = kern in ln
    if not infn: continue
    if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m: off2line[int(m.group(1), 16)] = cur
agg = collections.defaultdict(lambda: [0, 0, 0.0])
ts = ti = 0
for r in ins:
    off = int(r[0], 16) - base
    key = off2line.get(off, ("?", 0))
    s, i = int(r[si] or 0), int(r[ie] or 0)
    agg[key][0] += s; agg[key][1] += i; agg[key][2] += i * float(r[at] or 0)
    ts += s; ti += i
print(f"total samples {ts}, warp instructions {ti}")
for k, v in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
    print(f"{k[0]}:{k[1]:<5d} samples {100*v[0]/max(ts,1):5.1f}%  inst {100*v[1]/max(ti,1):5.1f}%  avg-threads {v[2]/max(v[1],1):5.1f}")
