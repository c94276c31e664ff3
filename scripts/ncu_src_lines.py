#!/usr/bin/env python
"""Per-CUDA-source-line summary of an ncu report captured with --import-source on:
   python scripts/ncu_src_lines.py <rep> [top]   (uses ncu's own line mapping: --page source --print-source cuda,sass)"""
import csv, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
fname, hdr, recs = "?", None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Name":
        fname = r[1].split("/")[-1]
    elif len(r) > 10 and r[0] == "Line No":
        hdr = r
    elif hdr and len(r) > 10 and r[0] not in ("", "Line No"):
        if r[-1] == "" and len(r) > len(hdr):
            r = r[:-1]
        g = lambda name: r[len(r) - (len(hdr) - hdr.index(name))]      # from the end: the source text may hold quotes and commas
        num = lambda name: float(g(name)) if g(name) not in ("", "-") else 0.0
        recs.append((fname, int(r[0]), r[1].strip(), num("# Samples"), num("Instructions Executed"), num("Thread Instructions Executed"),
                     num("stall_long_sb"), num("stall_short_sb"), num("stall_wait"), num("stall_branch_resolving") + num("stall_no_inst")))
ts, ti = sum(x[3] for x in recs), sum(x[4] for x in recs)
print(f"total samples {ts:.0f}, warp instructions {ti:.0f}, thread instr / warp instr {sum(x[5] for x in recs) / max(ti, 1):.2f}")
for x in sorted(recs, key=lambda x: -x[3])[:top]:
    print(f"{x[0]}:{x[1]:<4d} smp {100 * x[3] / ts:5.1f}%  inst {100 * x[4] / ti:5.1f}%  thr {x[5] / max(x[4], 1):4.1f}  long_sb {100 * x[6] / ts:4.1f}% short_sb {100 * x[7] / ts:4.1f}%  | {x[2][:110]}")
