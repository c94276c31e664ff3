#!/usr/bin/env python
"""Fills the @PLACEHOLDER@ numbers of DESIGN.md / README.md from the committed bench records:
   python scripts/fill_docs.py profiles/r02_bench_config3_final.json profiles/r02_bench_config3_final_reference_arm.json profiles/r02_bench_config3_8gpu.json"""
import json, sys
a, b, c = (json.load(open(p)) for p in sys.argv[1:4])
k = {x["kernel"].split("_kernel")[0]: x for x in a["roofline"]["kernels"]}
bsw, seed, sa = k["bsw_fast"], k["seed_collect"], k["sa_walk"]
v = {
    "VALUE": f"{a['value'] / 1e6:.2f}", "E2E": f"{a['e2e']['value'] / 1e6:.2f}", "REF": f"{b['value'] / 1e3:.0f}",
    "RATIO": f"{a['e2e']['value'] / b['value']:.0f}", "N8": f"{c['value'] / 1e6:.1f}", "GCUPS": f"{a['gcups']:.0f}",
    "SEEDMS": f"{seed['ms']:.0f}", "BSWMS": f"{bsw['ms']:.0f}", "CHAINMS": f"{k['chain_seeds']['ms']:.0f}",
    "C2AMS": f"{k['chain2aln']['ms']:.0f}", "SAMS": f"{sa['ms']:.0f}", "BSWFRAC": f"{bsw['frac']:.2f}",
    "SEEDGBS": f"{seed['achieved']:.0f}", "SEEDREQ": f"{seed['requested_gbs']:.0f}", "SEEDFRAC": f"{seed['frac']:.3f}",
    "SEEDREQFRAC": f"{seed['requested_frac']:.3f}", "SAGBS": f"{sa['achieved']:.0f}", "SAFRAC": f"{sa['frac']:.2f}",
}
for p in ("DESIGN.md", "README.md"):
    s = open(p).read()
    for key, val in v.items():
        s = s.replace(f"@{key}@", val)
    s = s.replace("ROUND2_HEADLINE", f"{v['VALUE']} M reads/s ({v['E2E']} M end to end)").replace("ROUND2_REFERENCE", f"{int(v['REF']) / 1e3:.2f} M reads/s")
    s = s.replace("ROUND2_8GPU", f"{v['N8']} M reads/s ({c['e2e']['value'] / 1e6:.1f} M end to end)").replace("ROUND2_GCUPS", f"{v['GCUPS']} GCUPS")
    open(p, "w").write(s)
print(v)
