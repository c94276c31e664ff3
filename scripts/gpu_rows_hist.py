"""Diagnostic: SA rows per read of the default bench workload (who is heavy in chaining)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from bwa_mem_scale_b200 import Context, FMI_search, pack_reads, seed_opt
class A: ref_mbp=100.0; reads=1_000_000; exact_frac=None
ref, idx, reads = bench.seed_inputs(A, 0)
ctx = Context(0); fmi = FMI_search(idx, ctx=ctx); fmi.build_fma()
flat, off, ln = pack_reads(reads)
d = [ctx.upload(x) for x in (flat, off, ln)]
s = fmi.seed_device(seed_opt(), d[0], d[1], d[2], None, len(ln), 151)
coff = np.zeros(s.n_smem + 1, np.int64); soff = np.zeros(len(ln) + 1, np.int64)
ctx.d2h(coff, s.d_coord_off); ctx.d2h(soff, s.d_smem_off)
rows = coff[soff[1:]] - coff[soff[:-1]]
print("reads", len(rows), "rows total", rows.sum(), "mean", rows.mean())
print("percentiles 50/90/99/99.9/99.99/max", np.percentile(rows, [50, 90, 99, 99.9, 99.99, 100]))
for t in (32, 64, 128, 256, 512, 1024, 2048, 4096):
    m = rows > t
    print(f"> {t:5d}: {m.mean()*100:7.3f}% of reads, {rows[m].sum()/rows.sum()*100:6.2f}% of rows, sum rows^2 share {((rows[m].astype(float))**2).sum()/((rows.astype(float))**2).sum()*100:6.2f}%")
w = rows.reshape(-1, 32)
print("per-warp max: mean", w.max(1).mean(), " mean of per-warp sum", w.sum(1).mean(), " efficiency sum/(32*max)", (w.sum(1) / (32.0 * np.maximum(w.max(1), 1))).mean())
print("sum over warps of max rows:", w.max(1).sum(), " vs total rows", rows.sum())
