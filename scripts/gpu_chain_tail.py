"""Diagnostic: chaining / filter kernel time when the heaviest reads are left out (is it a tail?)."""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from bwa_mem_scale_b200 import Context, Extender, FMI_search, chain_opt, ext_opt, fmindex, mem_core_device, pack_reads, seed_opt
class A: ref_mbp=100.0; reads=1_000_000; exact_frac=None
ref, idx, reads = bench.seed_inputs(A, 0)
ctx = Context(0); fmi = FMI_search(idx, ctx=ctx); fmi.build_fma()
ext = Extender(fmindex.text_with_revcomp(ref), len(ref), [0], [len(ref)], ctx=ctx)
flat, off, ln = pack_reads(reads)
d = [ctx.upload(x) for x in (flat, off, ln)]
s = fmi.seed_device(seed_opt(), d[0], d[1], d[2], None, len(ln), 151)
coff = np.zeros(s.n_smem + 1, np.int64); soff = np.zeros(len(ln) + 1, np.int64)
ctx.d2h(coff, s.d_coord_off); ctx.d2h(soff, s.d_smem_off)
rows = coff[soff[1:]] - coff[soff[:-1]]
reads = np.asarray(reads)
for cap in (10**9, 500, 400, 300, 200, 100, 32):
    sel = reads[rows <= cap]
    f, o, l = pack_reads(sel)
    dd = [ctx.upload(x) for x in (f, o, l)]
    for rep in range(2):
        ctx.set_profiling(rep == 1)
        mem_core_device(fmi, ext, seed_opt(), chain_opt(), ext_opt(), dd[0], dd[1], dd[2], len(l), 151)
    t = {k: ctx.kernel_time(k)[0] for k in (8, 7, 3, 0)}
    ctx.set_profiling(False)
    print(f"rows <= {cap:>10}: {len(l):7d} reads  chain {t[8]:6.2f} ms  c2a {t[7]:6.2f} ms  bsw {t[3]:6.2f} ms  seed {t[0]:6.2f} ms", flush=True)
    for p in dd: ctx.free(p)
