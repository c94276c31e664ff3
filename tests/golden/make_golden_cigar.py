"""Generates tests/golden/cigar_golden.npz from the UNMODIFIED reference (run in the build container only):
regions = what the reference's mem_kernel1_core + mem_chain2aln_across_reads_V2 leave for cigar_cases.golden_reads(),
records = the reference's own mem_reg2aln on every kept region (oracle/refwrap/ref_api_wrap.cpp: ref_reg2aln)."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import cigar_cases as gc        # noqa: E402
import fmi_cases as fc          # noqa: E402
import oracle_lib as ol         # noqa: E402
from bwa_mem_scale_b200 import fmindex  # noqa: E402


def main():
    ref, idx = fc.small_index()
    flat, off, ln = ol.pack_reads(gc.golden_reads(ref))
    out = dict(n_reads=np.int64(len(ln)))
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_index_files(prefix, ref, idx)
        rm = ol.RefMem(prefix)
        for name, kw in gc.OPT_CASES:
            o = ol.mem_opt(**kw)
            chains, seeds, regs, _ = rm.core(o, flat, off, ln)
            jobs = gc.jobs_from_regs(regs, chains, ol.CIGAR_JOB_DTYPE)
            outs, cigar, md = ol.reference_reg2aln(rm, o, flat, off, ln, jobs)
            assert outs["ok"].all()
            out[f"{name}_jobs"], out[f"{name}_outs"], out[f"{name}_cigar"], out[f"{name}_md"] = jobs, gc.out_matrix(outs), cigar, md
            print(name, len(jobs), "jobs", len(cigar), "ops", len(md), "MD bytes")
            if name == "default":
                # the same regions with an inflated local score: the global score stays below truesc - a, so mem_reg2aln
                # goes through its band-doubling retries (the kernel's second pass with full rows)
                jobs2 = jobs.copy()
                jobs2["truesc"] += 25
                outs, cigar, md = ol.reference_reg2aln(rm, o, flat, off, ln, jobs2)
                out["retry_jobs"], out["retry_outs"], out["retry_cigar"], out["retry_md"] = jobs2, gc.out_matrix(outs), cigar, md
    np.savez_compressed(gc.GOLDEN, **out)
    print("wrote", gc.GOLDEN, os.path.getsize(gc.GOLDEN), "bytes")


if __name__ == "__main__":
    main()
