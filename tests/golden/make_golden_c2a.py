"""Generates tests/golden/c2a_golden.npz from the REFERENCE ITSELF (oracle/_ref/libref_api_plain_sse41.so:
the reference's mem_kernel1_core and mem_chain2aln_across_reads_V2 run on the workloads of
tests/c2a_cases.py): the chains it builds (the input of the stage) and the alignment regions it leaves in
av_v (the output).  Run in the build container:  python tests/golden/make_golden_c2a.py"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import c2a_cases as cc  # noqa: E402
import oracle_lib as ol  # noqa: E402
from bwa_mem_scale_b200 import fmindex  # noqa: E402


def main():
    ref, ref_string = cc.genome()
    reads = cc.golden_reads(ref)
    flat, off, ln = ol.pack_reads(reads)
    out = {}
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_fasta(prefix + ".fa", ref)
        ol.ref_index(prefix, prefix + ".fa")
        rm = ol.RefMem(prefix)
        out["l_pac"] = np.array([rm.l_pac], np.int64)
        out["contig_off"], out["contig_len"] = rm.contig_off, rm.contig_len
        for name, kw in cc.OPT_CASES:
            chains, seeds, regs, reg_off = rm.core(ol.mem_opt(**kw), flat, off, ln)
            out[f"{name}_chains"], out[f"{name}_seeds"] = chains, seeds
            out[f"{name}_regs"] = cc.reg_matrix(regs).astype(np.int64)
            out[f"{name}_frac_rep"] = regs["frac_rep"]
    np.savez_compressed(os.path.join(HERE, "c2a_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()}, os.path.getsize(os.path.join(HERE, "c2a_golden.npz")))


def main_hard():
    """c2a_hard_golden.npz: the three-contig genome with gapped / boundary reads of c2a_cases.hard_reads."""
    ref, ref_string = cc.hard_genome()
    reads = cc.hard_reads(ref)
    flat, off, ln = ol.pack_reads(reads)
    out = {}
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_fasta_contigs(prefix + ".fa", ref, cc.HARD_CONTIGS)
        ol.ref_index(prefix, prefix + ".fa")
        assert np.array_equal(np.fromfile(prefix + ".0123", np.uint8), ref_string)
        rm = ol.RefMem(prefix)
        out["l_pac"] = np.array([rm.l_pac], np.int64)
        out["contig_off"], out["contig_len"] = rm.contig_off, rm.contig_len
        assert list(rm.contig_len) == list(cc.HARD_CONTIGS)
        for name, kw in cc.OPT_CASES_HARD:
            chains, seeds, regs, reg_off = rm.core(ol.mem_opt(**kw), flat, off, ln)
            out[f"{name}_chains"], out[f"{name}_seeds"] = chains, seeds
            out[f"{name}_regs"] = cc.reg_matrix(regs).astype(np.int64)
            out[f"{name}_frac_rep"] = regs["frac_rep"]
    np.savez_compressed(os.path.join(HERE, "c2a_hard_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()}, os.path.getsize(os.path.join(HERE, "c2a_hard_golden.npz")))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "hard":
        main_hard()
    else:
        main()
        main_hard()
