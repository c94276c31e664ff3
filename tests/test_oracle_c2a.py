"""CPU tests: the oracle's restatement of mem_chain2aln_across_reads_V2 (oracle/c2a_oracle.c) against
golden (chains -> regions) records produced by the reference itself, and against the reference library
when oracle/_ref exists."""
import os

import numpy as np
import pytest

import c2a_cases as cc
import oracle_lib as ol
from bwa_mem_scale_b200 import fmindex, synth

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_golden.npz"))


@pytest.mark.parametrize("case", cc.OPT_CASES, ids=[c[0] for c in cc.OPT_CASES])
def test_oracle_matches_golden(case):
    name, kw = case
    ref, ref_string = cc.genome()
    flat, off, ln = ol.pack_reads(cc.golden_reads(ref))
    chains, seeds = GOLD[f"{name}_chains"], GOLD[f"{name}_seeds"]
    got = ol.oracle_chain2aln(ol.mem_opt(**kw), int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"],
                              ref_string, flat, off, ln, chains, seeds)
    assert np.array_equal(cc.reg_matrix(got), GOLD[f"{name}_regs"])
    assert np.array_equal(got["frac_rep"], GOLD[f"{name}_frac_rep"])
    g = GOLD[f"{name}_regs"]
    assert ((g[:, 2] == -1) & (g[:, 3] == -1)).sum() > 100            # the containment filter purged regions


HARD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_hard_golden.npz"))


@pytest.mark.parametrize("case", cc.OPT_CASES_HARD, ids=[c[0] for c in cc.OPT_CASES_HARD])
def test_oracle_matches_hard_golden(case):
    """Three contigs, gapped reads, reads over contig ends / the strand boundary / the text ends: the records the
    reference left for them carry band retries (w doubled) and empty left windows, so those paths are pinned."""
    name, kw = case
    ref, ref_string = cc.hard_genome()
    flat, off, ln = ol.pack_reads(cc.hard_reads(ref))
    chains, seeds = HARD[f"{name}_chains"], HARD[f"{name}_seeds"]
    o = ol.mem_opt(**kw)
    got = ol.oracle_chain2aln(o, int(HARD["l_pac"][0]), HARD["contig_off"], HARD["contig_len"], ref_string, flat, off,
                              ln, chains, seeds)
    g = HARD[f"{name}_regs"]
    assert np.array_equal(cc.reg_matrix(got), g)
    assert np.array_equal(got["frac_rep"], HARD[f"{name}_frac_rep"])
    W = cc.REG_FIELDS.index("w")
    live = g[:, 2] != -1
    if o["w"] <= 12:
        assert (g[live, W] == 2 * o["w"]).sum() > 10                  # the band-doubling retry ran (bwamem.cpp:3248)
    assert len(set(g[live, cc.REG_FIELDS.index("rid")])) == 3          # regions on every contig
    l_pac = int(HARD["l_pac"][0])
    assert np.any((seeds["qbeg"] > 0) & ((seeds["rbeg"] == 0) | (seeds["rbeg"] == l_pac)))   # left job with len1 == 0


def test_oracle_matches_reference_lib(tmp_path):
    if ol.ref_api("plain", "sse41") is None or ol.ref_binary("plain", "sse41") is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    ref, ref_string = cc.genome()
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta(prefix + ".fa", ref)
    ol.ref_index(prefix, prefix + ".fa")
    rm = ol.RefMem(prefix)
    reads, _, _ = synth.simulate_reads(ref, 1500, seed=4, err=0.05, read_len=250)
    flat, off, ln = ol.pack_reads(reads)
    o = ol.mem_opt()
    chains, seeds, regs, _ = rm.core(o, flat, off, ln)
    got = ol.oracle_chain2aln(o, rm.l_pac, rm.contig_off, rm.contig_len, ref_string, flat, off, ln, chains, seeds)
    assert np.array_equal(cc.reg_matrix(got), cc.reg_matrix(regs))
    assert (regs["w"] == 200).sum() > 0                                # the doubled band was needed somewhere
