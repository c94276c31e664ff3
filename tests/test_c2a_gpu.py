"""GPU parity tests (through the C ABI): chains -> alignment regions on CUDA vs golden reference records,
the oracle, and (when oracle/_ref is present) the reference's own code run live on larger inputs."""
import os

import numpy as np
import pytest

import c2a_cases as cc
import oracle_lib as ol
from bwa_mem_scale_b200 import Extender, ext_opt, fmindex, synth

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_golden.npz"))


def to_ext_opt(o):
    return ext_opt(*[o[k] for k in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5", "pen_clip3", "w", "zdrop")])


@pytest.fixture(scope="module")
def extender(gpu_ctx):
    ref, ref_string = cc.genome()
    return ref, ref_string, Extender(ref_string, int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)


@pytest.mark.parametrize("case", cc.OPT_CASES, ids=[c[0] for c in cc.OPT_CASES])
def test_gpu_chain2aln_matches_golden(extender, case):
    name, kw = case
    ref, ref_string, ext = extender
    flat, off, ln = ol.pack_reads(cc.golden_reads(ref))
    chains, seeds = GOLD[f"{name}_chains"], GOLD[f"{name}_seeds"]
    got = ext.mem_chain2aln_across_reads_V2(to_ext_opt(ol.mem_opt(**kw)), flat, off, ln, chains, seeds)
    assert np.array_equal(cc.reg_matrix(got), GOLD[f"{name}_regs"])
    assert np.array_equal(got["frac_rep"], GOLD[f"{name}_frac_rep"])


HARD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_hard_golden.npz"))


@pytest.mark.parametrize("case", cc.OPT_CASES_HARD, ids=[c[0] for c in cc.OPT_CASES_HARD])
def test_gpu_chain2aln_matches_hard_golden(gpu_ctx, extender, case):
    """Three contigs, gapped reads, boundary reads (reference records; band retries and empty left windows inside)."""
    name, kw = case
    ref, ref_string = cc.hard_genome()
    ext = Extender(ref_string, int(HARD["l_pac"][0]), HARD["contig_off"], HARD["contig_len"], ctx=gpu_ctx)
    flat, off, ln = ol.pack_reads(cc.hard_reads(ref))
    chains, seeds = HARD[f"{name}_chains"], HARD[f"{name}_seeds"]
    o = ol.mem_opt(**kw)
    got = ext.mem_chain2aln_across_reads_V2(to_ext_opt(o), flat, off, ln, chains, seeds)
    assert np.array_equal(cc.reg_matrix(got), HARD[f"{name}_regs"])
    assert np.array_equal(got["frac_rep"], HARD[f"{name}_frac_rep"])
    if o["w"] <= 12:
        assert (got["w"][got["qb"] != -1] == 2 * o["w"]).sum() > 10
    ref0, rs0 = cc.genome()     # restore the module fixture's reference replica on the shared context
    Extender(rs0, int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)


def test_gpu_chain2aln_empty_and_subset(extender):
    ref, ref_string, ext = extender
    flat, off, ln = ol.pack_reads(cc.golden_reads(ref))
    chains, seeds = GOLD["default_chains"], GOLD["default_seeds"]
    o = to_ext_opt(ol.mem_opt())
    assert len(ext.mem_chain2aln_across_reads_V2(o, flat, off, ln, chains[:0], seeds[:0])) == 0
    # a prefix of the chains (whole reads) gives the same regions for those reads
    cut = int(np.searchsorted(chains["seqid"], 100))
    ns = int(chains["seed_off"][cut])
    got = ext.mem_chain2aln_across_reads_V2(o, flat, off, ln, chains[:cut], seeds[:ns])
    assert np.array_equal(cc.reg_matrix(got), GOLD["default_regs"][:ns])


def test_gpu_chain2aln_vs_reference_live(gpu_ctx, tmp_path):
    """Reference run live (oracle/_ref travels to the GPU box): 5 k reads, 250 bp, 5 % errors."""
    if ol.ref_api("plain", "sse41") is None or ol.ref_binary("plain", "sse41") is None:
        pytest.skip("oracle/_ref not present")
    ref = synth.repeat_reference(2_000_000, 9, n_families=30, repeat_frac=0.2, divergence=0.03)
    prefix = str(tmp_path / "ref")
    contigs = (900_000, 600_000, 300_000, 200_000)
    fmindex.write_fasta_contigs(prefix + ".fa", ref, contigs)
    ol.ref_index(prefix, prefix + ".fa")
    rm = ol.RefMem(prefix)
    assert list(rm.contig_len) == list(contigs)
    ref_string = fmindex.text_with_revcomp(ref)
    ext = Extender(ref_string, rm.l_pac, rm.contig_off, rm.contig_len, ctx=gpu_ctx)
    ends = np.cumsum(contigs)
    pins = [int(e) - k for e in ends[:-1] for k in range(5, 150, 9)] + [len(ref) - k for k in range(5, 150, 9)]
    gapped = synth.gapped_reads(ref_string, 5000, seed=3, starts=pins)
    for reads, okw in ((synth.simulate_reads(ref, 5000, seed=1, read_len=151, err=0.01)[0], {}),
                       (synth.simulate_reads(ref, 5000, seed=2, read_len=250, err=0.05)[0], {}),
                       (gapped, {}), (gapped, dict(w=10, zdrop=20)), (gapped, dict(w=4))):
        flat, off, ln = ol.pack_reads(reads)
        o = ol.mem_opt(**okw)
        chains, seeds, regs, _ = rm.core(o, flat, off, ln)
        got = ext.mem_chain2aln_across_reads_V2(to_ext_opt(o), flat, off, ln, chains, seeds)
        assert np.array_equal(cc.reg_matrix(got), cc.reg_matrix(regs)), okw
        if okw:
            assert (regs["w"][regs["qb"] != -1] == 2 * o["w"]).sum() > 50      # band-doubling retries ran
    ref0, rs0 = cc.genome()     # restore the module fixture's reference replica on the shared context
    Extender(rs0, int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)


def test_gpu_pipeline_long_reads_in_place_windows(gpu_ctx):
    """600-base reads with indels on the three-contig genome: extension queries longer than the shared-memory SW kernel
    takes (len2 > 512) go through the generic kernel, and every job reads its query and its reference window IN PLACE
    (read buffer / resident text, backwards for left extensions) — regions must equal the oracle's for the same chains."""
    from bwa_mem_scale_b200 import FMI_search, REG_DTYPE, chain_opt, mem_core_device, pack_reads, seed_opt
    from test_chain_hostsim import hard_index
    ref, idx = hard_index()
    hard = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_hard_golden.npz"))
    text = fmindex.text_with_revcomp(ref)
    l_pac, coff, clen = int(hard["l_pac"][0]), hard["contig_off"], hard["contig_len"]
    fmi = FMI_search(idx, ctx=gpu_ctx)
    ext = Extender(text, l_pac, coff, clen, ctx=gpu_ctx)
    reads = synth.gapped_reads(ref, 600, seed=5, read_len=600, err=0.02) + synth.gapped_reads(ref, 200, seed=6, read_len=540, err=0.01)
    flat, off, ln = pack_reads(reads)
    d = [gpu_ctx.upload(a) for a in (flat, off, ln)]
    sres, cres, ares = mem_core_device(fmi, ext, seed_opt(), chain_opt(), ext_opt(), d[0], d[1], d[2], len(ln), int(ln.max()))
    chains, seeds = np.zeros(cres.n_chains, ol.CHAIN_DTYPE), np.zeros(cres.n_seeds, ol.SEED_DTYPE)
    regs = np.zeros(ares.n_regs, REG_DTYPE)
    gpu_ctx.d2h(chains, cres.d_chains); gpu_ctx.d2h(seeds, cres.d_seeds); gpu_ctx.d2h(regs, ares.d_regs)
    for p in d:
        gpu_ctx.free(p)
    want = ol.oracle_chain2aln(ol.mem_opt(), l_pac, coff, clen, text, flat, off, ln, chains, seeds)
    assert len(regs) > 1000 and np.array_equal(cc.reg_matrix(regs), cc.reg_matrix(want))
    assert ares.seq_bytes == 0                                        # nothing was materialised
    # some jobs had queries beyond the shared-memory kernel's reach
    s = seeds
    assert (s["qbeg"] > 512).any() or ((np.repeat(600, len(s)) - s["qbeg"] - s["len"]) > 512).any()
