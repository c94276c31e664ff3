"""GPU parity tests (through the C ABI): chaining on CUDA vs the chains the reference built, and the
whole device pipeline reads -> SMEMs -> SA -> chains -> alignment regions vs the reference's regions."""
import os

import numpy as np
import pytest

import c2a_cases as cc
import fmi_cases as fc
import oracle_lib as ol
from bwa_mem_scale_b200 import (Extender, FMI_search, REG_DTYPE, chain_opt, ext_opt, fmindex, mem_core_device,
                                seed_opt, synth)
from test_chain_hostsim import seeds_of

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_golden.npz"))


def opts(o):
    return (seed_opt(o["min_seed_len"], o["split_factor"], o["split_width"], o["max_occ"], o["max_mem_intv"]),
            chain_opt(o["w"], o["max_chain_gap"], o["max_occ"], o["min_seed_len"], o["min_chain_weight"],
                      o["max_chain_extend"], o["mask_level"], o["drop_ratio"]),
            ext_opt(*[o[k] for k in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5", "pen_clip3", "w", "zdrop")]))


@pytest.fixture(scope="module")
def small(gpu_ctx):
    ref, idx = fc.small_index()
    ref_string = fmindex.text_with_revcomp(ref)
    ext = Extender(ref_string, int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)
    return ref, idx, ext


@pytest.mark.parametrize("case", cc.OPT_CASES, ids=[c[0] for c in cc.OPT_CASES])
def test_gpu_chaining_matches_golden(small, case):
    name, kw = case
    ref, idx, ext = small
    flat, off, ln = ol.pack_reads(cc.golden_reads(ref))
    o = ol.mem_opt(**kw)
    sm, co = seeds_of(idx, o, flat, off, ln)
    chains, seeds = ext.mem_chain_seeds(opts(o)[1], sm, co, ln)
    assert np.array_equal(chains, GOLD[f"{name}_chains"])
    assert np.array_equal(seeds, GOLD[f"{name}_seeds"])


def run_device_pipeline(ctx, idx, ext, o, reads):
    fmi = FMI_search(idx, ctx=ctx)
    flat, off, ln = ol.pack_reads(reads)
    d_reads, d_off, d_len = ctx.upload(flat), ctx.upload(off), ctx.upload(ln)
    so, co, eo = opts(o)
    sres, cres, ares = mem_core_device(fmi, ext, so, co, eo, d_reads, d_off, d_len, len(ln), int(ln.max()))
    regs = np.zeros(ares.n_regs, REG_DTYPE)
    if ares.n_regs:
        ctx.d2h(regs, ares.d_regs)
    chains = np.zeros(cres.n_chains, ol.CHAIN_DTYPE)
    if cres.n_chains:
        ctx.d2h(chains, cres.d_chains)
    for p in (d_reads, d_off, d_len):
        ctx.free(p)
    return chains, regs, ares


@pytest.mark.parametrize("case", cc.OPT_CASES, ids=[c[0] for c in cc.OPT_CASES])
def test_gpu_reads_to_regions_matches_golden(small, gpu_ctx, case):
    """Nothing but reads go in: seeding, SA, chaining, extension and the filters all run on the device."""
    name, kw = case
    ref, idx, ext = small
    chains, regs, ares = run_device_pipeline(gpu_ctx, idx, ext, ol.mem_opt(**kw), cc.golden_reads(ref))
    assert np.array_equal(chains, GOLD[f"{name}_chains"])
    assert np.array_equal(cc.reg_matrix(regs), GOLD[f"{name}_regs"])
    assert ares.n_left > 0 and ares.n_right > 0 and ares.n_pairs8 + ares.n_pairs16 + ares.n_pairs_scalar == ares.n_left + ares.n_right


HARD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_hard_golden.npz"))


@pytest.mark.parametrize("case", cc.OPT_CASES_HARD, ids=[c[0] for c in cc.OPT_CASES_HARD])
def test_gpu_reads_to_regions_hard_golden(small, gpu_ctx, case):
    """Device pipeline on the three-contig genome with gapped / boundary reads: chains and regions equal the
    reference's, and the band-doubling retry path really ran (counted by the launcher)."""
    from test_chain_hostsim import hard_index
    name, kw = case
    ref, idx = hard_index()
    ext = Extender(fmindex.text_with_revcomp(ref), int(HARD["l_pac"][0]), HARD["contig_off"], HARD["contig_len"], ctx=gpu_ctx)
    o = ol.mem_opt(**kw)
    chains, regs, ares = run_device_pipeline(gpu_ctx, idx, ext, o, cc.hard_reads(ref))
    assert np.array_equal(chains, HARD[f"{name}_chains"])
    assert np.array_equal(cc.reg_matrix(regs), HARD[f"{name}_regs"])
    if o["w"] <= 12:
        assert ares.n_retry_left + ares.n_retry_right > 10
        assert (regs["w"][regs["qb"] != -1] == 2 * o["w"]).sum() > 10
    assert ares.n_pairs8 > 0 and ares.n_pairs16 > 0
    rs0 = fmindex.text_with_revcomp(fc.small_index()[0])   # restore the module fixture's replica
    Extender(rs0, int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)


def test_gpu_reads_to_regions_vs_reference_live(gpu_ctx, tmp_path):
    """Config-1 shape (1 Mbp random reference, 10 k x 151 bp, 1 % errors) against the reference run live."""
    if ol.ref_api("plain", "sse41") is None or ol.ref_binary("plain", "sse41") is None:
        pytest.skip("oracle/_ref not present")
    ref = synth.random_reference(1_000_000, 42)
    idx = fmindex.build_fm_index(ref)
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta(prefix + ".fa", ref)
    ol.ref_index(prefix, prefix + ".fa")
    rm = ol.RefMem(prefix)
    ext = Extender(fmindex.text_with_revcomp(ref), rm.l_pac, rm.contig_off, rm.contig_len, ctx=gpu_ctx)
    reads, _, _ = synth.simulate_reads(ref, 10000, seed=1)
    flat, off, ln = ol.pack_reads(reads)
    o = ol.mem_opt()
    want_c, want_s, want_r, _ = rm.core(o, flat, off, ln)
    chains, regs, ares = run_device_pipeline(gpu_ctx, idx, ext, o, reads)
    assert np.array_equal(chains, want_c)
    assert np.array_equal(cc.reg_matrix(regs), cc.reg_matrix(want_r))
    rs0 = fmindex.text_with_revcomp(fc.small_index()[0])   # restore the module fixture's replica
    Extender(rs0, int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)
