"""CPU tests: the per-thread chaining code of the CUDA kernels (csrc/chain_core.cuh: B-tree and
introsort replay, test_and_merge, mem_chain_flt), run on the host, against the chains the reference
itself built (golden records; live against oracle/_ref when present)."""
import ctypes as C
import os

import numpy as np
import pytest

import c2a_cases as cc
import fmi_cases as fc
import hostsim_lib
import oracle_lib as ol
from bwa_mem_scale_b200.lib import ChainOpt

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_golden.npz"))


def chain_opt_of(o):
    return ChainOpt(o["w"], o["max_chain_gap"], o["max_occ"], o["min_seed_len"], o["min_chain_weight"],
                    o["max_chain_extend"], o["mask_level"], o["drop_ratio"])


def hostsim_chains(o, l_pac, contig_off, contig_len, sm, co, ln):
    lib = hostsim_lib.load()
    oc, os_ = np.zeros(len(co) + 1, ol.CHAIN_DTYPE), np.zeros(len(co) + 1, ol.SEED_DTYPE)
    ns = C.c_int64(0)
    copt = chain_opt_of(o)
    contig_off, contig_len = np.ascontiguousarray(contig_off, np.int64), np.ascontiguousarray(contig_len, np.int32)
    nc = lib.hostsim_chain_seeds(C.byref(copt), C.c_int64(int(l_pac)), C.c_int(len(contig_off)),
                                 C.c_void_p(contig_off.ctypes.data), C.c_void_p(contig_len.ctypes.data),
                                 C.c_void_p(sm.ctypes.data), C.c_int64(len(sm)), C.c_void_p(co.ctypes.data),
                                 C.c_void_p(ln.ctypes.data), C.c_int32(len(ln)), C.c_void_p(oc.ctypes.data),
                                 C.c_void_p(os_.ctypes.data), C.byref(ns))
    return oc[:nc], os_[:ns.value]


def seeds_of(idx, o, flat, off, ln):
    so = ol.default_seed_opt(min_seed_len=o["min_seed_len"], split_factor=o["split_factor"],
                             split_width=o["split_width"], max_mem_intv=o["max_mem_intv"])
    sm, _ = ol.oracle_collect_smem(idx, so, flat, off, ln)
    co, _ = ol.oracle_sa_lookup(idx, sm, o["max_occ"])
    return sm, co


@pytest.mark.parametrize("case", cc.OPT_CASES, ids=[c[0] for c in cc.OPT_CASES])
def test_chaining_matches_golden(case):
    name, kw = case
    ref, idx = fc.small_index()
    flat, off, ln = ol.pack_reads(cc.golden_reads(ref))
    o = ol.mem_opt(**kw)
    sm, co = seeds_of(idx, o, flat, off, ln)
    chains, seeds = hostsim_chains(o, GOLD["l_pac"][0], GOLD["contig_off"], GOLD["contig_len"], sm, co, ln)
    assert np.array_equal(chains, GOLD[f"{name}_chains"])
    assert np.array_equal(seeds, GOLD[f"{name}_seeds"])
    # the replayed structures were really exercised: multi-level tree and the quicksort part of introsort
    assert np.bincount(chains["seqid"]).max() > 40


HARD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_hard_golden.npz"))
_hard = {}


def hard_index():
    if not _hard:
        from bwa_mem_scale_b200 import fmindex
        _hard["ref"] = cc.hard_genome()[0]
        _hard["idx"] = fmindex.build_fm_index(_hard["ref"], device="cpu")
    return _hard["ref"], _hard["idx"]


@pytest.mark.parametrize("case", cc.OPT_CASES_HARD, ids=[c[0] for c in cc.OPT_CASES_HARD])
def test_chaining_matches_hard_golden(case):
    """Multi-contig chaining: seeds that bridge two contigs or the strand boundary are dropped (bns_intv2rid < 0,
    reference src/bwamem.cpp:872-876), chains never span contigs."""
    name, kw = case
    ref, idx = hard_index()
    flat, off, ln = ol.pack_reads(cc.hard_reads(ref))
    o = ol.mem_opt(**kw)
    sm, co = seeds_of(idx, o, flat, off, ln)
    chains, seeds = hostsim_chains(o, HARD["l_pac"][0], HARD["contig_off"], HARD["contig_len"], sm, co, ln)
    assert np.array_equal(chains, HARD[f"{name}_chains"])
    assert np.array_equal(seeds, HARD[f"{name}_seeds"])
    assert set(chains["rid"]) == {0, 1, 2}


def test_chaining_matches_reference_live(tmp_path):
    if ol.ref_api("plain", "sse41") is None or ol.ref_binary("plain", "sse41") is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    from bwa_mem_scale_b200 import fmindex, synth
    ref, idx = fc.small_index()
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta(prefix + ".fa", ref)
    ol.ref_index(prefix, prefix + ".fa")
    rm = ol.RefMem(prefix)
    for reads, kw in ((fc.small_reads(ref), dict(max_occ=50, min_seed_len=10, split_width=50, max_mem_intv=100)),
                      (fc.small_reads(ref), dict(w=10, mask_level=0.3, drop_ratio=0.8, max_chain_extend=3)),
                      (list(synth.simulate_reads(ref, 1200, seed=4, err=0.05, read_len=250)[0]), {})):
        flat, off, ln = ol.pack_reads(reads)
        o = ol.mem_opt(**kw)
        want_c, want_s, _, _ = rm.core(o, flat, off, ln)
        sm, co = seeds_of(idx, o, flat, off, ln)
        chains, seeds = hostsim_chains(o, rm.l_pac, rm.contig_off, rm.contig_len, sm, co, ln)
        assert np.array_equal(chains, want_c) and np.array_equal(seeds, want_s), kw
