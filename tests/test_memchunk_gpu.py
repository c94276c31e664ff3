"""GPU parity tests of the worker-level entry (bwagpu_mem_process_chunk / _device / bwagpu_mem_fetch, C ABI section W):
one call per chunk must leave, per read, exactly the regions the reference's worker_bwt + worker_aln leave after the
compaction loop (reference src/bwamem.cpp:1449-1460) — checked against the reference-generated golden records."""
import os

import numpy as np
import pytest

import c2a_cases as cc
import fmi_cases as fc
import oracle_lib as ol
from bwa_mem_scale_b200 import (Extender, FMI_search, MemChunk, PerfectMatch, PerfectTable, chain_opt, ext_opt, fmindex,
                                mem_opt, seed_opt)
from test_chain_hostsim import hard_index

pytestmark = pytest.mark.gpu
HARD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_hard_golden.npz"))
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "c2a_golden.npz"))


def opts(o, **kw):
    return mem_opt(seed_opt(o["min_seed_len"], o["split_factor"], o["split_width"], o["max_occ"], o["max_mem_intv"]),
                   chain_opt(o["w"], o["max_chain_gap"], o["max_occ"], o["min_seed_len"], o["min_chain_weight"],
                             o["max_chain_extend"], o["mask_level"], o["drop_ratio"]),
                   ext_opt(*[o[k] for k in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5", "pen_clip3", "w", "zdrop")]),
                   **kw)


def expected(gold, name, nseq):
    """Golden raw regions -> per-read compacted lists."""
    g, chains = gold[f"{name}_regs"], gold[f"{name}_chains"]
    Q, E, CH = (cc.REG_FIELDS.index(f) for f in ("qb", "qe", "chain"))
    keep = g[:, E] > g[:, Q]
    read_of = chains["seqid"][g[:, CH]]
    off = np.zeros(nseq + 1, np.int64)
    np.add.at(off, read_of[keep] + 1, 1)
    return g[keep], np.cumsum(off)


@pytest.mark.parametrize("case", cc.OPT_CASES_HARD, ids=[c[0] for c in cc.OPT_CASES_HARD])
def test_chunk_entry_matches_hard_golden(gpu_ctx, case):
    name, kw = case
    ref, idx = hard_index()
    FMI_search(idx, ctx=gpu_ctx)
    Extender(fmindex.text_with_revcomp(ref), int(HARD["l_pac"][0]), HARD["contig_off"], HARD["contig_len"], ctx=gpu_ctx)
    flat, off, ln = ol.pack_reads(cc.hard_reads(ref))
    mc = MemChunk(gpu_ctx)
    # the golden records come from the tap's mem_kernel1_core calls, 512 reads at a time like worker_bwt; no 512-read slice
    # of this read set totals <= 1 SMEM, so the single-SMEM rule never fires and one call-wide batch gives the same chains
    # (the rule itself is covered by test_chunk_entry_batch_rule_and_empty)
    res = mc.process(opts(ol.mem_opt(**kw), batch_size=0), flat, off, ln)
    regs, reg_off, emf_ret, perfect = mc.fetch(res, len(ln))
    want, want_off = expected(HARD, name, len(ln))
    assert np.array_equal(reg_off, want_off)
    got = cc.reg_matrix(regs)
    CH = cc.REG_FIELDS.index("chain")
    assert np.array_equal(np.delete(got, CH, 1), np.delete(want, CH, 1))
    assert res.n_regs == len(want) and res.n_regs_raw == len(HARD[f"{name}_regs"]) and res.n_chains == len(HARD[f"{name}_chains"])
    assert res.sw_cells > 0 and res.n_sw_left + res.n_sw_right == res.n_pairs8 + res.n_pairs16 + res.n_pairs_scalar
    assert not emf_ret.any() and not perfect["flags"].any()
    # device-resident form: same bytes
    d_reads, d_off, d_len = gpu_ctx.upload(flat), gpu_ctx.upload(off), gpu_ctx.upload(ln)
    res2 = mc.process_device(opts(ol.mem_opt(**kw), batch_size=0), d_reads, d_off, d_len, len(ln), int(ln.max()))
    regs2, reg_off2, _, _ = mc.fetch(res2, len(ln))
    assert np.array_equal(regs2, regs) and np.array_equal(reg_off2, reg_off) and res2.sw_cells == res.sw_cells
    for p in (d_reads, d_off, d_len):
        gpu_ctx.free(p)


def test_chunk_entry_batch_rule_and_empty(gpu_ctx):
    """mem_chain_seeds' loop bound: a BATCH_SIZE slice whose SMEMs total one yields no chains (src/bwamem.cpp:819)."""
    ref, idx = hard_index()
    FMI_search(idx, ctx=gpu_ctx)
    Extender(fmindex.text_with_revcomp(ref), int(HARD["l_pac"][0]), HARD["contig_off"], HARD["contig_len"], ctx=gpu_ctx)
    reads = cc.hard_reads(ref)[:512] + [ref[30000:30019].copy()]
    flat, off, ln = ol.pack_reads(reads)
    mc = MemChunk(gpu_ctx)
    o = ol.mem_opt()
    whole = mc.fetch(mc.process(opts(o, batch_size=0), flat, off, ln), len(ln))
    sliced = mc.fetch(mc.process(opts(o, batch_size=512), flat, off, ln), len(ln))
    assert whole[1][-1] - whole[1][-2] >= 1 and sliced[1][-1] - sliced[1][-2] == 0
    assert np.array_equal(whole[1][:-1], sliced[1][:-1]) and np.array_equal(whole[0][:sliced[1][-1]], sliced[0])
    # empty chunk, and a chunk of reads that cannot seed
    res = mc.process(opts(o), flat[:0], off[:0], ln[:0])
    assert res.n_regs == 0
    f2, o2, l2 = ol.pack_reads([np.full(40, 4, np.uint8), ref[5:15].copy(), np.zeros(0, np.uint8)])
    r2 = mc.fetch(mc.process(opts(o), f2, o2, l2), 3)
    assert len(r2[0]) == 0 and not r2[1].any()


def test_chunk_entry_with_emf(gpu_ctx):
    """EMF in front: answered reads carry perfect[] / ret and no regions; the others are unchanged."""
    ref, idx = fc.small_index()
    fmi = FMI_search(idx, ctx=gpu_ctx)
    Extender(fmindex.text_with_revcomp(ref), int(GOLD["l_pac"][0]), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)
    from bwa_mem_scale_b200 import synth
    reads = list(synth.simulate_reads(ref, 1500, seed=9, err=0.004)[0]) + cc.golden_reads(ref)[:300]
    flat, off, ln = ol.pack_reads(reads)
    table = PerfectTable.build(ref, 151, 1.1)
    pm = PerfectMatch(table, gpu_ctx)
    mc = MemChunk(gpu_ctx)
    o = ol.mem_opt()
    plain = mc.fetch(mc.process(opts(o, batch_size=0), flat, off, ln), len(ln))
    res = mc.process(opts(o, batch_size=0, use_emf=True), flat, off, ln)
    regs, reg_off, ret, perfect = mc.fetch(res, len(ln))
    pm.unload()
    want_ret, want_perf = ol.oracle_emf_probe(table, fmindex.text_with_revcomp(ref), flat, off, ln)
    assert np.array_equal(ret, want_ret)
    assert np.array_equal(perfect["flags"], want_perf[:, 0]) and np.array_equal(perfect["location"], want_perf[:, 1])
    hit = (ret == 3) | (ret == 4)
    assert 600 < hit.sum() == res.n_emf_hits
    n_per = np.diff(reg_off)
    assert not n_per[hit].any()
    p_per = np.diff(plain[1])
    assert np.array_equal(n_per[~hit], p_per[~hit])
    keep = np.repeat(~hit, p_per)
    a, b = plain[0][keep].copy(), regs.copy()
    a["chain"] = 0; b["chain"] = 0
    assert np.array_equal(a, b)


def test_gpu_worker_contexts_share_one_index(gpu_ctx):
    """bwagpu_ctx_share_index: two contexts on one GPU, one index replica, one host thread each; same results as
    one context, a borrower can not touch the index, and the owner still works after the borrower is gone."""
    from concurrent.futures import ThreadPoolExecutor
    from bwa_mem_scale_b200 import BwaGpuError
    ref, idx = hard_index()
    Extender(fmindex.text_with_revcomp(ref), int(HARD["l_pac"][0]), HARD["contig_off"], HARD["contig_len"], ctx=gpu_ctx)
    fmi = FMI_search(idx, ctx=gpu_ctx)
    fmi.build_fma()
    flat, off, ln = ol.pack_reads(cc.hard_reads(ref))
    opt = opts(ol.mem_opt(), batch_size=0)
    whole = MemChunk(gpu_ctx)
    want = whole.fetch(whole.process(opt, flat, off, ln), len(ln))
    w = gpu_ctx.worker()
    try:
        with pytest.raises(BwaGpuError):
            FMI_search(idx, ctx=w)                                        # a borrower loads nothing
        half = len(ln) // 2
        parts = [(0, half), (half, len(ln))]

        def run(k):
            c0, c1 = parts[k]
            mc = MemChunk([gpu_ctx, w][k])
            out = []
            for _ in range(3):                                             # overlapping calls on the two streams
                lo = int(off[c0])
                hi = int(off[c1 - 1] + ln[c1 - 1])
                r = mc.process(opt, flat[lo:hi], off[c0:c1] - lo, ln[c0:c1])
                out = mc.fetch(r, c1 - c0)
            return out
        with ThreadPoolExecutor(2) as ex:
            a, b = list(ex.map(run, range(2)))
        CH = cc.REG_FIELDS.index("chain")                                  # chain ids count from 0 in every call
        strip = lambda regs: np.delete(cc.reg_matrix(regs), CH, 1)
        assert np.array_equal(strip(np.concatenate([a[0], b[0]])), strip(want[0]))
        assert np.array_equal(np.concatenate([a[1][:-1], b[1] + a[1][-1]]), want[1])
    finally:
        w.close()
    again = whole.fetch(whole.process(opt, flat, off, ln), len(ln))
    assert np.array_equal(again[0], want[0])
