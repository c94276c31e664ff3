"""GPU tests at bench scale through size-independent properties (the oracle would take minutes here):
the device pipeline reads -> regions on a 20 Mbp reference with 300 k reads, checked for
 * shard invariance: two half batches stitched by bwa_mem_scale_b200.shard == the whole batch
   (what the multi-GPU path relies on; a checksum of checksums over ranks),
 * idempotence: the same call twice gives the same bytes,
 * structural invariants of SMEMs, chains and regions,
 * a 3 000-read sample cut from the middle of the batch == the oracle's regions for those reads,
 * the EMF in front: reads it answers drop out, every other read's regions are unchanged."""
import hashlib

import numpy as np
import pytest

import c2a_cases as cc
import oracle_lib as ol
from bwa_mem_scale_b200 import (Extender, FMI_search, PerfectMatch, PerfectTable, REG_DTYPE, SMEM_DTYPE, chain_opt,
                                ext_opt, fmindex, mem_core_device, seed_opt, shard, synth)

pytestmark = pytest.mark.gpu
N_BP, N_READS = 20_000_000, 300_000


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.fixture(scope="module")
def world(gpu_ctx):
    ref = synth.repeat_reference(N_BP, 43, n_families=20, repeat_frac=0.10, divergence=0.02)
    idx = fmindex.build_fm_index(ref)
    reads, _, _ = synth.simulate_reads(ref, N_READS, seed=77, err=0.01)
    text = fmindex.text_with_revcomp(ref)
    fmi = FMI_search(idx, ctx=gpu_ctx)
    fmi.build_fma()
    ext = Extender(text, len(ref), [0], [len(ref)], ctx=gpu_ctx)
    yield ref, idx, text, reads, fmi, ext
    fmi.use_fma(False)


def run(ctx, fmi, ext, flat, off, ln, emf=None):
    d_reads, d_off, d_len = ctx.upload(flat), ctx.upload(off), ctx.upload(ln)
    d_emf = (ctx.malloc(4 * len(ln)), ctx.malloc(8 * len(ln)), ctx.malloc(len(ln))) if emf is not None else None
    sres, cres, ares = mem_core_device(fmi, ext, seed_opt(), chain_opt(), ext_opt(), d_reads, d_off, d_len, len(ln),
                                       int(ln.max()), emf, d_emf)
    sm = np.zeros(sres.n_smem, SMEM_DTYPE)
    chains, seeds = np.zeros(cres.n_chains, ol.CHAIN_DTYPE), np.zeros(cres.n_seeds, ol.SEED_DTYPE)
    regs = np.zeros(ares.n_regs, REG_DTYPE)
    for a, p in ((sm, sres.d_smems), (chains, cres.d_chains), (seeds, cres.d_seeds), (regs, ares.d_regs)):
        if len(a):
            ctx.d2h(a, p)
    skip = None
    if emf is not None:
        skip = np.zeros(len(ln), np.uint8)
        ctx.d2h(skip, d_emf[2])
    for p in (d_reads, d_off, d_len) + (d_emf or ()):
        ctx.free(p)
    return sm, chains, seeds, regs, skip


def test_scale_properties_and_shard_invariance(gpu_ctx, world):
    ref, idx, text, reads, fmi, ext = world
    flat, off, ln = ol.pack_reads(reads)
    sm, chains, seeds, regs, _ = run(gpu_ctx, fmi, ext, flat, off, ln)
    # idempotence
    sm2, chains2, seeds2, regs2, _ = run(gpu_ctx, fmi, ext, flat, off, ln)
    assert [digest(x) for x in (sm, chains, seeds, regs)] == [digest(x) for x in (sm2, chains2, seeds2, regs2)]
    # shard invariance (two ranks' worth of work, stitched on the host)
    parts = []
    for r in range(2):
        f_r, o_r, l_r, first = shard.shard_reads(flat, off, ln, 2, r)
        parts.append((first,) + run(gpu_ctx, fmi, ext, f_r, o_r, l_r)[:4])
    firsts = [p[0] for p in parts]
    assert digest(shard.merge_smems([p[1] for p in parts], firsts)) == digest(sm)
    c, s, g = shard.merge_regions([p[2] for p in parts], [p[3] for p in parts], [p[4] for p in parts], firsts)
    assert (digest(c), digest(s), digest(g)) == (digest(chains), digest(seeds), digest(regs))
    # structure: canonical SMEM order, spans inside the read, intervals inside the text
    key = (sm["rid"].astype(np.int64) << 32) | (sm["m"].astype(np.int64) << 16) | sm["n"]
    assert np.all(np.diff(key) >= 0) and np.all(sm["m"] <= sm["n"]) and np.all(sm["n"] < 151)
    assert np.all(sm["s"] >= 1) and np.all(sm["k"] + sm["s"] <= idx.ref_seq_len)
    # chains grouped by read, seed lists contiguous, every seed inside its read and inside the text
    assert np.all(np.diff(chains["seqid"]) >= 0)
    assert np.array_equal(chains["seed_off"], np.concatenate([[0], np.cumsum(chains["n"])[:-1]]))
    assert len(seeds) == int(chains["n"].sum()) == len(regs)
    assert np.all(seeds["qbeg"] >= 0) and np.all(seeds["qbeg"] + seeds["len"] <= 151)
    assert np.all(seeds["rbeg"] >= 0) and np.all(seeds["rbeg"] + seeds["len"] <= 2 * len(ref))
    # regions: purged ones carry qb == qe == -1; the others contain their seed and stay on one strand
    live = ~((regs["qb"] == -1) & (regs["qe"] == -1))
    r = regs[live]
    assert 0.3 < live.mean() < 0.9
    assert np.all((r["qb"] >= 0) & (r["qb"] < r["qe"]) & (r["qe"] <= 151) & (r["rb"] < r["re"]))
    assert np.all((r["re"] <= len(ref)) | (r["rb"] >= len(ref)))
    assert np.all(r["score"] >= r["seedlen0"]) and np.all(r["seedcov"] > 0) and np.all(r["rid"] == 0)
    best = np.zeros(len(ln), np.int32)
    np.maximum.at(best, chains["seqid"][r["chain"]], r["score"])
    assert (best >= 126).mean() > 0.95          # 1 % substitutions, -5 per mismatch: <= 5 mismatches covers 99 %

    # a sample from the middle of the batch against the oracle (batch composition must not matter)
    lo = N_READS // 2
    sub = reads[lo:lo + 3000]
    f3, o3, l3 = ol.pack_reads(sub)
    in_sub = (chains["seqid"] >= lo) & (chains["seqid"] < lo + 3000)
    c_sub = chains[in_sub].copy()
    s_lo, s_hi = int(c_sub["seed_off"][0]), int(c_sub["seed_off"][-1] + c_sub["n"][-1])
    c_sub["seqid"] -= lo
    c_sub["seed_off"] -= s_lo
    want = ol.oracle_chain2aln(ol.mem_opt(), len(ref), np.array([0], np.int64), np.array([len(ref)], np.int32), text,
                               f3, o3, l3, c_sub, seeds[s_lo:s_hi])
    got = regs[s_lo:s_hi].copy()
    got["chain"] -= int(np.nonzero(in_sub)[0][0])
    assert np.array_equal(cc.reg_matrix(got), cc.reg_matrix(want))
    want_sm, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(), f3, o3, l3)
    got_sm = sm[(sm["rid"] >= lo) & (sm["rid"] < lo + 3000)].copy()
    got_sm["rid"] -= lo
    import fmi_cases as fc
    assert np.array_equal(fc.smem_matrix(got_sm), fc.smem_matrix(want_sm))


def test_scale_emf_in_front(gpu_ctx, world):
    ref, idx, text, reads, fmi, ext = world
    flat, off, ln = ol.pack_reads(reads[:100000])
    table = PerfectTable.build(ref, 151, 1.1)
    pm = PerfectMatch(table, gpu_ctx)
    sm0, ch0, sd0, rg0, _ = run(gpu_ctx, fmi, ext, flat, off, ln)
    sm1, ch1, sd1, rg1, skip = run(gpu_ctx, fmi, ext, flat, off, ln, emf=pm)
    pm.unload()
    assert 0.15 < skip.mean() < 0.30                       # (1 - 0.01)^151 = 22 % of the reads are exact
    keep = ~skip.astype(bool)
    assert not np.any(skip[ch1["seqid"]]) and not np.any(skip[sm1["rid"]])
    assert digest(sm0[keep[sm0["rid"]]]) == digest(sm1)
    m0 = keep[ch0["seqid"]]
    a, b = ch0[m0].copy(), ch1.copy()
    a["seed_off"] = 0; b["seed_off"] = 0
    assert digest(a) == digest(b)
    r0 = rg0[m0[rg0["chain"]]].copy(); r1 = rg1.copy()
    r0["chain"] = 0; r1["chain"] = 0
    assert digest(r0) == digest(r1)
