"""GPU parity tests (through the C ABI): CUDA ERT seeding, and reads -> regions with ERT seeding in front, against
records produced by the unmodified reference (tests/golden/make_golden_ert.py) — and, when oracle/_ref and a
reference-built index are present on the box (scripts/ert_box_check.py builds one), live against the reference."""
import ctypes as C
import os

import numpy as np
import pytest

import c2a_cases as cc
import ert_cases as ec
import oracle_lib as ol
from bwa_mem_scale_b200 import (BwaGpuError, Context, ErtIndex, ErtSeeding, Extender, REG_DTYPE, chain_opt, ext_opt,
                                fmindex, mem_core_ert_device, pack_reads, seed_opt)

pytestmark = pytest.mark.gpu
GOLD = np.load(ec.GOLDEN)


@pytest.fixture(scope="module")
def tiny_ert(gpu_ctx):
    ref = ec.tiny_genome()
    ext = Extender(fmindex.text_with_revcomp(ref), len(ref), GOLD["contig_off"], GOLD["contig_len"], ctx=gpu_ctx)
    ert = ErtSeeding(ErtIndex(GOLD["mlt_table"], sparse_idx=GOLD["kmer_idx"], sparse_val=GOLD["kmer_val"]), gpu_ctx)
    flat, off, ln = pack_reads(ec.golden_reads(ref))
    yield ref, ext, ert, flat, off, ln
    gpu_ctx.check(gpu_ctx.lib.bwagpu_index_unload_ert(gpu_ctx.h), "bwagpu_index_unload_ert")


def sopt_of(kw, max_occ=500):
    o = ol.mem_opt(**kw)
    return seed_opt(o["min_seed_len"], o["split_factor"], o["split_width"], max_occ, o["max_mem_intv"])


@pytest.mark.parametrize("case", ec.OPT_CASES, ids=[c[0] for c in ec.OPT_CASES])
def test_gpu_ert_seeding_matches_reference_records(tiny_ert, case):
    name, kw = case
    ref, ext, ert, flat, off, ln = tiny_ert
    mems, hits, mem_off = ert.seed(sopt_of(kw), flat, off, ln)
    assert np.array_equal(ec.mem_matrix(mems), ec.mem_matrix(GOLD[f"{name}_mems"]))
    assert np.array_equal(hits, GOLD[f"{name}_hits"])
    assert np.array_equal(mem_off, GOLD[f"{name}_mem_off"])


def test_gpu_ert_skip_empty_and_errors(tiny_ert, gpu_ctx):
    ref, ext, ert, flat, off, ln = tiny_ert
    opt = sopt_of({})
    skip = np.zeros(len(ln), np.uint8)
    skip[::3] = 1
    mems, hits, mem_off = ert.seed(opt, flat, off, ln, skip)
    gm, gh, go = GOLD["default_mems"], GOLD["default_hits"], GOLD["default_mem_off"]
    keep = skip[gm["rid"]] == 0
    assert np.array_equal(ec.mem_matrix(mems)[:, :4], ec.mem_matrix(gm[keep])[:, :4])
    want_hits = np.concatenate([gh[m["hit_off"]: m["hit_off"] + m["hitcount"]] for m in gm[keep]])
    assert np.array_equal(hits, want_hits)
    assert np.all(np.diff(mem_off)[skip == 1] == 0)
    m0, h0, o0 = ert.seed(opt, flat[:0], off[:0], ln[:0])
    assert len(m0) == 0 and len(h0) == 0 and list(o0) == [0]
    long_read = np.zeros(152, np.uint8)
    f2, o2, l2 = pack_reads([long_read])
    with pytest.raises(BwaGpuError):
        ert.seed(opt, f2, o2, l2)                       # longer than READ_LEN: the reference asserts
    with pytest.raises(BwaGpuError):
        ert.seed(seed_opt(min_seed_len=15), flat, off, ln)   # below k-mer + x-mer: the reference asserts


@pytest.mark.parametrize("case", ec.OPT_CASES[:2], ids=[c[0] for c in ec.OPT_CASES[:2]])
def test_gpu_ert_pipeline_matches_reference_chains_and_regions(tiny_ert, gpu_ctx, case):
    """reads -> ERT MEMs -> chains -> banded SW -> regions on the device vs mem_kernel1_core_ert + V2."""
    name, kw = case
    ref, ext, ert, flat, off, ln = tiny_ert
    o = ol.mem_opt(**kw)
    d_reads, d_off, d_len = gpu_ctx.upload(flat), gpu_ctx.upload(off), gpu_ctx.upload(ln)
    sres, cres, ares = mem_core_ert_device(ert, ext, sopt_of(kw), chain_opt(**{k: o[k] for k in (
        "w", "max_chain_gap", "max_occ", "min_seed_len", "min_chain_weight", "max_chain_extend", "mask_level",
        "drop_ratio")}), ext_opt(**{k: o[k] for k in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5",
                                                      "pen_clip3", "w", "zdrop")}),
        d_reads, d_off, d_len, len(ln), int(ln.max()))
    chains, seeds = np.zeros(cres.n_chains, ol.CHAIN_DTYPE), np.zeros(cres.n_seeds, ol.SEED_DTYPE)
    regs = np.zeros(ares.n_regs, REG_DTYPE)
    gpu_ctx.d2h(chains, cres.d_chains); gpu_ctx.d2h(seeds, cres.d_seeds); gpu_ctx.d2h(regs, ares.d_regs)
    assert np.array_equal(chains, GOLD[f"{name}_chains"])
    assert np.array_equal(seeds, GOLD[f"{name}_seeds"])
    assert np.array_equal(cc.reg_matrix(regs), GOLD[f"{name}_regs"])
    assert sres.n_hits == int(GOLD[f"{name}_mems"]["hitcount"].sum()) and sres.n_index_bytes > 0
    for d in (d_reads, d_off, d_len):
        gpu_ctx.free(d)


def test_gpu_ert_requires_index(gpu_ctx):
    ctx = Context(0)
    try:
        flat, off, ln = pack_reads([np.zeros(30, np.uint8)])
        n = C.c_int64(0)
        rc = ctx.lib.bwagpu_seed_ert(ctx.h, C.byref(seed_opt()), flat.ctypes.data, off.ctypes.data, ln.ctypes.data, None,
                                     1, None, 0, None, 0, None, C.byref(n), C.byref(n))
        assert rc == -3 and b"no ERT index" in ctx.lib.bwagpu_last_error(ctx.h)
    finally:
        ctx.close()


def test_gpu_kmer_table_completion_from_fm_index(tiny_ert, gpu_ctx):
    """Entries of absent k-mers recomputed on the device from the FM-index == the reference-built table (the bits
    seeding reads: type, LEP mask, hit count, pointer width); entries of occurring k-mers stay as loaded."""
    from bwa_mem_scale_b200 import FMI_search
    ref, ext, ert, flat, off, ln = tiny_ert
    FMI_search(fmindex.build_fm_index(ref, device="cpu"), ctx=gpu_ctx)
    idx, val = GOLD["kmer_idx"], GOLD["kmer_val"]
    rng = np.random.default_rng(5)
    extra = rng.integers(0, 1 << 30, 20000).astype(np.uint32)
    assert np.array_equal(ert.kmer_entries(idx), val)
    assert not np.any(ert.kmer_entries(extra[~np.isin(extra, idx)]))          # sparse load: everything else is 0
    ert.complete_from_fm()
    got = ert.kmer_entries(idx)
    present = (val & 3) != 0
    assert np.array_equal(got[present], val[present])
    assert np.array_equal(got[~present], val[~present] & np.uint64(0xFFFFFF))
    mems, hits, mem_off = ert.seed(sopt_of({}), flat, off, ln)
    assert np.array_equal(ec.mem_matrix(mems), ec.mem_matrix(GOLD["default_mems"]))
    assert np.array_equal(hits, GOLD["default_hits"])


BENCH_DATA = os.path.join(ec.HERE, "golden", "ert_small_genome.npz")


def test_gpu_ert_live_against_reference_at_scale(gpu_ctx):
    """60 k fresh reads on the 300 kbp reference-built index (k-mer table completed on the device): MEMs and hits vs the
    unmodified reference's get_seeds / reseed / last, chains and regions vs its mem_kernel1_core_ert + V2."""
    import tempfile
    from bwa_mem_scale_b200 import FMI_search, synth
    assert ol.ref_binary() is not None, "oracle/_ref missing: run __graft_entry__.build() where /root/reference exists"
    d = np.load(BENCH_DATA)
    ref = np.ascontiguousarray(d["ref"])
    ref_string = fmindex.text_with_revcomp(ref)
    idx = fmindex.build_fm_index(ref, device="cpu")
    FMI_search(idx, ctx=gpu_ctx)
    ext = Extender(ref_string, len(ref), [0], [len(ref)], ctx=gpu_ctx)
    ert = ErtSeeding(ErtIndex(d["mlt_table"], sparse_idx=d["kmer_idx"], sparse_val=d["kmer_val"]), gpu_ctx)
    ert.complete_from_fm()
    dense, mlt = ert.dense_table(), np.ascontiguousarray(d["mlt_table"])
    reads, _, _ = synth.simulate_reads(ref, 60000, seed=4242, err=0.015)
    rng = np.random.default_rng(8)
    extra = ec._sim(rng, ref, 1500, 151, 0.02, 0.01) + [ref[1000 + k:1151 + k].copy() for k in range(0, 300, 7)]
    flat, off, ln = pack_reads(list(reads) + extra)
    tap = ol.RefErt((dense, mlt, ref_string))
    want = tap.seed(flat, off, ln)
    got = ert.seed(sopt_of({}), flat, off, ln)
    assert np.array_equal(ec.mem_matrix(got[0]), ec.mem_matrix(want[0]))
    assert np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
    sub = list(reads[:15000]) + extra
    flat, off, ln = pack_reads(sub)
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_index_files(prefix, ref, idx)
        rm = ol.RefMem(prefix)
        rm.use_ert_tables(dense, mlt)
        chains, seeds, regs, _ = rm.core(ol.mem_opt(), flat, off, ln)
    d_reads, d_off, d_len = gpu_ctx.upload(flat), gpu_ctx.upload(off), gpu_ctx.upload(ln)
    sres, cres, ares = mem_core_ert_device(ert, ext, sopt_of({}), chain_opt(), ext_opt(), d_reads, d_off, d_len, len(ln), int(ln.max()))
    g_chains, g_seeds = np.zeros(cres.n_chains, ol.CHAIN_DTYPE), np.zeros(cres.n_seeds, ol.SEED_DTYPE)
    g_regs = np.zeros(ares.n_regs, REG_DTYPE)
    gpu_ctx.d2h(g_chains, cres.d_chains); gpu_ctx.d2h(g_seeds, cres.d_seeds); gpu_ctx.d2h(g_regs, ares.d_regs)
    assert np.array_equal(g_chains, chains) and np.array_equal(g_seeds, seeds)
    assert np.array_equal(cc.reg_matrix(g_regs), cc.reg_matrix(regs))
    for p in (d_reads, d_off, d_len):
        gpu_ctx.free(p)
    gpu_ctx.check(gpu_ctx.lib.bwagpu_index_unload_ert(gpu_ctx.h), "bwagpu_index_unload_ert")
