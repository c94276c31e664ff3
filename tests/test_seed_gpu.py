"""GPU parity tests (through the C ABI): CUDA FM-index seeding vs the oracle / golden digests."""
import hashlib
import os

import numpy as np
import pytest

import fmi_cases as fc
import oracle_lib as ol
from bwa_mem_scale_b200 import FMI_search, fmindex, seed_opt, synth

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(fc.HERE, "golden", "fmi_golden.npz"))


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


@pytest.fixture(scope="module")
def small_fmi(gpu_ctx):
    ref, idx = fc.small_index()
    return ref, idx, FMI_search(idx, ctx=gpu_ctx)


@pytest.mark.parametrize("case", fc.OPT_CASES, ids=[c[0] for c in fc.OPT_CASES])
def test_gpu_seeding_matches_golden_and_oracle(small_fmi, case):
    name, kw, max_occ = case
    ref, idx, fmi = small_fmi
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    oo = ol.default_seed_opt(**kw)
    got = fmi.collect_smem(seed_opt(oo.min_seed_len, oo.split_factor, oo.split_width, max_occ, oo.max_mem_intv),
                           flat, off, ln)
    sa = fmi.get_sa_entries_prefetch(got, max_occ)
    assert list(GOLD[f"{name}_counts"]) == [len(got), len(sa)]
    assert np.array_equal(sha(fc.smem_matrix(got)), GOLD[f"{name}_smem_sha256"])
    assert np.array_equal(sha(sa), GOLD[f"{name}_sa_sha256"])
    want, _ = ol.oracle_collect_smem(idx, oo, flat, off, ln)
    assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(want))
    assert np.array_equal(sa, ol.oracle_sa_lookup(idx, want, max_occ)[0])


@pytest.mark.parametrize("cap", [2, 5])
def test_gpu_seeding_overflow_list(small_fmi, cap):
    """Reads that outgrow the common SMEM slab are re-run from a list with private slabs (several rounds when the
    truncated first pass under-reports): forced here with a tiny slab, results must not change."""
    name, kw, max_occ = fc.OPT_CASES[1]                      # short_seeds: up to dozens of SMEMs per read
    ref, idx, fmi = small_fmi
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    oo = ol.default_seed_opt(**kw)
    os.environ["BWAGPU_SEED_CAP"] = str(cap)
    try:
        got = fmi.collect_smem(seed_opt(oo.min_seed_len, oo.split_factor, oo.split_width, max_occ, oo.max_mem_intv), flat, off, ln)
    finally:
        os.environ.pop("BWAGPU_SEED_CAP")
    assert np.array_equal(sha(fc.smem_matrix(got)), GOLD[f"{name}_smem_sha256"])


def test_gpu_seeding_skip_empty_and_overflow(small_fmi, gpu_ctx):
    ref, idx, fmi = small_fmi
    reads = fc.small_reads(ref)[:300]
    flat, off, ln = ol.pack_reads(reads)
    opt = seed_opt()
    skip = np.zeros(len(reads), np.uint8)
    skip[::3] = 1
    got = fmi.collect_smem(opt, flat, off, ln, skip)
    want, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(), flat, off, ln, skip)
    assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(want))
    assert len(fmi.collect_smem(opt, flat[:0], off[:0], ln[:0])) == 0
    assert len(fmi.get_sa_entries_prefetch(got[:0], 500)) == 0
    # a read made of a short tandem repeat yields far more SMEMs than the default per-read
    # capacity of the kernel's scratch: exercises the overflow-and-retry path
    unit = np.array([0, 1, 2, 3, 0, 0, 1], np.uint8)
    rep = np.tile(unit, 60)
    ref2 = np.concatenate([synth.random_reference(5000, 3), rep, synth.random_reference(5000, 4)])
    idx2 = fmindex.build_fm_index(ref2, device="cpu")
    fmi2 = FMI_search(idx2, ctx=gpu_ctx)
    r0 = rep[5:305].copy()
    r0[::23] = (r0[::23] + 1) & 3          # periodic mismatches: hundreds of SMEMs in one read
    rd = [r0, ref2[100:251].copy(), rep[:200].copy()]
    f2, o2, l2 = ol.pack_reads(rd)
    o = ol.default_seed_opt(min_seed_len=5, split_width=1000, max_mem_intv=1000)
    got2 = fmi2.collect_smem(seed_opt(5, 1.5, 1000, 500, 1000), f2, o2, l2)
    want2, _ = ol.oracle_collect_smem(idx2, o, f2, o2, l2)
    assert np.bincount(want2["rid"]).max() > 200
    assert np.array_equal(fc.smem_matrix(got2), fc.smem_matrix(want2))
    assert np.array_equal(fmi2.get_sa_entries_prefetch(got2, 500), ol.oracle_sa_lookup(idx2, want2, 500)[0])
    FMI_search(idx, ctx=gpu_ctx)      # restore the module's index replica on the shared context


def test_gpu_seeding_config1_shape(gpu_ctx):
    """BASELINE config 1 shape: 1 Mbp random reference, 10 k x 151 bp reads, 1 % substitutions."""
    ref = synth.random_reference(1_000_000, 42)
    idx = fmindex.build_fm_index(ref)
    reads, _, _ = synth.simulate_reads(ref, 10000, seed=1)
    fmi = FMI_search(idx, ctx=gpu_ctx)
    flat, off, ln = ol.pack_reads(reads)
    got = fmi.collect_smem(seed_opt(), flat, off, ln)
    want, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(), flat, off, ln)
    assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(want))
    assert np.array_equal(fmi.get_sa_entries_prefetch(got, 500), ol.oracle_sa_lookup(idx, want, 500)[0])
    # size-independent properties: canonical order, spans inside the read, interval inside the text
    key = (got["rid"].astype(np.int64) << 32) | (got["m"].astype(np.int64) << 16) | got["n"]
    assert np.all(np.diff(key) >= 0)
    assert np.all(got["n"] < 151) and np.all(got["m"] <= got["n"])
    assert np.all(got["s"] >= 1) and np.all(got["k"] + got["s"] <= idx.ref_seq_len)


# ---- FMA tables (SMEM_ACCEL) -------------------------------------------------------------------
def test_gpu_fma_tables_match_reference_bytes(gpu_ctx):
    """Tables built on the device == the bytes the reference's `smem-table` command writes."""
    ref, idx = fc.small_index()
    fmi = FMI_search(idx, ctx=gpu_ctx)      # (earlier tests left another index on the shared context)
    fmi.build_fma()
    a, b = fmi.get_fma()
    assert np.array_equal(a.reshape(-1, 32)[GOLD["fma_all_sample_idx"]], GOLD["fma_all_sample"])
    assert np.array_equal(b.reshape(-1, 4)[GOLD["fma_last_sample_idx"]], GOLD["fma_last_sample"])
    assert np.array_equal(sha(a), GOLD["fma_all_sha256"])
    assert np.array_equal(sha(b), GOLD["fma_last_sha256"])
    rng = np.random.default_rng(5)
    ia, il = rng.integers(0, 1 << 22, 3000), rng.integers(0, 1 << 26, 3000)
    assert np.array_equal(a.reshape(-1, 32)[ia], ol.oracle_fma_entries(idx, "all", ia))
    assert np.array_equal(b.reshape(-1, 4)[il], ol.oracle_fma_entries(idx, "last", il))
    fmi.use_fma(False)


@pytest.mark.parametrize("mode", ["both", "all_only", "last_only", "uploaded"])
def test_gpu_fma_seeding_matches_golden_and_oracle(gpu_ctx, mode):
    ref, idx = fc.small_index()
    fmi = FMI_search(idx, ctx=gpu_ctx)
    fmi.build_fma()
    a, b = fmi.get_fma()
    fa, fl = a, b
    if mode == "all_only":
        fl = None
    if mode == "last_only":
        fa = None
    if mode != "both":
        fmi.load_fma(fa, fl)          # host tables in (what a caller holding the reference's files does)
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    try:
        for name, kw, max_occ in fc.OPT_CASES:
            oo = ol.default_seed_opt(**kw)
            got = fmi.collect_smem(seed_opt(oo.min_seed_len, oo.split_factor, oo.split_width, max_occ,
                                            oo.max_mem_intv), flat, off, ln)
            want, _ = ol.oracle_collect_smem(idx, oo, flat, off, ln, fma=(fa, fl))
            assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(want)), (name, mode)
            if mode in ("both", "uploaded"):
                assert len(got) == int(GOLD[f"fma_{name}_count"][0])
                assert np.array_equal(sha(fc.smem_matrix(got)), GOLD[f"fma_{name}_smem_sha256"])
    finally:
        fmi.load_fma(None, None)


def test_gpu_fma_config1_shape(gpu_ctx):
    """Config-1 shape with the tables on: on N-free 151-bp reads with default options the tables are
    a pure memoisation, so the SMEM list equals the plain one."""
    ref = synth.random_reference(1_000_000, 42)
    idx = fmindex.build_fm_index(ref)
    reads, _, _ = synth.simulate_reads(ref, 10000, seed=1)
    fmi = FMI_search(idx, ctx=gpu_ctx)
    flat, off, ln = ol.pack_reads(reads)
    plain = fmi.collect_smem(seed_opt(), flat, off, ln)
    fmi.build_fma()
    got = fmi.collect_smem(seed_opt(), flat, off, ln)
    assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(plain))
    want, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(), flat, off, ln)
    assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(want))
