"""CPU tests: the per-thread ERT seeding code of the CUDA kernels (csrc/ert_core.cuh), run on the host, against
MEMs / hits / chains the reference itself produced (golden records built by tests/golden/make_golden_ert.py from the
reference's own `index -a ert` tables; live against oracle/_ref when a reference-built index is at hand)."""
import os

import numpy as np
import pytest

import c2a_cases as cc
import ert_cases as ec
import hostsim_lib
import oracle_lib as ol
from bwa_mem_scale_b200 import fmindex
from test_chain_hostsim import hostsim_chains

GOLD = np.load(ec.GOLDEN)
LIVE_PREFIX = os.environ.get("BWAGPU_ERT_PREFIX", "/tmp/ert_tiny/g")


@pytest.fixture(scope="module")
def tiny():
    ref = ec.tiny_genome()
    kmer = np.zeros(1 << 30, np.uint64)            # untouched pages stay virtual
    kmer[GOLD["kmer_idx"].astype(np.int64)] = GOLD["kmer_val"]
    flat, off, ln = ol.pack_reads(ec.golden_reads(ref))
    assert len(ln) == int(GOLD["n_reads"])
    return ref, kmer, np.ascontiguousarray(GOLD["mlt_table"]), fmindex.text_with_revcomp(ref), flat, off, ln


@pytest.mark.parametrize("case", ec.OPT_CASES, ids=[c[0] for c in ec.OPT_CASES])
def test_ert_walker_matches_reference_records(tiny, case):
    name, kw = case
    ref, kmer, mlt, ref_string, flat, off, ln = tiny
    mems, hits, mem_off = hostsim_lib.ert_seed(hostsim_lib.load(), kmer, mlt, ref_string, flat, off, ln, **kw)
    assert np.array_equal(ec.mem_matrix(mems), ec.mem_matrix(GOLD[f"{name}_mems"]))
    assert np.array_equal(hits, GOLD[f"{name}_hits"])
    assert np.array_equal(mem_off, GOLD[f"{name}_mem_off"])
    # structure: per read sorted by (start, end), inside the read, at least min_seed_len long
    msl = kw.get("min_seed_len", 19)
    for r in range(len(ln)):
        m = mems[mem_off[r]:mem_off[r + 1]]
        assert np.all(m["rid"] == r) and np.all(m["end"] <= ln[r]) and np.all(m["end"] - m["start"] >= msl)
        key = m["start"].astype(np.int64) * 1000 + m["end"]
        assert np.all(np.diff(key) >= 0)


def test_ert_scratch_size_does_not_change_results(tiny):
    """The kernel re-runs reads that outgrow their scratch slab; any slab size must give the same answer."""
    ref, kmer, mlt, ref_string, flat, off, ln = tiny
    lib = hostsim_lib.load()
    a = hostsim_lib.ert_seed(lib, kmer, mlt, ref_string, flat, off, ln, per_read_mems=1, per_read_hits=1)
    b = hostsim_lib.ert_seed(lib, kmer, mlt, ref_string, flat, off, ln, per_read_mems=64, per_read_hits=4096)
    assert all(np.array_equal(x, y) for x, y in zip(a, b))


def test_every_hit_is_an_exact_match(tiny):
    ref, kmer, mlt, ref_string, flat, off, ln = tiny
    mems, hits = GOLD["default_mems"], GOLD["default_hits"]
    rng = np.random.default_rng(0)
    for i in rng.choice(len(mems), 400, replace=False):
        m = mems[i]
        q = flat[off[m["rid"]] + m["start"]: off[m["rid"]] + m["end"]]
        for h in hits[m["hit_off"]: m["hit_off"] + min(m["hitcount"], 5)]:
            assert np.array_equal(ref_string[h:h + len(q)], q)


def sampled(mems, hits, max_occ):
    """MEMs + hits -> the SMEM / coordinate form chaining takes (mem_chain_new's hit sampling, bwamem.cpp:985-987)."""
    sm = np.zeros(len(mems), ol.SMEM_DTYPE)
    sm["rid"], sm["m"], sm["n"], sm["s"] = mems["rid"], mems["start"], mems["end"] - 1, mems["hitcount"]
    co = []
    for m in mems:
        hc = int(m["hitcount"])
        step = hc // max_occ if hc > max_occ else 1
        co.append(hits[m["hit_off"]: m["hit_off"] + hc: step][:max_occ])
    return sm, (np.concatenate(co) if co else np.zeros(0, np.int64))


@pytest.mark.parametrize("case", ec.OPT_CASES[:2], ids=[c[0] for c in ec.OPT_CASES[:2]])
def test_chaining_of_ert_mems_matches_mem_kernel1_core_ert(tiny, case):
    """mem_chain_new + mem_chain_flt == the chaining code of the FM path fed with the sampled hits."""
    name, kw = case
    ref, kmer, mlt, ref_string, flat, off, ln = tiny
    o = ol.mem_opt(**kw)
    sm, co = sampled(GOLD[f"{name}_mems"], GOLD[f"{name}_hits"], o["max_occ"])
    chains, seeds = hostsim_chains(o, len(ref), GOLD["contig_off"], GOLD["contig_len"], sm, co, ln)
    assert np.array_equal(chains, GOLD[f"{name}_chains"])
    assert np.array_equal(seeds, GOLD[f"{name}_seeds"])


def test_child_and_leaf_offsets_match_reference_tables():
    """ert_child_off / ert_leaf_off are closed forms of the reference's code_table / leaf_table."""
    if ol.ref_binary() is None:
        pytest.skip("oracle/_ref not built")
    lib = hostsim_lib.load()
    tap = ol.C.CDLL(os.path.join(ol.ORACLE_DIR, "_ref", "libref_api_plain_sse41.so"))
    tap.ref_ert_tables.restype = ol.C.c_void_p
    tabs = [np.ctypeslib.as_array(ol.C.cast(tap.ref_ert_tables(w), ol.C.POINTER(ol.C.c_uint8)), shape=(3, 256, 4))
            for w in (0, 1)]
    for pw in (2, 3, 4):
        for code in range(256):
            types = [(code >> (2 * d)) & 3 for d in range(4)]
            if 2 in types and sum(t != 0 for t in types) > 1:
                continue                     # a UNIFORM child is always an only child
            for c in range(4):
                if types[c] == 1:
                    assert lib.hostsim_ert_offsets(pw, code, c, 0) == tabs[0][pw - 2, code, c]
                if types[c] == 3:
                    assert lib.hostsim_ert_offsets(pw, code, c, 1) == tabs[1][pw - 2, code, c]


@pytest.mark.skipif(not (os.path.exists(LIVE_PREFIX + ".kmer_table") and ol.ref_binary() is not None),
                    reason="no reference-built ERT index at hand (tests/golden/make_golden_ert.py builds one)")
def test_ert_walker_live_against_reference():
    tap = ol.RefErt(LIVE_PREFIX)
    kmer, mlt = tap.kmer_table(), tap.mlt_table()
    ref_string = np.fromfile(LIVE_PREFIX + ".0123", np.uint8)
    ref = ref_string[:len(ref_string) // 2]
    rng = np.random.default_rng(99)
    reads = ec._sim(rng, ref, 3000, 151, 0.02) + ec._sim(rng, ref, 1000, 151, 0.02, 0.01) + ec._sim(rng, ref, 1000, 80, 0.01)
    flat, off, ln = ol.pack_reads(reads)
    for kw in ({}, dict(min_seed_len=21, split_width=4, max_mem_intv=8)):
        want = tap.seed(flat, off, ln, **kw)
        got = hostsim_lib.ert_seed(hostsim_lib.load(), kmer, mlt, ref_string, flat, off, ln, **kw)
        assert np.array_equal(ec.mem_matrix(got[0]), ec.mem_matrix(want[0]))
        assert np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
    tap.close()


def test_kmer_entry_bits_from_fm_index(tiny):
    """csrc/ert_build.cuh: LEP mask, type and reseeding hit count of a k-mer entry follow from occurrence counts in
    the FM-index; checked on every reference-built entry the golden reads touch (absent and occurring k-mers)."""
    import ctypes as C
    from test_seed_hostsim import desc_of
    ref = tiny[0]
    idx = fmindex.build_fm_index(ref, device="cpu")
    kidx, want = np.ascontiguousarray(GOLD["kmer_idx"]), GOLD["kmer_val"]
    out = np.zeros(len(kidx), np.uint64)
    d = desc_of(idx)
    hostsim_lib.load().hostsim_ert_kmer_lep(C.byref(d), C.c_void_p(kidx.ctypes.data), C.c_int64(len(kidx)),
                                            C.c_void_p(out.ctypes.data))
    lep, hits = out & np.uint64(0xFFFF), (out >> np.uint64(16)).astype(np.int64)
    assert np.array_equal(lep, (want >> np.uint64(2)) & np.uint64(0x3FFF))
    typ = np.where(hits == 0, 0, np.where(hits == 1, 1, np.where(hits <= 256, 2, 3)))
    assert np.array_equal(typ, (want & np.uint64(3)).astype(np.int64))
    assert np.array_equal(np.where(hits < 20, hits, 0), ((want >> np.uint64(17)) & np.uint64(0x1F)).astype(np.int64))
    assert (typ == 0).sum() > 1000 and (typ == 3).sum() > 0
