"""CPU tests for the EMF row: the oracle's restatement of find_perfect_match_entry
(oracle/emf_oracle.c) and the host table builder (bwagpu_emf_build) against golden data produced by the
reference's own `perfect-index` + find_perfect_match_entry, and against the reference run live when
oracle/_ref exists."""
import hashlib
import os
import tempfile

import numpy as np
import pytest

import emf_cases as ec
import oracle_lib as ol
from bwa_mem_scale_b200 import PerfectTable

GOLD = np.load(os.path.join(ec.HERE, "golden", "emf_golden.npz"))
CASES = [("l101", ec.SEED_LEN, 1.1), ("l151", 151, 1.02), ("l32", 32, 1.5)]
_cache = {}


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def workload():
    if "g" not in _cache:
        g = ec.genome()
        _cache["g"] = g
        _cache["text"] = np.concatenate([g[0], 3 - g[0][::-1]]).astype(np.uint8)
        _cache["reads"] = ol.pack_reads(ec.reads(*g))
    return _cache["g"], _cache["text"], _cache["reads"]


def table(name, L, slack):
    if name not in _cache:
        g = workload()[0]
        _cache[name] = PerfectTable.build(g[0], L, slack, g[1], g[2], g[3], g[4])
    return _cache[name]


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_builder_writes_the_reference_table(case):
    name, L, slack = case
    t = table(*case)
    assert [t.seed_len, t.seq_len, t.num_seed_entry, len(t.loc_table), t.num_seed_used, t.num_seed_key] == list(GOLD[f"{name}_hdr"])
    assert np.array_equal(sha(t.loc_table), GOLD[f"{name}_loc_sha256"])
    assert np.array_equal(sha(t.seed_table), GOLD[f"{name}_seed_sha256"])


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_probe_matches_golden(case):
    name, L, slack = case
    _, text, (flat, off, ln) = workload()
    t = table(*case)
    ret, perfect = ol.oracle_emf_probe(t, text, flat, off, ln)
    assert np.array_equal(ret, GOLD[f"{name}_ret"]) and np.array_equal(perfect, GOLD[f"{name}_perfect"])
    ret, perfect = ol.oracle_emf_probe(t, text, flat, off, ln, num_seed_load=t.num_seed_entry // 2)
    assert np.array_equal(ret, GOLD[f"{name}_half_ret"]) and np.array_equal(perfect, GOLD[f"{name}_half_perfect"])


def test_golden_exercises_every_branch():
    for name, _, _ in CASES:
        assert np.all(np.bincount(GOLD[f"{name}_ret"], minlength=6) > 0), name      # all six FIND_PERFECT_* codes
    r, p = GOLD["l101_ret"], GOLD["l101_perfect"]
    assert ((p[:, 0] >> 2) > 0).sum() > 50                 # answers that carry a multi-location index
    assert (r[:4000] == 5).sum() > 0 and (r == 5).sum() >= 2
    assert np.all(p[(r != 3) & (r != 4)] == 0)             # perfect stays {0,0} unless FW/RC matched
    assert np.all((p[r == 3, 0] & 3) == 1) and np.all((p[r == 4, 0] & 3) == 3)
    assert (GOLD["l101_half_ret"] != r).sum() > 100        # the partial load changes answers


def test_table_file_round_trip_and_no_table():
    _, text, (flat, off, ln) = workload()
    t = table(*CASES[0])
    with tempfile.TemporaryDirectory() as d:
        t.save(d + "/x.perfect.101")
        u = PerfectTable.load(d + "/x.perfect.101")
    assert np.array_equal(u.loc_table, t.loc_table) and np.array_equal(u.seed_table, t.seed_table)
    assert (u.seed_len, u.seq_len, u.num_seed_used, u.num_seed_key) == (t.seed_len, t.seq_len, t.num_seed_used, t.num_seed_key)
    ret, perfect = ol.oracle_emf_probe(None, text, flat, off, ln)
    assert np.all(ret == 0) and np.all(perfect == 0)


def test_live_against_reference():
    """Fresh genome + reads: builder bytes vs the reference's perfect-index file, oracle vs the
    reference's find_perfect_match_entry."""
    if ol.ref_binary("emf") is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    from bwa_mem_scale_b200 import synth
    ref = synth.repeat_reference(120000, 21, n_families=6, repeat_frac=0.3, divergence=0.0, unit_lo=200, unit_hi=900)
    co, cl = np.array([0, 70000], np.int64), np.array([70000, 50000], np.int32)
    ho, hl = np.array([30000], np.int64), np.array([500], np.int32)
    with tempfile.TemporaryDirectory() as d:
        text = ol.write_emf_inputs(d + "/r", ref, co, cl, ho, hl)
        path = ol.ref_perfect_index(d + "/r", 151, 1.1)
        rt = PerfectTable.load(path)
        t = PerfectTable.build(ref, 151, 1.1, co, cl, ho, hl)
        assert np.array_equal(t.loc_table, rt.loc_table) and np.array_equal(t.seed_table, rt.seed_table)
        reads, _, _ = synth.simulate_reads(ref, 3000, seed=3, err=0.004)
        reads = list(reads) + [ref[p:p + 180].copy() for p in range(100, 60000, 977)]
        flat, off, ln = ol.pack_reads(reads)
        r = ol.RefEmf(path, d + "/r.0123")
        want = r.probe(flat, off, ln)
        r.close()
    got = ol.oracle_emf_probe(t, text, flat, off, ln)
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])
    assert (want[0] >= 3).sum() > 500


def test_live_seeding_skips_answered_reads_like_the_reference():
    """Reads the EMF answers are left out of mem_collect_smem (reference src/bwamem.cpp:672-689, and the
    compacted max_intv index of round 3, src/FMI_search.cpp:1679-1739): the oracle's `skip` argument vs
    the reference's PERFECT_MATCH build with seq_[i].perfect.exist set."""
    import ctypes as C
    import subprocess
    import fmi_cases as fc
    from bwa_mem_scale_b200 import fmindex
    lib = ol.ref_api("emf")
    if lib is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    lib.ref_collect_smem_skip.restype = C.c_int64
    lib.ref_collect_smem_skip.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_int64, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p]
    ref, idx = fc.small_index()
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    rng = np.random.default_rng(3)
    with tempfile.TemporaryDirectory() as d:
        fmindex.write_index_files(d + "/g", ref, idx)
        subprocess.run([ol.ref_binary("emf"), "smem-table", d + "/g"], check=True, stdout=subprocess.DEVNULL,
                       stderr=subprocess.DEVNULL)
        fma = (np.fromfile(d + "/g.all_smem.11", np.uint32), np.fromfile(d + "/g.last_smem.13", np.uint32))
        h = lib.ref_fmi_open((d + "/g").encode())
        for name, kw, _ in fc.OPT_CASES:
            for frac in (0.3, 0.9):
                skip = (rng.random(len(ln)) < frac).astype(np.uint8)
                oo = ol.default_seed_opt(**kw)
                cap = int(ln.sum()) * 4
                out = np.zeros(cap, ol.SMEM_DTYPE)
                tot = lib.ref_collect_smem_skip(h, oo.min_seed_len, oo.split_factor, oo.split_width, oo.max_mem_intv,
                                                flat.ctypes.data, off.ctypes.data, ln.ctypes.data, len(ln),
                                                out.ctypes.data, cap, skip.ctypes.data)
                got, _ = ol.oracle_collect_smem(idx, oo, flat, off, ln, skip, fma=fma)
                assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(out[:tot])), (name, frac)
        lib.ref_fmi_close(h)
