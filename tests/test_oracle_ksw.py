"""CPU tests: the oracle's restatement of ksw_align2 (oracle/ksw_oracle.c) against golden records
produced by the reference's own ksw.cpp, and against the reference library when oracle/_ref exists."""
import os

import numpy as np
import pytest

import ksw_cases as kc
import oracle_lib as ol

GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ksw_golden.npz"))


def as_matrix(aln):
    return np.stack([aln[f] for f in aln.dtype.names], axis=1).astype(np.int32)


@pytest.mark.parametrize("case", kc.CASES, ids=[c[0] for c in kc.CASES])
def test_oracle_matches_golden(case):
    name, _, (a, b, od, ed, oi, ei) = case
    pairs, ref, qer = kc.make(case)
    got = ol.oracle_ksw(pairs, ref, qer, ol.score_matrix(a, b), od, ed, oi, ei)
    assert np.array_equal(as_matrix(got), GOLD[name])


def test_oracle_matches_reference_lib():
    lib = ol.ref_ksw()
    if lib is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    from bwa_mem_scale_b200 import synth
    for seed, kw in enumerate([dict(), dict(xtra=0x40000 | 0x80000 | 19), dict(n_rate=0.02, sub_rate=0.05),
                               dict(qlen_lo=1, qlen_hi=200, win_lo=1, win_hi=300)]):
        pairs, ref, qer = synth.rescue_jobs(400, seed=100 + seed, **kw)
        assert np.array_equal(as_matrix(ol.oracle_ksw(pairs, ref, qer)), as_matrix(ol.reference_ksw(lib, pairs, ref, qer)))


def test_golden_exercises_every_output():
    g = GOLD["matesw_u8"]
    assert (g[:, 3] > 0).sum() > 20 and (g[:, 5] >= 0).sum() > 200        # score2 / tb found
    assert (GOLD["u8_overflow"][:, 0] == 255).sum() > 50                    # 8-bit overflow reported as 255
    assert (GOLD["score_only"][:, 5] == -1).all()                           # no XSTART -> no start


def pe_rescue_jobs():
    """Real mate-rescue jobs recorded from a paired-end run of the reference (tests/golden/make_golden_pe_jobs.py)."""
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "pe_rescue_jobs.npz"))
    return np.ascontiguousarray(g["pairs"]), g["ref"], g["qer"], g["kswr"]


def test_oracle_matches_harvested_pe_jobs():
    """mem_matesw's real ksw_align2 calls (reference src/bwamem_pair.cpp:200) and the records they returned."""
    pairs, ref, qer, want = pe_rescue_jobs()
    assert len(pairs) > 3000 and (want[:, 0] >= 30).sum() > 2000 and (want[:, 0] < 30).sum() > 300
    assert np.all(pairs["h0"] & ol.KSW_XSTART) and np.all(pairs["len1"] >= 500)      # insert-size windows
    got = ol.oracle_ksw(pairs, ref, qer)
    assert np.array_equal(as_matrix(got), want)
