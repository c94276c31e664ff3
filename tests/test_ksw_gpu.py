"""GPU parity tests (through the C ABI): CUDA mate-rescue SW vs golden reference records / the oracle."""
import os

import numpy as np
import pytest

import ksw_cases as kc
import oracle_lib as ol
from bwa_mem_scale_b200 import MateSW, synth

pytestmark = pytest.mark.gpu
GOLD = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ksw_golden.npz"))


def as_matrix(aln):
    return np.stack([aln[f] for f in aln.dtype.names], axis=1).astype(np.int32)


@pytest.mark.parametrize("case", kc.CASES, ids=[c[0] for c in kc.CASES])
def test_gpu_ksw_matches_golden(gpu_ctx, case):
    name, _, (a, b, od, ed, oi, ei) = case
    pairs, ref, qer = kc.make(case)
    sw = MateSW(ol.score_matrix(a, b), od, ed, oi, ei, ctx=gpu_ctx)
    before = (pairs.copy(), ref.copy(), qer.copy())
    got = sw.mem_sam_pe_batch(pairs, ref, qer)
    assert np.array_equal(as_matrix(got), GOLD[name])
    # the reference reverses its inputs in place and restores them; we never touch them
    assert np.array_equal(pairs, before[0]) and np.array_equal(ref, before[1]) and np.array_equal(qer, before[2])


def test_gpu_ksw_mixed_batch_and_single(gpu_ctx):
    """8-bit and 16-bit jobs interleaved in one batch (mem_sam_pe_batch_pre emits them mixed)."""
    p1, r1, q1 = synth.rescue_jobs(700, seed=21)
    p2, r2, q2 = synth.rescue_jobs(700, seed=22, xtra=0x40000 | 0x80000 | 19)
    p2 = p2.copy(); p2["idr"] += len(r1); p2["idq"] += len(q1)
    pairs = np.concatenate([p1, p2]); ref = np.concatenate([r1, r2]); qer = np.concatenate([q1, q2])
    perm = np.random.default_rng(1).permutation(len(pairs))
    pairs = np.ascontiguousarray(pairs[perm])
    sw = MateSW(ctx=gpu_ctx)
    got = sw.mem_sam_pe_batch(pairs, ref, qer)
    want = ol.oracle_ksw(pairs, ref, qer)
    assert np.array_equal(as_matrix(got), as_matrix(want))
    assert len(sw.mem_sam_pe_batch(pairs[:0], ref, qer)) == 0
    i = 5
    one = sw.ksw_align2(int(pairs[i]["len2"]), qer[pairs[i]["idq"]:], int(pairs[i]["len1"]), ref[pairs[i]["idr"]:],
                        int(pairs[i]["h0"]))
    assert tuple(one) == tuple(want[i])


def test_gpu_ksw_large_batch_properties(gpu_ctx):
    """20 k jobs: oracle on a sample, plus size-independent properties on all of them."""
    pairs, ref, qer = synth.rescue_jobs(20000, seed=31)
    sw = MateSW(ctx=gpu_ctx)
    got = sw.mem_sam_pe_batch(pairs, ref, qer)
    want = ol.oracle_ksw(pairs[:3000], ref, qer)
    assert np.array_equal(as_matrix(got[:3000]), as_matrix(want))
    assert np.all(got["score"] >= 0) and np.all(got["te"] < pairs["len1"]) and np.all(got["qe"] < pairs["len2"])
    ok = got["tb"] >= 0
    assert ok.sum() > 10000
    assert np.all(got["tb"][ok] <= got["te"][ok]) and np.all(got["qb"][ok] <= got["qe"][ok])
    assert np.all(got["score2"] <= got["score"])


def test_gpu_ksw_harvested_pe_jobs(gpu_ctx):
    """Row a19 on REAL work: every ksw_align2 job mem_matesw issued in a paired-end run of the reference
    (2 x 151 bp, insert N(400,50), half of the second mates damaged), answers as the reference returned them."""
    from test_oracle_ksw import pe_rescue_jobs
    pairs, ref, qer, want = pe_rescue_jobs()
    got = MateSW(ctx=gpu_ctx).mem_sam_pe_batch(pairs, ref, qer)
    assert np.array_equal(as_matrix(got), want)
