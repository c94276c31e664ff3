"""GPU parity tests (through the C ABI): CUDA banded SW vs the oracle / golden vectors."""
import numpy as np
import pytest

import oracle_lib as ol
from bsw_cases import CASES, GOLDEN, results_matrix
from bwa_mem_scale_b200 import BandedPairWiseSW, synth

pytestmark = pytest.mark.gpu
F = synth.RESULT_FIELDS


def make_bsw(ctx, **prm):
    P = ol.default_params(**prm)
    return P, BandedPairWiseSW(P.o_del, P.e_del, P.o_ins, P.e_ins, P.zdrop, P.end_bonus, None,
                               P.w_match, P.w_mismatch, 1, ctx=ctx)


def run_kind(bsw, kind, pairs, ref, qer, w):
    got = pairs.copy()
    {0: bsw.scalarBandedSWAWrapper, 8: bsw.getScores8, 16: bsw.getScores16}[kind](
        got, ref, qer, len(got), 1, w)
    return got


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gpu_matches_golden(gpu_ctx, case):
    name, gen, prm, kind, w = case
    pairs, ref, qer = synth.extension_pairs(**gen)
    _, bsw = make_bsw(gpu_ctx, **prm)
    got = run_kind(bsw, kind, pairs, ref, qer, w)
    assert np.array_equal(results_matrix(got, F), GOLDEN[name].astype(np.int64))
    # inputs untouched
    for f in ("idr", "idq", "id", "len1", "len2", "h0", "seqid", "regid"):
        assert np.array_equal(got[f], pairs[f])


@pytest.mark.parametrize("kind", [0, 8, 16])
@pytest.mark.parametrize("w", [16, 25, 50, 75, 100])
def test_gpu_matches_oracle_band_sweep(gpu_ctx, kind, w):
    pairs, ref, qer = synth.extension_pairs(20000, seed=200 + w, n_rate=0.002, indel_rate=0.01,
                                            tiny_frac=0.02, eight_bit=(kind == 8))
    P, bsw = make_bsw(gpu_ctx)
    got = run_kind(bsw, kind, pairs, ref, qer, w)
    want, _ = ol.oracle_bsw(P, kind, pairs, ref, qer, w)
    assert np.array_equal(results_matrix(got, F), results_matrix(want, F))


def test_gpu_empty_and_degenerate(gpu_ctx):
    P, bsw = make_bsw(gpu_ctx)
    empty = np.zeros(0, dtype=synth.SEQPAIR_DTYPE)
    bsw.getScores16(empty, np.zeros(1, np.uint8), np.zeros(1, np.uint8), 0, 1, 100)
    pairs = np.zeros(3, dtype=synth.SEQPAIR_DTYPE)
    pairs["len2"] = [10, 1, 5]
    pairs["len1"] = [0, 1, 3]
    pairs["h0"] = [37, 19, 50]
    pairs["idq"] = [0, 10, 11]
    qer = np.array([0] * 10 + [0] + [4, 4, 1, 2, 3], np.uint8)
    ref = np.array([0, 4, 1, 2], np.uint8)
    pairs["idr"] = [0, 0, 1]
    for kind in (0, 8, 16):
        got = run_kind(bsw, kind, pairs, ref, qer, 100)
        want, _ = ol.oracle_bsw(P, kind, pairs, ref, qer, 100)
        assert np.array_equal(results_matrix(got, F), results_matrix(want, F))


def test_gpu_device_resident_matches_host_path(gpu_ctx):
    pairs, ref, qer = synth.extension_pairs(5000, seed=77, indel_rate=0.01)
    P, bsw = make_bsw(gpu_ctx)
    host = run_kind(bsw, 16, pairs, ref, qer, 100)
    dp, dr, dq = gpu_ctx.upload(pairs), gpu_ctx.upload(ref), gpu_ctx.upload(qer)
    bsw.getScoresDevice(16, dp, dr, ref.nbytes, dq, qer.nbytes, len(pairs), 100)
    gpu_ctx.synchronize()
    back = pairs.copy()
    gpu_ctx.d2h(back, dp)
    for p in (dp, dr, dq):
        gpu_ctx.free(p)
    assert np.array_equal(results_matrix(back, F), results_matrix(host, F))


def test_gpu_long_sequences(gpu_ctx):
    # lengths well past the short-read shapes: exercises the generic path
    pairs, ref, qer = synth.extension_pairs(300, seed=5, max_len2=900, min_len2=400,
                                            indel_rate=0.01, h0_hi=400)
    P, bsw = make_bsw(gpu_ctx)
    for kind in (0, 16):
        got = run_kind(bsw, kind, pairs, ref, qer, 100)
        want, _ = ol.oracle_bsw(P, kind, pairs, ref, qer, 100)
        assert np.array_equal(results_matrix(got, F), results_matrix(want, F))
