"""CPU tests: the oracle's banded-SW restatement against (a) golden vectors produced by the
reference's own SSE4.1/AVX2 kernels and (b) the reference libs directly when oracle/_ref exists."""
import numpy as np
import pytest

import oracle_lib as ol
from bsw_cases import CASES, GOLDEN, results_matrix
from bwa_mem_scale_b200 import synth

F = synth.RESULT_FIELDS


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_oracle_matches_golden(case):
    name, gen, prm, kind, w = case
    pairs, ref, qer = synth.extension_pairs(**gen)
    out, cells = ol.oracle_bsw(ol.default_params(**prm), kind, pairs, ref, qer, w)
    assert np.array_equal(results_matrix(out, F), GOLDEN[name].astype(np.int64))
    if kind == 0:
        assert cells > 0


@pytest.mark.parametrize("isa", ["sse41", "avx2"])
@pytest.mark.parametrize("kind", [0, 8, 16])
def test_oracle_matches_reference_lib(isa, kind):
    lib = ol.ref_bsw(isa)
    if lib is None:
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    for seed, prm, w in [(21, {}, 100), (22, dict(e_del=2, e_ins=2, zdrop=15), 25),
                         (23, dict(o_del=4, o_ins=8, e_del=2, e_ins=1, zdrop=30, end_bonus=0), 64)]:
        pairs, ref, qer = synth.extension_pairs(4000, seed=seed, n_rate=0.003, indel_rate=0.02,
                                                tiny_frac=0.05, eight_bit=(kind == 8))
        P = ol.default_params(**prm)
        want = ol.reference_bsw(lib, P, kind, pairs, ref, qer, w)
        got, _ = ol.oracle_bsw(P, kind, pairs, ref, qer, w)
        assert np.array_equal(results_matrix(got, F), results_matrix(want, F))


def test_scalar_equals_vector_for_unit_gap_extension():
    # with e_del = e_ins = 1 the three kernels agree
    pairs, ref, qer = synth.extension_pairs(3000, seed=31, n_rate=0.002, indel_rate=0.02)
    P = ol.default_params()
    a, _ = ol.oracle_bsw(P, 0, pairs, ref, qer, 100)
    b, _ = ol.oracle_bsw(P, 16, pairs, ref, qer, 100)
    assert np.array_equal(results_matrix(a, F), results_matrix(b, F))


def test_degenerate_pairs():
    # len1 == 0 (seed at a contig end): score stays h0, all ends 0, gscore -1
    pairs = np.zeros(2, dtype=synth.SEQPAIR_DTYPE)
    pairs["len2"] = [10, 1]
    pairs["len1"] = [0, 1]
    pairs["h0"] = [37, 19]
    pairs["idq"] = [0, 10]
    pairs["idr"] = [0, 0]
    qer = np.zeros(16, np.uint8)
    ref = np.zeros(16, np.uint8)
    for kind in (0, 16, 8):
        out, _ = ol.oracle_bsw(ol.default_params(), kind, pairs, ref, qer, 100)
        assert out["score"][0] == 37 and out["tle"][0] == 0 and out["qle"][0] == 0
        assert out["gscore"][0] == -1 and out["gtle"][0] == 0
        assert out["score"][1] == 20 and out["qle"][1] == 1 and out["tle"][1] == 1
        assert out["gscore"][1] == 20 and out["gtle"][1] == 1
