"""bwagpu_bsw_get_scores_noalloc (no counterpart in the reference):
the allocation-free banded-SW entry point must give exactly what bwagpu_bsw_get_scores_device gives, on two calls
in a row that share the static scratch array.  Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as ol
from bsw_cases import results_matrix
from bwa_mem_scale_b200 import BswParams, synth

pytestmark = pytest.mark.gpu
F = synth.RESULT_FIELDS
NOALLOC_MAX_LEN2 = 1020


def _params():
    P = ol.default_params()
    return BswParams(P.o_del, P.e_del, P.o_ins, P.e_ins, P.zdrop, P.end_bonus, P.w_match, P.w_mismatch)


def _both(ctx, prm, kind, pairs, ref, qer, w):
    """Results of the existing device entry point and of the no-alloc one on the same uploaded inputs."""
    lib = ctx.lib
    dr, dq = ctx.upload(ref), ctx.upload(qer)
    dp_old, dp_new = ctx.upload(pairs), ctx.upload(pairs)
    ctx.check(lib.bwagpu_bsw_get_scores_device(ctx.h, C.byref(prm), kind, C.c_void_p(dp_old), C.c_void_p(dr), ref.nbytes,
                                               C.c_void_p(dq), qer.nbytes, len(pairs), w), "bsw_get_scores_device")
    ctx.check(lib.bwagpu_bsw_get_scores_noalloc(ctx.h, C.byref(prm), kind, C.c_void_p(dp_new), C.c_void_p(dr),
                                                C.c_void_p(dq), len(pairs), int(pairs["len2"].max()), w),
              "bsw_get_scores_noalloc")
    ctx.synchronize()
    old, new = pairs.copy(), pairs.copy()
    ctx.d2h(old, dp_old)
    ctx.d2h(new, dp_new)
    for p in (dp_old, dp_new, dr, dq):
        ctx.free(p)
    return old, new


def test_noalloc_matches_device_entry_twice(gpu_ctx):
    prm = _params()
    # call 1: 16-bit kind, long-ish queries (leaves wide column state behind in the shared array)
    pairs, ref, qer = synth.extension_pairs(6000, seed=901, indel_rate=0.01, n_rate=0.002, max_len2=600, min_len2=30,
                                            h0_hi=300)
    old, new = _both(gpu_ctx, prm, 16, pairs, ref, qer, 100)
    assert np.array_equal(results_matrix(new, F), results_matrix(old, F))
    # call 2, straight after: other pairs, other kinds, narrower band — stale scratch contents must not matter
    for kind, seed, w in ((8, 902, 50), (0, 903, 100), (16, 904, 25)):
        pairs, ref, qer = synth.extension_pairs(20000, seed=seed, indel_rate=0.01, n_rate=0.002, tiny_frac=0.02,
                                                eight_bit=(kind == 8))
        old, new = _both(gpu_ctx, prm, kind, pairs, ref, qer, w)
        assert np.array_equal(results_matrix(new, F), results_matrix(old, F)), kind
        for f in ("idr", "idq", "id", "len1", "len2", "h0", "seqid", "regid"):
            assert np.array_equal(new[f], pairs[f])


def test_noalloc_rejects_oversize_and_marks_unbounded_pairs(gpu_ctx):
    prm = _params()
    lib = gpu_ctx.lib
    pairs, ref, qer = synth.extension_pairs(256, seed=905)
    dp, dr, dq = gpu_ctx.upload(pairs), gpu_ctx.upload(ref), gpu_ctx.upload(qer)
    rc = lib.bwagpu_bsw_get_scores_noalloc(gpu_ctx.h, C.byref(prm), 16, C.c_void_p(dp), C.c_void_p(dr), C.c_void_p(dq),
                                           len(pairs), NOALLOC_MAX_LEN2 + 1, 100)
    assert rc == -5 and b"exceeds the fixed scratch" in lib.bwagpu_last_error(gpu_ctx.h)
    # a bound below the longest query: those pairs are left uncomputed (-1), the others are exact
    bound = int(np.median(pairs["len2"]))
    gpu_ctx.check(lib.bwagpu_bsw_get_scores_noalloc(gpu_ctx.h, C.byref(prm), 16, C.c_void_p(dp), C.c_void_p(dr),
                                                    C.c_void_p(dq), len(pairs), bound, 100), "noalloc")
    gpu_ctx.synchronize()
    got = pairs.copy()
    gpu_ctx.d2h(got, dp)
    for p in (dp, dr, dq):
        gpu_ctx.free(p)
    want, _ = ol.oracle_bsw(ol.default_params(), 16, pairs, ref, qer, 100)
    over = pairs["len2"] > bound
    assert over.any() and (~over).any()
    assert np.all(results_matrix(got, F)[over] == -1)
    assert np.array_equal(results_matrix(got, F)[~over], results_matrix(want, F)[~over])
