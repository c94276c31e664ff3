"""CPU tests: the per-thread seeding code of the CUDA kernels (csrc/fmi_core.cuh), single-stepped on
the host, against the oracle — checks the state machine's logic before it ever reaches a GPU."""
import ctypes as C

import numpy as np
import pytest

import fmi_cases as fc
import hostsim_lib
import oracle_lib as ol
from bwa_mem_scale_b200.lib import FmIndexDesc, SeedOpt


def desc_of(idx):
    d = FmIndexDesc()
    d.reference_seq_len = idx.ref_seq_len
    for i, v in enumerate(idx.count_loaded()):
        d.count[i] = int(v)
    d.cp_occ, d.sa_ms_byte, d.sa_ls_word = idx.cp_occ.ctypes.data, idx.sa_ms_byte.ctypes.data, idx.sa_ls_word.ctypes.data
    d.sentinel_index = idx.sentinel_index
    return d


@pytest.mark.parametrize("case", fc.OPT_CASES, ids=[c[0] for c in fc.OPT_CASES])
def test_state_machine_matches_oracle(case):
    name, kw, max_occ = case
    lib = hostsim_lib.load()
    ref, idx = fc.small_index()
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    oo = ol.default_seed_opt(**kw)
    want, n_ext_oracle = ol.oracle_collect_smem(idx, oo, flat, off, ln)
    o = SeedOpt(oo.min_seed_len, oo.split_factor, oo.split_width, max_occ, oo.max_mem_intv)
    d = desc_of(idx)
    out = np.zeros(len(want) + 16, ol.SMEM_DTYPE)
    n_ext = C.c_int64(0)
    tot = lib.hostsim_collect_smem(C.byref(d), C.byref(o), C.c_void_p(flat.ctypes.data),
                                   C.c_void_p(off.ctypes.data), C.c_void_p(ln.ctypes.data), None,
                                   C.c_int32(len(ln)), C.c_void_p(out.ctypes.data), C.c_int64(len(out)),
                                   C.c_int(1024), C.byref(n_ext), None, None)
    assert tot == len(want)
    assert np.array_equal(fc.smem_matrix(out[:tot]), fc.smem_matrix(want))
    # the kernel skips the reference's duplicated prevArray entry: never more extensions than it
    assert 0 < n_ext.value <= n_ext_oracle
    wsa, _ = ol.oracle_sa_lookup(idx, want, max_occ)
    co = np.zeros(len(wsa) + 8, np.int64)
    n = lib.hostsim_sa_lookup(C.byref(d), C.c_void_p(out.ctypes.data), C.c_int64(tot), C.c_int32(max_occ),
                              C.c_void_p(co.ctypes.data))
    assert n == len(wsa) and np.array_equal(co[:n], wsa)


@pytest.mark.parametrize("mode", ["both", "all_only", "last_only"])
def test_state_machine_fma_matches_oracle(mode):
    """The kernels' FMA path (fma_forward / fma_last_jump) against the oracle's FMA mode."""
    from test_oracle_fmi import fma_tables
    lib = hostsim_lib.load()
    ref, idx = fc.small_index()
    reads = fc.small_reads(ref)
    flat, off, ln = ol.pack_reads(reads)
    fa, fl = fma_tables(idx, reads)
    if mode == "all_only":
        fl = None
    if mode == "last_only":
        fa = None
    d = desc_of(idx)
    for name, kw, max_occ in fc.OPT_CASES:
        oo = ol.default_seed_opt(**kw)
        want, _ = ol.oracle_collect_smem(idx, oo, flat, off, ln, fma=(fa, fl))
        o = SeedOpt(oo.min_seed_len, oo.split_factor, oo.split_width, max_occ, oo.max_mem_intv)
        out = np.zeros(len(want) + 16, ol.SMEM_DTYPE)
        n_ext = C.c_int64(0)
        tot = lib.hostsim_collect_smem(C.byref(d), C.byref(o), C.c_void_p(flat.ctypes.data),
                                       C.c_void_p(off.ctypes.data), C.c_void_p(ln.ctypes.data), None,
                                       C.c_int32(len(ln)), C.c_void_p(out.ctypes.data), C.c_int64(len(out)),
                                       C.c_int(4096), C.byref(n_ext),
                                       C.c_void_p(fa.ctypes.data) if fa is not None else None,
                                       C.c_void_p(fl.ctypes.data) if fl is not None else None)
        assert tot == len(want), (name, mode)
        assert np.array_equal(fc.smem_matrix(out[:tot]), fc.smem_matrix(want)), (name, mode)
