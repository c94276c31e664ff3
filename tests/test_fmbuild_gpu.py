"""GPU tests of bwagpu_index_build_fm: the FM-index built in HBM from the resident text must have the bytes of the
index the reference's `index` sub-command writes (tests/test_oracle_fmi.py pins fmindex.build_fm_index to those bytes on
the CPU; tests/golden/fmi_golden.npz holds a sha of the reference-built file)."""
import os

import numpy as np
import pytest

import c2a_cases as cc
import fmi_cases as fc
from bwa_mem_scale_b200 import Extender, FMI_search, fmindex, synth

pytestmark = pytest.mark.gpu


def build_on_device(ctx, ref, P=None):
    Extender(fmindex.text_with_revcomp(ref), len(ref), [0], [len(ref)], ctx=ctx)
    if P is None:
        os.environ.pop("BWAGPU_FMB_P", None)
    else:
        os.environ["BWAGPU_FMB_P"] = str(P)
    try:
        return FMI_search(None, ctx=ctx).get_index()
    finally:
        os.environ.pop("BWAGPU_FMB_P", None)


def same(a, b):
    return (a.ref_seq_len == b.ref_seq_len and a.sentinel_index == b.sentinel_index and np.array_equal(a.count, b.count)
            and np.array_equal(a.cp_occ, b.cp_occ) and np.array_equal(a.sa_ms_byte, b.sa_ms_byte)
            and np.array_equal(a.sa_ls_word, b.sa_ls_word))


@pytest.mark.parametrize("P", [0, 1, 3])
def test_device_builder_matches_reference_bytes_repeats(gpu_ctx, P):
    """300 kbp with 40 % repeats, a 400-base poly-A run and a 600-base dinucleotide run: runs of equal 21-symbol keys
    hundreds of suffixes long, ~30 refinement rounds; one bucket, 4 and 64 buckets."""
    ref, want = fc.small_index()
    got = build_on_device(gpu_ctx, ref, P)
    assert same(got, want)
    assert got.tobytes() == want.tobytes()


def test_device_builder_other_genomes(gpu_ctx):
    for ref in (cc.hard_genome()[0], synth.random_reference(1_000_000, 42), synth.random_reference(1000, 3),
                np.zeros(700, np.uint8), np.array([0, 1, 2, 3, 3, 2, 1, 0, 0, 0], np.uint8),
                synth.repeat_reference(3_000_000, 43, n_families=20)):
        want = fmindex.build_fm_index(ref, device="cpu")
        assert same(build_on_device(gpu_ctx, ref, 2), want), len(ref)
        if len(ref) <= 1000:
            assert same(build_on_device(gpu_ctx, ref, 6), want), len(ref)


def test_device_built_index_seeds_like_the_uploaded_one(gpu_ctx):
    """Seeding on the device-built replica (no host copy in between) gives the golden SMEMs."""
    import oracle_lib as ol
    from bwa_mem_scale_b200 import seed_opt
    ref, idx = fc.small_index()
    Extender(fmindex.text_with_revcomp(ref), len(ref), [0], [len(ref)], ctx=gpu_ctx)
    fmi = FMI_search(None, ctx=gpu_ctx)
    flat, off, ln = ol.pack_reads(fc.small_reads(ref))
    got = fmi.collect_smem(seed_opt(), flat, off, ln)
    want, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(), flat, off, ln)
    assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(want))
