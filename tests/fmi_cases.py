"""Shared FM-seeding test workloads (also used by tests/golden/make_golden_fmi.py)."""
import os

import numpy as np

from bwa_mem_scale_b200 import fmindex, synth

HERE = os.path.dirname(os.path.abspath(__file__))
SMEM_FIELDS = ("rid", "m", "n", "k", "l", "s")

# (name, seed-opt kwargs, max_occ)
OPT_CASES = [
    ("default", {}, 500),
    ("short_seeds", dict(min_seed_len=10, split_width=50, max_mem_intv=100), 50),
    ("no_round3", dict(max_mem_intv=0, split_factor=1.0), 500),
]


def small_genome():
    """300 kbp with heavy repeat content, a poly-A run and a dinucleotide run."""
    ref = synth.repeat_reference(300000, 7, n_families=8, repeat_frac=0.4, divergence=0.01,
                                 unit_lo=100, unit_hi=600)
    ref[1000:1400] = 0
    ref[5000:5600] = np.tile(np.array([0, 1], np.uint8), 300)
    return ref


def small_reads(ref):
    """Ragged reads (10-200 bp, Ns, both strands) plus hand-picked edge cases."""
    reads = synth.ragged_reads(ref, 3000, 5)
    reads += [ref[1100:1251].copy(), ref[5100:5251].copy(), np.full(50, 4, np.uint8),
              ref[0:151].copy(), ref[3:154].copy(), ref[5:100].copy(),
              (3 - ref[-151:][::-1]).astype(np.uint8), ref[-151:].copy(),
              np.array([0], np.uint8), np.array([4], np.uint8)]
    return reads


def smem_matrix(sm):
    return np.stack([sm[f].astype(np.int64) for f in SMEM_FIELDS], axis=1)


_cache = {}


def small_index():
    """FM-index of small_genome() (built by our builder; test_oracle_fmi checks it is byte-identical
    to the reference's)."""
    if "idx" not in _cache:
        _cache["ref"] = small_genome()
        _cache["idx"] = fmindex.build_fm_index(_cache["ref"], device="cpu")
    return _cache["ref"], _cache["idx"]
