"""Shared chain -> alignment-region test workloads (also used by tests/golden/make_golden_c2a.py)."""
import numpy as np

import fmi_cases as fc
from bwa_mem_scale_b200 import fmindex, synth


def genome():
    ref, _ = fc.small_index()
    return ref, fmindex.text_with_revcomp(ref)


def golden_reads(ref):
    """400 simulated 151-bp reads + the ragged edge-case reads of the seeding tests."""
    reads, _, _ = synth.simulate_reads(ref, 400, seed=21, err=0.02)
    return [r for r in reads] + fc.small_reads(ref)[-10:] + fc.small_reads(ref)[:150]


# (name, mem_opt overrides)
OPT_CASES = [
    ("default", {}),
    ("narrow_band", dict(w=10, zdrop=20)),
    ("penalties", dict(a=2, b=5, o_del=5, e_del=2, o_ins=7, e_ins=1, pen_clip5=3, pen_clip3=8)),
]
REG_FIELDS = ("rb", "re", "qb", "qe", "rid", "score", "truesc", "w", "seedcov", "seedlen0", "chain")


def reg_matrix(regs):
    return np.stack([regs[f].astype(np.int64) for f in REG_FIELDS], axis=1)
