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


# ---- "hard" workload: three contigs, reads with indels / N runs, reads over contig ends and over the
# forward / reverse-complement boundary (reference src/bwamem.cpp:2909-2913, 3028-3051,
# 3248, 3320, 3388) ----
HARD_CONTIGS = (70000, 50000, 40000)


def hard_genome():
    ref = synth.repeat_reference(sum(HARD_CONTIGS), 17, n_families=6, repeat_frac=0.25, divergence=0.02,
                                 unit_lo=150, unit_hi=700)
    return ref, fmindex.text_with_revcomp(ref)


def hard_reads(ref):
    text = fmindex.text_with_revcomp(ref)
    l_pac = len(ref)
    ends = np.cumsum(HARD_CONTIGS)
    # reads that straddle each contig end, the strand boundary, and both ends of the text
    pins = [int(e) - k for e in ends[:-1] for k in (20, 60, 75, 100, 131)]
    pins += [l_pac - k for k in (15, 40, 76, 110, 140)] + [0, 1, 2 * l_pac - 167 - 1]
    reads = synth.gapped_reads(text, len(pins), seed=31, small_indel_frac=0.0, long_indel_frac=0.0, n_run_frac=0.0,
                               err=0.01, starts=pins)
    reads += synth.gapped_reads(ref, 700, seed=32)
    # overhanging reads: the alignment starts at text position 0 / l_pac with unaligned query in front (left
    # extension jobs with an empty reference window, len1 == 0)
    rng = np.random.default_rng(33)
    for base in (0, l_pac):
        for k in (12, 25, 40):
            reads.append(np.concatenate([rng.integers(0, 4, k).astype(np.uint8), text[base:base + 151 - k]]))
    return reads


OPT_CASES_HARD = [
    ("default", {}),
    ("narrow_band", dict(w=10, zdrop=20)),
    ("tiny_band", dict(w=4, zdrop=100)),
    ("penalties", dict(a=2, b=5, o_del=5, e_del=2, o_ins=7, e_ins=1, pen_clip5=3, pen_clip3=8, w=12)),
]
