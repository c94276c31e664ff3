"""Shared ERT-seeding test workloads (also used by tests/golden/make_golden_ert.py).

The ERT index itself cannot be rebuilt in the test suite (the reference's `index -a ert` enumerates all 4^15
k-mers: minutes of CPU and an 8 GiB table even for a toy genome), so the golden file carries, for a 30 kbp
genome, the reference-built tree blob verbatim and the k-mer table entries that the golden reads touch."""
import os

import numpy as np

from bwa_mem_scale_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "ert_golden.npz")

# (name, seed-opt kwargs).  min_seed_len stays >= 19 = k-mer + x-mer: below that the reference itself asserts
# (src/ertseeding.cpp:2847)
OPT_CASES = [
    ("default", {}),
    ("long_seeds", dict(min_seed_len=25, split_width=50, max_mem_intv=100)),
    ("no_last", dict(max_mem_intv=0, split_factor=1.0)),
    ("tight", dict(min_seed_len=22, split_width=3, max_mem_intv=5)),
]


def tiny_genome():
    """30 kbp: 40 % repeat families, a poly-A run (a FREQUENT k-mer: > 256 hits) and a dinucleotide run."""
    ref = synth.repeat_reference(30000, 11, n_families=4, repeat_frac=0.4, divergence=0.01, unit_lo=100, unit_hi=400)
    ref[1000:1320] = 0
    ref[5000:5600] = np.tile(np.array([0, 1], np.uint8), 300)
    return ref


def _sim(rng, ref, n, L, err, nfrac=0.0):
    out = []
    for _ in range(n):
        p = int(rng.integers(0, len(ref) - L))
        x = ref[p:p + L].copy()
        e = rng.random(L) < err
        x[e] = (x[e] + rng.integers(1, 4, int(e.sum()))) % 4
        if nfrac:
            x[rng.random(L) < nfrac] = 4
        if rng.random() < 0.5:
            x = np.where(x < 4, 3 - x, 4)[::-1]
        out.append(np.ascontiguousarray(x, np.uint8))
    return out


def golden_reads(ref):
    """Ragged reads (1-151 bp, Ns, both strands), simulated 151-mers at several error rates, reads inside the
    poly-A / dinucleotide runs, reads at both ends of the text, and degenerate ones."""
    rng = np.random.default_rng(17)
    reads = [r for r in synth.ragged_reads(ref, 600, 5) if len(r) <= 151]
    reads += _sim(rng, ref, 700, 151, 0.01) + _sim(rng, ref, 200, 151, 0.05) + _sim(rng, ref, 200, 151, 0.0)
    reads += _sim(rng, ref, 200, 151, 0.01, 0.01) + _sim(rng, ref, 200, 60, 0.02) + _sim(rng, ref, 100, 25, 0.0)
    reads += [ref[1000 + k:1151 + k].copy() for k in range(0, 300, 11)]
    reads += [ref[4900 + k:5051 + k].copy() for k in range(0, 700, 17)]
    reads += [ref[0:151].copy(), ref[3:154].copy(), ref[5:100].copy(), (3 - ref[-151:][::-1]).astype(np.uint8),
              ref[-151:].copy(), np.full(50, 4, np.uint8), np.array([0], np.uint8), np.array([4], np.uint8),
              ref[100:119].copy(), ref[100:118].copy(), np.zeros(151, np.uint8)]
    x = ref[2000:2151].copy(); x[0] = 4; reads.append(x)
    x = ref[2000:2151].copy(); x[150] = 4; reads.append(x)
    x = ref[2000:2151].copy(); x[14] = 4; x[16] = 4; reads.append(x)
    return reads


def mem_matrix(m):
    return np.stack([m[f].astype(np.int64) for f in ("rid", "start", "end", "hitcount", "hit_off")], axis=1)
