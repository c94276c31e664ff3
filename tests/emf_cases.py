"""Shared EMF (exact-match filter) test workload (also used by tests/golden/make_golden_emf.py)."""
import os

import numpy as np

from bwa_mem_scale_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
SEED_LEN = 101            # table seed length; reads of 101..160 bp exercise seedmatch_further


def genome():
    """(ref_fwd, contig_off, contig_len, hole_off, hole_len): 4 contigs (one shorter than a seed),
    two holes (their bases are whatever the .0123 text holds there), exact and reverse-complement
    duplicates (multi-location lists), and a 700-bp poly-A run (a list past LOC_MANY = 256)."""
    rng = np.random.default_rng(77)
    a = synth.random_reference(30000, 1)
    a[4000:4700] = 0                                       # poly-A: 600 identical seeds
    a[8000:8400] = a[2000:2400]                            # forward duplicate
    a[12000:12400] = a[2000:2400]                          # ... three copies
    a[15000:15300] = (3 - a[6000:6300][::-1])              # reverse-complement duplicate
    a[20000:20250] = (3 - a[2050:2300][::-1])              # rc copy of a multi-copy region
    a[26000:26400] = np.tile(np.array([0, 1, 2], np.uint8), 134)[:400]   # short tandem repeat
    b = synth.random_reference(20000, 2)
    b[9000:9300] = a[2100:2400]                            # duplicate across contigs
    c = synth.random_reference(80, 3)                      # shorter than a seed: contributes nothing
    d = synth.random_reference(10000, 4)
    d[0:200] = a[2000:2200]                                # duplicate at the very start of a contig
    d[-150:] = (3 - a[2000:2150][::-1])                    # rc duplicate at the very end of the text
    ref = np.concatenate([a, b, c, d]).astype(np.uint8)
    lens = np.array([len(a), len(b), len(c), len(d)], np.int32)
    offs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int64)
    hole_off = np.array([offs[1] + 5000, offs[1] + 12000], np.int64)
    hole_len = np.array([300, 10], np.int32)
    del rng
    return ref, offs, lens, hole_off, hole_len


def reads(ref, offs, lens, hole_off, hole_len, n=4000, seed=9):
    """Reads of 60..160 bp cut from the text (both strands), half of them with one substitution, a few
    with an N; plus hand-made cases around contig ends, holes, duplicates and the text's two ends."""
    rng = np.random.default_rng(seed)
    out = []
    L = len(ref)
    for i in range(n):
        ln = int(rng.choice([SEED_LEN, SEED_LEN, 151, 120, 160, 60, 102]))
        p = int(rng.integers(0, L - ln))
        r = ref[p:p + ln].copy()
        if rng.random() < 0.5:
            r = (3 - r[::-1]).astype(np.uint8)
        u = rng.random()
        if u < 0.35:
            q = int(rng.integers(0, ln))
            r[q] = (r[q] + 1 + rng.integers(0, 3)) & 3
        elif u < 0.40:
            r[int(rng.integers(0, ln))] = 4
        out.append(r)

    def fw(p, ln):
        return ref[p:p + ln].copy()

    def rc(p, ln):
        return (3 - ref[p:p + ln][::-1]).astype(np.uint8)

    for ln in (SEED_LEN, 130, 160):
        for p in (2000, 2050, 2100, 8000, 12000, 12399 - ln, 15000, 6000, 20000, 4000, 4100, 4699 - ln, 26000, 26001,
                  int(offs[1]) + 9000, int(offs[3]), int(offs[3]) + 40, L - ln, L - ln - 1, 0, 1,
                  int(offs[1]) - ln, int(offs[1]) - ln + 1, int(offs[1]) - 50,          # contig boundary
                  int(hole_off[0]) - ln, int(hole_off[0]) - ln + 1, int(hole_off[0]) + 100,
                  int(hole_off[0] + hole_len[0]), int(hole_off[0] + hole_len[0]) - 1,
                  int(hole_off[1]) - 30, int(hole_off[1] + hole_len[1]), int(offs[2]) - 20):
            if 0 <= p and p + ln <= L:
                out += [fw(p, ln), rc(p, ln)]
    # seed matches a multi-copy region but the tail matches none of the copies -> SEED_ONLY_MATCHED
    r = fw(2000, 150); r[140] = (r[140] + 1) & 3; out.append(r)
    r = rc(2050, 150); r[120] = (r[120] + 2) & 3; out.append(r)
    # tail matches only a later copy: a[8000:8400] copy continues differently after 8400
    out += [fw(8300, 150), rc(8300, 150), fw(12300, 140), rc(11950, 150), fw(1950, 150)]
    out += [np.zeros(SEED_LEN, np.uint8), np.zeros(160, np.uint8), np.full(SEED_LEN, 3, np.uint8),
            np.full(SEED_LEN, 4, np.uint8), np.zeros(SEED_LEN - 1, np.uint8), np.array([0], np.uint8)]
    return out
