"""Shared mate-rescue SW test workloads (also used by tests/golden/make_golden_ksw.py)."""
from bwa_mem_scale_b200 import synth

X_DEF = None                                  # mem_matesw's flags: XSUBO | XSTART | XBYTE(len2*a<250) | min_seed_len
# (name, rescue_jobs kwargs, (a, b, o_del, e_del, o_ins, e_ins))
CASES = [
    ("matesw_u8", dict(n_jobs=600, seed=1), (1, 4, 6, 1, 6, 1)),
    ("matesw_i16", dict(n_jobs=400, seed=2, xtra=0x40000 | 0x80000 | 19), (1, 4, 6, 1, 6, 1)),
    ("score_only", dict(n_jobs=300, seed=3, xtra=0), (1, 4, 6, 1, 6, 1)),
    ("score_only_u8", dict(n_jobs=300, seed=4, xtra=0x10000), (1, 4, 6, 1, 6, 1)),
    ("with_n", dict(n_jobs=300, seed=5, n_rate=0.01), (1, 4, 6, 1, 6, 1)),
    ("ragged", dict(n_jobs=500, seed=6, qlen_lo=10, qlen_hi=300, win_lo=1, win_hi=400), (1, 4, 6, 1, 6, 1)),
    ("u8_overflow", dict(n_jobs=200, seed=7, qlen_lo=300, qlen_hi=400, xtra=0x10000 | 0x40000 | 0x80000 | 19,
                         sub_rate=0.0, indel_rate=0), (1, 4, 6, 1, 6, 1)),
    ("tiny", dict(n_jobs=400, seed=8, qlen_lo=1, qlen_hi=20, win_lo=1, win_hi=30), (1, 4, 6, 1, 6, 1)),
    ("noisy", dict(n_jobs=300, seed=9, sub_rate=0.1, indel_rate=0.02), (1, 4, 6, 1, 6, 1)),
    ("penalties", dict(n_jobs=300, seed=10), (2, 5, 7, 2, 5, 3)),
    ("a3", dict(n_jobs=300, seed=11, a=3), (3, 2, 4, 1, 9, 2)),
    ("long_query", dict(n_jobs=40, seed=12, qlen_lo=900, qlen_hi=1500, win_lo=1200, win_hi=2500,
                        xtra=0x40000 | 0x80000 | 30), (1, 4, 6, 1, 6, 1)),
]


def make(case):
    name, kw, pen = case
    return synth.rescue_jobs(**kw)
