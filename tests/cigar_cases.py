"""Shared CIGAR-generation test workloads (also used by tests/golden/make_golden_cigar.py)."""
import os

import numpy as np

import fmi_cases as fc
from bwa_mem_scale_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden", "cigar_golden.npz")
OPT_CASES = [
    ("default", {}),
    ("penalties", dict(a=2, b=5, o_del=5, e_del=2, o_ins=7, e_ins=1, pen_clip5=3, pen_clip3=8)),
    ("narrow_band", dict(w=10, zdrop=20)),
]
# golden record sets: the option cases + "retry" (default options, truesc inflated so that the band-doubling retries run)
GOLDEN_CASES = OPT_CASES + [("retry", {})]


def _indel(rng, x, n_events):
    x = list(x)
    for _ in range(n_events):
        p = int(rng.integers(10, len(x) - 10))
        k = int(rng.integers(1, 6))
        if rng.random() < 0.5:
            del x[p:p + k]
        else:
            x[p:p] = list(rng.integers(0, 4, k))
    return np.array(x[:151], np.uint8)


def golden_reads(ref):
    """Substitution-only reads, reads with 1-3 indels of 1-5 bases (both strands), a few with N, ragged edge cases."""
    rng = np.random.default_rng(31)
    reads, _, _ = synth.simulate_reads(ref, 300, seed=33, err=0.02)
    out = [r for r in reads]
    base, _, _ = synth.simulate_reads(ref, 400, seed=34, err=0.01)
    for i, r in enumerate(base):
        # simulate from a longer window so that deletions do not shorten the read below 151 where possible
        out.append(_indel(rng, r, 1 + i % 3))
    for r in base[:40]:
        x = r.copy()
        x[rng.integers(0, len(x), 2)] = 4
        out.append(x)
    return out + fc.small_reads(ref)[-10:] + fc.small_reads(ref)[:100]


def jobs_from_regs(regs, chains, job_dtype):
    """One CIGAR job per region mem_chain2aln_across_reads_V2 kept (purged ones carry qb == qe == -1)."""
    keep = ~((regs["qb"] == -1) & (regs["qe"] == -1))
    r = regs[keep]
    jobs = np.zeros(len(r), job_dtype)
    jobs["rb"], jobs["re"], jobs["qb"], jobs["qe"] = r["rb"], r["re"], r["qb"], r["qe"]
    jobs["truesc"], jobs["w"] = r["truesc"], r["w"]
    jobs["read"] = chains["seqid"][r["chain"]]
    return jobs


def out_matrix(o):
    return np.stack([o[f].astype(np.int64) for f in ("pos", "cigar_off", "md_off", "is_rev", "n_cigar", "nm", "l_md", "ok")], axis=1)
