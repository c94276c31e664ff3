"""Sharding of one chunk of reads across the GPUs of a box.

The path has no exchange step: every GPU holds a full replica of the index, rank r takes the
contiguous slice of the chunk `shard_bounds(nseq, world)[r:r+2]`, and the per-rank results are
stitched back in input order on the host (the reference's own unit of work is a batch of reads too:
worker_bwt / worker_aln run per BATCH_SIZE slice, reference src/bwamem.cpp:1489-1543, and results are
consumed by read index).  No collective is involved; only ids and offsets are rebased.
"""
from __future__ import annotations

import numpy as np


def shard_bounds(nseq: int, world: int) -> np.ndarray:
    """First read of every rank's slice (world + 1 entries; slices differ by at most one read)."""
    return np.array([nseq * r // world for r in range(world + 1)], np.int64)


def shard_reads(flat: np.ndarray, off: np.ndarray, ln: np.ndarray, world: int, rank: int):
    """(flat_r, off_r, len_r, first_read) of rank `rank`; offsets rebased to the slice's own buffer."""
    b = shard_bounds(len(ln), world)
    lo, hi = int(b[rank]), int(b[rank + 1])
    if hi == lo:
        return flat[:0], off[:0], ln[:0], lo
    start = int(off[lo])
    end = int((off[lo:hi] + ln[lo:hi]).max())
    return flat[start:end], (off[lo:hi] - start).astype(np.int64), ln[lo:hi], lo


def merge_smems(parts, firsts) -> np.ndarray:
    """SMEM arrays of the ranks (each in the canonical rid, m, n order) -> the array of the whole chunk."""
    out = []
    for p, f in zip(parts, firsts):
        p = p.copy()
        p["rid"] += np.uint32(f)
        out.append(p)
    return np.concatenate(out) if out else np.zeros(0)


def merge_regions(chains, seeds, regs, firsts):
    """Per-rank (chains, seeds, regs) -> whole-chunk arrays: read ids, seed offsets and chain indices rebased."""
    oc, os_, orr = [], [], []
    n_ch = n_sd = 0
    for c, s, r, f in zip(chains, seeds, regs, firsts):
        c, r = c.copy(), r.copy()
        c["seqid"] += f
        c["seed_off"] += n_sd
        r["chain"] += n_ch
        oc.append(c); os_.append(s); orr.append(r)
        n_ch += len(c); n_sd += len(s)
    return np.concatenate(oc), np.concatenate(os_), np.concatenate(orr)
