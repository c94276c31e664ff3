"""Deterministic synthetic inputs for the seeding + extension hot path.

Shapes follow BASELINE.json `configs` / SURVEY.md §8(d): i.i.d. ACGT references, 151-bp reads with
1 % substitutions and random strand, and seed-extension `SeqPair` sets for the isolated banded-SW
sweep (config 5).  Everything is seeded numpy; nothing here touches the GPU.
"""
from __future__ import annotations

import numpy as np

# byte-for-byte mirror of SeqPair (reference src/bandedSWA.h:90-99)
SEQPAIR_DTYPE = np.dtype(
    [("idr", "<i4"), ("idq", "<i4"), ("id", "<i4"), ("len1", "<i4"), ("len2", "<i4"),
     ("h0", "<i4"), ("seqid", "<i4"), ("regid", "<i4"), ("score", "<i4"), ("tle", "<i4"),
     ("gtle", "<i4"), ("qle", "<i4"), ("gscore", "<i4"), ("max_off", "<i4")])
assert SEQPAIR_DTYPE.itemsize == 56

RESULT_FIELDS = ("score", "qle", "tle", "gtle", "gscore", "max_off")


def random_reference(n_bp: int, seed: int) -> np.ndarray:
    """Uniform i.i.d. reference, values 0..3 (A,C,G,T)."""
    return np.random.default_rng(seed).integers(0, 4, size=n_bp, dtype=np.uint8)


def simulate_reads(ref: np.ndarray, n_reads: int, read_len: int = 151, err: float = 0.01,
                   seed: int = 1, n_frac: float = 0.0) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
    """Reads as an (n, read_len) uint8 matrix of codes 0..4, plus true start and strand.

    Each base is substituted (to a different base) with probability `err`; strand flipped with
    probability 0.5.  `n_frac` sprinkles N (code 4).
    """
    rng = np.random.default_rng(seed)
    start = rng.integers(0, len(ref) - read_len + 1, size=n_reads)
    idx = start[:, None] + np.arange(read_len)[None, :]
    reads = ref[idx]
    sub = rng.random(reads.shape) < err
    reads = np.where(sub, (reads + rng.integers(1, 4, size=reads.shape)) & 3, reads).astype(np.uint8)
    rev = rng.random(n_reads) < 0.5
    rc = (3 - reads[:, ::-1])
    reads = np.where(rev[:, None], rc, reads).astype(np.uint8)
    if n_frac > 0:
        reads = np.where(rng.random(reads.shape) < n_frac, 4, reads).astype(np.uint8)
    return reads, start, rev


def extension_pairs(n_pairs: int, seed: int = 7, max_len2: int = 150, min_len2: int = 1,
                    sub_rate: float = 0.01, indel_rate: float = 0.001, n_rate: float = 0.0,
                    h0_lo: int = 19, h0_hi: int = 151, eight_bit: bool = False,
                    tiny_frac: float = 0.0, a: int = 1, o: int = 6, e: int = 1, w: int = 100,
                    end_bonus: int = 5):
    """Seed-extension jobs shaped like those mem_chain2aln_across_reads_V2 emits.

    Query = the read part beyond the seed (len2), target = reference window of len2 + max-gap
    (cal_max_gap, reference src/bwamem.cpp:94-104) that begins with a mutated copy of the query.
    Returns (pairs[SEQPAIR_DTYPE], ref_buf u8, qer_buf u8).  With `eight_bit` every pair satisfies
    the 8-bit routing rule len1,len2 < 128 and h0 + min(len)*a < 128 (bwamem.cpp:2684-2694).
    """
    rng = np.random.default_rng(seed)
    if eight_bit:
        max_len2 = min(max_len2, 100)
    len2 = rng.integers(min_len2, max_len2 + 1, size=n_pairs)
    # max gap the extension could afford (cal_max_gap): (len*a - o)/e + 1, clipped to [1, 2w]
    gap = np.maximum(((len2 * a + end_bonus - o) / e + 1.0).astype(np.int64), 1)
    gap = np.minimum(gap, w << 1)
    len1 = len2 + gap
    if tiny_frac > 0:
        tiny = rng.random(n_pairs) < tiny_frac
        len1 = np.where(tiny, rng.integers(0, 5, size=n_pairs), len1)
    if eight_bit:
        len1 = np.minimum(len1, 127)
        h0 = rng.integers(h0_lo, 128, size=n_pairs)
        h0 = np.minimum(h0, 127 - np.minimum(len1, len2) * a)
        h0 = np.maximum(h0, 1)
    else:
        h0 = rng.integers(h0_lo, h0_hi + 1, size=n_pairs)

    pairs = np.zeros(n_pairs, dtype=SEQPAIR_DTYPE)
    offq = np.concatenate(([0], np.cumsum(len2)))
    offr = np.concatenate(([0], np.cumsum(len1)))
    qer = rng.integers(0, 4, size=int(offq[-1]), dtype=np.uint8)
    ref = rng.integers(0, 4, size=int(offr[-1]), dtype=np.uint8)
    # copy query into the head of the target with substitutions and (rare) indels
    for i in range(n_pairs):
        q = qer[offq[i]:offq[i + 1]]
        t = ref[offr[i]:offr[i + 1]]
        m = q.copy()
        s = rng.random(len(m)) < sub_rate
        m[s] = (m[s] + rng.integers(1, 4, size=int(s.sum()))) & 3
        if indel_rate > 0 and len(m) > 2:
            k = rng.random(len(m)) < indel_rate
            if k.any():
                keep = []
                for pos in range(len(m)):
                    if k[pos]:
                        if rng.random() < 0.5:
                            continue                       # deletion from the target copy
                        keep.append(rng.integers(0, 4))    # insertion into the target copy
                    keep.append(m[pos])
                m = np.asarray(keep, dtype=np.uint8)
        n = min(len(m), len(t))
        t[:n] = m[:n]
    if n_rate > 0:
        qer[rng.random(len(qer)) < n_rate] = 4
        ref[rng.random(len(ref)) < n_rate] = 4
    pairs["idr"] = offr[:-1]
    pairs["idq"] = offq[:-1]
    pairs["id"] = np.arange(n_pairs)
    pairs["len1"] = len1
    pairs["len2"] = len2
    pairs["h0"] = h0
    pairs["seqid"] = np.arange(n_pairs) // 3
    pairs["regid"] = np.arange(n_pairs) % 3
    return pairs, ref, qer


def extension_pairs_fast(n_pairs: int, seed: int = 7, max_len2: int = 150, min_len2: int = 1,
                         sub_rate: float = 0.01, indel_frac: float = 0.1, h0_lo: int = 19,
                         h0_hi: int = 151, eight_bit: bool = False, a: int = 1, o: int = 6,
                         e: int = 1, w: int = 100, end_bonus: int = 5):
    """Vectorised variant of `extension_pairs` for bench-sized sets (millions of pairs): same
    length/h0 distributions; a fraction `indel_frac` of pairs carries one 1-3 bp indel."""
    rng = np.random.default_rng(seed)
    if eight_bit:
        max_len2 = min(max_len2, 100)
    len2 = rng.integers(min_len2, max_len2 + 1, size=n_pairs)
    gap = np.maximum(((len2 * a + end_bonus - o) / e + 1.0).astype(np.int64), 1)
    gap = np.minimum(gap, w << 1)
    len1 = len2 + gap
    if eight_bit:
        len1 = np.minimum(len1, 127)
        h0 = np.maximum(np.minimum(rng.integers(h0_lo, 128, size=n_pairs),
                                   127 - np.minimum(len1, len2) * a), 1)
    else:
        h0 = rng.integers(h0_lo, h0_hi + 1, size=n_pairs)
    offq = np.concatenate(([0], np.cumsum(len2)))
    offr = np.concatenate(([0], np.cumsum(len1)))
    qer = rng.integers(0, 4, size=int(offq[-1]), dtype=np.uint8)
    ref = rng.integers(0, 4, size=int(offr[-1]), dtype=np.uint8)
    # target position p (within its pair) copies query position p + shift(p); shift models one indel
    pid = np.repeat(np.arange(n_pairs), len1)
    pos = np.arange(int(offr[-1])) - offr[:-1][pid]
    has = rng.random(n_pairs) < indel_frac
    ipos = (rng.random(n_pairs) * len2).astype(np.int64)
    ilen = rng.integers(1, 4, size=n_pairs) * np.where(rng.random(n_pairs) < 0.5, 1, -1)
    shift = np.where(has[pid] & (pos >= ipos[pid]), ilen[pid], 0)
    src = pos + shift
    ok = (src >= 0) & (src < len2[pid]) & (pos < len2[pid] + 3)
    vals = qer[np.minimum(offq[:-1][pid] + np.clip(src, 0, None), len(qer) - 1)]
    sub = rng.random(len(vals)) < sub_rate
    vals = np.where(sub, (vals + rng.integers(1, 4, size=len(vals))) & 3, vals).astype(np.uint8)
    ref = np.where(ok, vals, ref).astype(np.uint8)
    pairs = np.zeros(n_pairs, dtype=SEQPAIR_DTYPE)
    pairs["idr"] = offr[:-1]
    pairs["idq"] = offq[:-1]
    pairs["id"] = np.arange(n_pairs)
    pairs["len1"] = len1
    pairs["len2"] = len2
    pairs["h0"] = h0
    pairs["seqid"] = np.arange(n_pairs) // 3
    pairs["regid"] = np.arange(n_pairs) % 3
    return pairs, ref, qer


def repeat_reference(n_bp: int, seed: int, n_families: int = 50, repeat_frac: float = 0.10,
                     divergence: float = 0.02, unit_lo: int = 200, unit_hi: int = 2000) -> np.ndarray:
    """Random reference in which `repeat_frac` of the sequence is drawn from `n_families` repeat
    families, each copy mutated at `divergence`: keeps SMEM intervals
    wider than 1 for longer, as on real genomes."""
    rng = np.random.default_rng(seed)
    ref = rng.integers(0, 4, size=n_bp, dtype=np.uint8)
    fams = [rng.integers(0, 4, size=int(rng.integers(unit_lo, unit_hi + 1)), dtype=np.uint8)
            for _ in range(n_families)]
    target = int(n_bp * repeat_frac)
    placed = 0
    while placed < target:
        f = fams[int(rng.integers(0, n_families))]
        if len(f) >= n_bp:
            break
        cp = f.copy()
        mut = rng.random(len(cp)) < divergence
        cp[mut] = (cp[mut] + rng.integers(1, 4, size=int(mut.sum()))) & 3
        if rng.random() < 0.5:
            cp = (3 - cp[::-1]).astype(np.uint8)
        p = int(rng.integers(0, n_bp - len(cp)))
        ref[p:p + len(cp)] = cp
        placed += len(cp)
    return ref


def ragged_reads(ref: np.ndarray, n_reads: int, seed: int, len_lo: int = 10, len_hi: int = 200,
                 err: float = 0.01, n_frac: float = 0.002):
    """Reads of mixed lengths (incl. shorter than the minimum seed length) as a list of arrays."""
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n_reads):
        L = int(rng.integers(len_lo, len_hi + 1))
        s = int(rng.integers(0, len(ref) - L + 1))
        r = ref[s:s + L].copy()
        sub = rng.random(L) < err
        r[sub] = (r[sub] + rng.integers(1, 4, size=int(sub.sum()))) & 3
        if rng.random() < 0.5:
            r = (3 - r[::-1]).astype(np.uint8)
        r[rng.random(L) < n_frac] = 4
        out.append(r)
    return out


def rescue_jobs(n_jobs: int, seed: int = 11, qlen_lo: int = 151, qlen_hi: int = 151, win_lo: int = 200,
                win_hi: int = 900, sub_rate: float = 0.02, indel_rate: float = 0.003, n_rate: float = 0.0,
                hit_frac: float = 0.7, xtra: int | None = None, a: int = 1, min_seed_len: int = 19,
                repeat_frac: float = 0.2):
    """Mate-rescue jobs shaped like those mem_matesw hands to ksw_align2 (reference
    src/bwamem_pair.cpp:170-260): query = a whole read (len2), target = an insert-size window of the
    reference (len1) that contains a mutated copy of the read in `hit_frac` of the jobs (and a second,
    more diverged copy in some, so that score2/te2 are exercised).  h0 carries `xtra` as in
    mem_sam_pe_batch_pre (KSW_XSUBO | KSW_XSTART | KSW_XBYTE when len2*a < 250 | min_seed_len*a).
    Returns (pairs[SEQPAIR_DTYPE], ref_buf, qer_buf)."""
    rng = np.random.default_rng(seed)
    pairs = np.zeros(n_jobs, SEQPAIR_DTYPE)
    refs, qers = [], []
    ro = qo = 0
    for i in range(n_jobs):
        ql = int(rng.integers(qlen_lo, qlen_hi + 1))
        wl = int(rng.integers(max(win_lo, 1), win_hi + 1))
        win = rng.integers(0, 4, wl).astype(np.uint8)
        q = rng.integers(0, 4, ql).astype(np.uint8)
        if rng.random() < hit_frac and wl > 8:
            for rep, rate in enumerate((sub_rate, 6 * sub_rate)):
                if rep == 1 and rng.random() > repeat_frac:
                    break
                src = q.copy()
                sub = rng.random(ql) < rate
                src = np.where(sub, (src + rng.integers(1, 4, ql)) & 3, src).astype(np.uint8)
                if rng.random() < indel_rate * ql and ql > 20:      # one indel of 1-6 bases
                    p, g = int(rng.integers(5, ql - 5)), int(rng.integers(1, 7))
                    src = (np.concatenate([src[:p], src[p + g:]]) if rng.random() < 0.5 else
                           np.concatenate([src[:p], rng.integers(0, 4, g).astype(np.uint8), src[p:]]))
                cut = int(rng.integers(0, max(len(src) // 3, 1))) if rng.random() < 0.2 else 0
                src = src[cut:]
                pos = int(rng.integers(0, max(wl - len(src), 0) + 1))
                seg = src[:max(wl - pos, 0)]
                win[pos:pos + len(seg)] = seg
        if n_rate > 0:
            q = np.where(rng.random(ql) < n_rate, 4, q).astype(np.uint8)
            win = np.where(rng.random(wl) < n_rate, 4, win).astype(np.uint8)
        x = xtra if xtra is not None else (0x40000 | 0x80000 | (0x10000 if ql * a < 250 else 0) | (min_seed_len * a))
        pairs[i]["idr"], pairs[i]["idq"], pairs[i]["id"] = ro, qo, i
        pairs[i]["len1"], pairs[i]["len2"], pairs[i]["h0"] = wl, ql, x
        pairs[i]["seqid"], pairs[i]["regid"] = i, i
        refs.append(win); qers.append(q)
        ro += wl; qo += ql
    return pairs, np.concatenate(refs + [np.zeros(64, np.uint8)]), np.concatenate(qers + [np.zeros(64, np.uint8)])


def gapped_reads(text: np.ndarray, n_reads: int, seed: int, read_len: int = 151, err: float = 0.02,
                 small_indel_frac: float = 0.30, long_indel_frac: float = 0.10, n_run_frac: float = 0.05,
                 starts=None):
    """Reads with gaps, as a list of arrays: each read is cut from `text` (any 0..3 array: the concatenated
    contigs, or the forward + reverse-complement text when reads across the strand boundary are wanted), then
    a fraction carries one 1-3 base insertion or deletion, a fraction one 5-15 base one (wide enough to push
    the banded extension's max_off past 3/4 of a narrow band, which triggers the band-doubling retry of
    mem_chain2aln_across_reads_V2, reference src/bwamem.cpp:3248), a fraction a run of 2-6 N; substitutions at
    `err`; random strand.  `starts` pins the first reads to given text positions (contig ends etc.)."""
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n_reads):
        span = read_len + 16
        s = int(starts[i]) if starts is not None and i < len(starts) else int(rng.integers(0, len(text) - span))
        src = text[s:s + span].copy()
        u = rng.random()
        if u < small_indel_frac + long_indel_frac:
            g = int(rng.integers(1, 4)) if u < small_indel_frac else int(rng.integers(5, 16))
            p = int(rng.integers(25, read_len - 25))
            if rng.random() < 0.5:
                src = np.concatenate([src[:p], src[p + g:]])                                   # deletion
            else:
                src = np.concatenate([src[:p], rng.integers(0, 4, g).astype(np.uint8), src[p:]])   # insertion
        r = src[:read_len].copy()
        sub = rng.random(len(r)) < err
        r[sub] = (r[sub] + rng.integers(1, 4, size=int(sub.sum()))) & 3
        if rng.random() < n_run_frac:
            p, g = int(rng.integers(0, len(r) - 6)), int(rng.integers(2, 7))
            r[p:p + g] = 4
        if rng.random() < 0.5:
            r = np.where(r > 3, 4, 3 - r)[::-1].astype(np.uint8)
        out.append(np.ascontiguousarray(r, np.uint8))
    return out


def simulate_pairs(ref: np.ndarray, n_pairs: int, read_len: int = 151, insert_mean: float = 400.0,
                   insert_sd: float = 50.0, err: float = 0.01, seed: int = 1):
    """Paired-end reads in FR orientation (BASELINE config 4): fragment length ~ N(insert_mean, insert_sd), read 1 from
    the fragment's left end on the forward strand, read 2 the reverse complement of its right end; the whole pair is
    flipped with probability 0.5.  Returns two (n, read_len) uint8 matrices."""
    rng = np.random.default_rng(seed)
    frag = np.clip(rng.normal(insert_mean, insert_sd, n_pairs).astype(np.int64), read_len, None)
    start = (rng.random(n_pairs) * (len(ref) - frag)).astype(np.int64)
    idx = np.arange(read_len)[None, :]
    a = ref[start[:, None] + idx]
    b = 3 - ref[(start + frag)[:, None] - 1 - idx]
    flip = rng.random(n_pairs) < 0.5
    r1 = np.where(flip[:, None], b, a)
    r2 = np.where(flip[:, None], a, b)
    out = []
    for r in (r1, r2):
        sub = rng.random(r.shape) < err
        out.append(np.where(sub, (r + rng.integers(1, 4, size=r.shape)) & 3, r).astype(np.uint8))
    return out[0], out[1]
