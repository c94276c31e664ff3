"""FM-index container and builder, byte-compatible with the reference's `<prefix>.bwt.2bit.64`.

The layout is the one `FMI_search::build_fm_index` writes and `FMI_search::load_index` reads
(reference src/FMI_search.cpp:611-771, 1251-1351):

    int64  ref_seq_len (= 2*l_pac + 1)
    int64  count[5]                      (cumulative base counts, BEFORE the +1 applied at load)
    CP_OCC cp_occ[(ref_seq_len >> 6) + 1]   64 B each: int64 cp_count[4]; uint64 one_hot_bwt_str[4]
    int8   sa_ms_byte[(ref_seq_len >> 3) + 1]
    uint32 sa_ls_word[(ref_seq_len >> 3) + 1]
    int64  sentinel_index

The reference index can be loaded verbatim (`FMIndex.load`).  `build_fm_index` exists so that
synthetic benchmarks can make an index on the GPU box without the reference binary: a prefix-doubling
suffix sort written with torch ops (runs on CUDA when available, on CPU otherwise).  The reference's
own builder (SA-IS) is out of scope; tests check that the bytes produced here are identical to its.
"""
from __future__ import annotations

import os
from dataclasses import dataclass

import numpy as np

CP_SHIFT = 6
SA_COMPX = 3


@dataclass
class FMIndex:
    ref_seq_len: int            # 2*l_pac + 1 (text + sentinel)
    count: np.ndarray           # int64[5], file values (load adds 1 to each: FMI_search.cpp:1304-1307)
    cp_occ: np.ndarray          # int64[(n>>6)+1, 8]: 4 counts then 4 one-hot words (bit pattern)
    sa_ms_byte: np.ndarray      # int8[(n>>3)+1]
    sa_ls_word: np.ndarray      # uint32[(n>>3)+1]
    sentinel_index: int

    @property
    def l_pac(self) -> int:
        return (self.ref_seq_len - 1) // 2

    def count_loaded(self) -> np.ndarray:
        """count[] as the search code sees it after load_index (each entry +1)."""
        return (self.count + 1).astype(np.int64)

    def nbytes(self) -> int:
        return self.cp_occ.nbytes + self.sa_ms_byte.nbytes + self.sa_ls_word.nbytes

    def tobytes(self) -> bytes:
        return b"".join([
            np.int64(self.ref_seq_len).tobytes(), self.count.astype("<i8").tobytes(),
            self.cp_occ.astype("<i8").tobytes(), self.sa_ms_byte.astype(np.int8).tobytes(),
            self.sa_ls_word.astype("<u4").tobytes(), np.int64(self.sentinel_index).tobytes()])

    def save(self, prefix: str) -> None:
        with open(prefix + ".bwt.2bit.64", "wb") as f:
            f.write(self.tobytes())

    @staticmethod
    def load(prefix: str) -> "FMIndex":
        path = prefix + ".bwt.2bit.64"
        with open(path, "rb") as f:
            n = int(np.fromfile(f, "<i8", 1)[0])
            count = np.fromfile(f, "<i8", 5)
            n_cp = (n >> CP_SHIFT) + 1
            cp = np.fromfile(f, "<i8", n_cp * 8).reshape(n_cp, 8)
            n_sa = (n >> SA_COMPX) + 1
            ms = np.fromfile(f, np.int8, n_sa)
            ls = np.fromfile(f, "<u4", n_sa)
            sent = int(np.fromfile(f, "<i8", 1)[0])
        if len(ls) != n_sa:
            raise ValueError(f"{path}: truncated index file")
        return FMIndex(n, count, cp, ms, ls, sent)


def text_with_revcomp(ref_fwd: np.ndarray) -> np.ndarray:
    """The `.0123` text: forward strand then its reverse complement (FMI_search.cpp:587-607)."""
    ref_fwd = np.asarray(ref_fwd, dtype=np.uint8)
    return np.concatenate([ref_fwd, (3 - ref_fwd[::-1]).astype(np.uint8)])


def write_0123(prefix: str, ref_fwd: np.ndarray) -> np.ndarray:
    t = text_with_revcomp(ref_fwd)
    t.tofile(prefix + ".0123")
    return t


def suffix_array(text, device=None):
    """Suffix array of `text` (values 0..3) with the implicit end-of-text smallest, by prefix
    doubling with torch sorts.  Returns an int64 torch tensor of length len(text)."""
    import torch
    if device is None:
        device = "cuda" if torch.cuda.is_available() else "cpu"
    t = torch.as_tensor(np.asarray(text), device=device).to(torch.int64) + 1   # 0 = past the end
    n = t.numel()
    K = 27                                        # 5**27 < 2**63
    pad = torch.zeros(n + K, dtype=torch.int64, device=device)
    pad[:n] = t
    key = torch.zeros(n, dtype=torch.int64, device=device)
    for j in range(K):
        key = key * 5 + pad[j:j + n]
    del pad, t

    def dense_rank(sorted_keys, order):
        new = torch.ones(n, dtype=torch.int64, device=device)
        new[1:] = (sorted_keys[1:] != sorted_keys[:-1]).to(torch.int64)
        r_sorted = torch.cumsum(new, 0) - 1
        rank = torch.empty(n, dtype=torch.int64, device=device)
        rank[order] = r_sorted
        return rank, int(r_sorted[-1].item()) + 1

    skey, order = torch.sort(key)
    rank, distinct = dense_rank(skey, order)
    del key, skey
    h = K
    while distinct < n:
        nxt = torch.zeros(n, dtype=torch.int64, device=device)
        if h < n:
            nxt[:n - h] = rank[h:] + 1
        key = rank * (n + 1) + nxt
        del nxt
        skey, order = torch.sort(key)
        rank, distinct = dense_rank(skey, order)
        del key, skey
        h *= 2
    return order


def build_fm_index(ref_fwd: np.ndarray, device=None) -> FMIndex:
    """FM-index of forward+reverse-complement text, identical to what the reference's `index`
    command writes for the same sequence."""
    import torch
    if device is None:
        device = "cuda" if torch.cuda.is_available() else "cpu"
    ref_fwd = np.asarray(ref_fwd, dtype=np.uint8)
    assert ref_fwd.max(initial=0) < 4, "reference must be 0..3 (Ns already replaced, as in .pac)"
    text = text_with_revcomp(ref_fwd)
    n_txt = len(text)
    n = n_txt + 1
    sa = suffix_array(text, device)
    # row 0 is the empty suffix (suffix_array[0] = pac_len upstream, FMI_search.cpp:838)
    sa_full = torch.empty(n, dtype=torch.int64, device=device)
    sa_full[0] = n_txt
    sa_full[1:] = sa
    del sa
    tt = torch.as_tensor(text, device=device)
    bwt = torch.full((n,), 4, dtype=torch.uint8, device=device)
    nz = sa_full != 0
    bwt[nz] = tt[sa_full[nz] - 1]
    sentinel_index = int(torch.nonzero(~nz)[0, 0].item())
    del tt, nz

    bc = np.bincount(ref_fwd, minlength=4).astype(np.int64)
    bc = bc + bc[::-1]                       # the reverse complement adds the complementary counts
    count = np.zeros(5, dtype=np.int64)
    count[1:] = np.cumsum(bc)

    # checkpointed occurrence table: one record per 64 BWT symbols
    n_cp = (n >> CP_SHIFT) + 1
    padded = torch.full((n_cp * 64,), 6, dtype=torch.uint8, device=device)   # DUMMY_CHAR
    padded[:n] = bwt
    blocks = padded.view(n_cp, 64)
    cp = torch.zeros((n_cp, 8), dtype=torch.int64, device=device)
    weights = torch.zeros(64, dtype=torch.int64, device=device)
    for j in range(64):                      # MSB = first symbol of the block (cpp:701-713)
        weights[j] = -(1 << 63) if j == 0 else (1 << (63 - j))
    for c in range(4):
        m = (blocks == c)
        per_block = m.sum(dim=1, dtype=torch.int64)
        cp[:, c] = torch.cumsum(per_block, 0) - per_block
        cp[:, 4 + c] = (m.to(torch.int64) * weights[None, :]).sum(dim=1)
    sa_s = sa_full[::8]
    n_sa = (n >> SA_COMPX) + 1
    assert sa_s.numel() == n_sa or sa_s.numel() == n_sa - 1
    ms = np.zeros(n_sa, dtype=np.int8)
    ls = np.zeros(n_sa, dtype=np.uint32)
    sa_np = sa_s.cpu().numpy()
    ms[:len(sa_np)] = ((sa_np >> 32) & 0xFF).astype(np.int8)
    ls[:len(sa_np)] = (sa_np & 0xFFFFFFFF).astype(np.uint32)
    return FMIndex(n, count, cp.cpu().numpy(), ms, ls, sentinel_index)


def write_bns_files(prefix: str, ref_fwd: np.ndarray, name: str = "chr1") -> None:
    """<prefix>.ann / .amb / .pac for a single N-free contig, byte-identical to what the reference's
    `index` sub-command writes (bns_fasta2bntseq + bns_dump, reference src/bntseq.cpp:256-381, 81-131):
    .ann/.amb are text, .pac packs 4 bases per byte (first base in the top bits) followed by an extra
    zero byte when l_pac is a multiple of 4 and a final byte holding l_pac % 4."""
    ref_fwd = np.asarray(ref_fwd, dtype=np.uint8)
    n = len(ref_fwd)
    with open(prefix + ".ann", "w") as f:
        f.write(f"{n} 1 11\n0 {name} (null)\n0 {n} 0\n")
    with open(prefix + ".amb", "w") as f:
        f.write(f"{n} 1 0\n")
    pad = (-n) % 4
    v = np.concatenate([ref_fwd, np.zeros(pad, np.uint8)]).reshape(-1, 4)
    packed = ((v[:, 0] << 6) | (v[:, 1] << 4) | (v[:, 2] << 2) | v[:, 3]).astype(np.uint8)
    tail = [0, 0] if n % 4 == 0 else [n % 4]
    np.concatenate([packed, np.array(tail, np.uint8)]).tofile(prefix + ".pac")


def write_index_files(prefix: str, ref_fwd: np.ndarray, idx: "FMIndex") -> None:
    """Everything FMI_search::load_index + bwa_idx_load_ele read for a plain FM-index run."""
    idx.save(prefix)
    write_0123(prefix, ref_fwd)
    write_bns_files(prefix, ref_fwd)


def write_fasta(path: str, ref_fwd: np.ndarray, name: str = "chr1", width: int = 80) -> None:
    letters = np.frombuffer(b"ACGTN", dtype=np.uint8)[np.asarray(ref_fwd, dtype=np.uint8)]
    with open(path, "wb") as f:
        f.write(b">" + name.encode() + b"\n")
        for i in range(0, len(letters), width * 4096):
            chunk = letters[i:i + width * 4096]
            rows = [chunk[j:j + width].tobytes() for j in range(0, len(chunk), width)]
            f.write(b"\n".join(rows) + b"\n")


def write_fasta_contigs(path: str, ref_fwd: np.ndarray, contig_len, width: int = 80) -> None:
    """Multi-contig FASTA: `ref_fwd` cut into consecutive contigs chr1, chr2, ... of the given lengths."""
    letters = np.frombuffer(b"ACGTN", dtype=np.uint8)[np.asarray(ref_fwd, dtype=np.uint8)]
    assert int(np.sum(contig_len)) == len(letters)
    o = 0
    with open(path, "wb") as f:
        for i, ln in enumerate(contig_len):
            f.write(b">chr%d\n" % (i + 1))
            chunk = letters[o:o + int(ln)]
            f.write(b"\n".join(chunk[j:j + width].tobytes() for j in range(0, len(chunk), width)) + b"\n")
            o += int(ln)


def write_fastq(path: str, reads: np.ndarray, lens=None) -> None:
    letters = np.frombuffer(b"ACGTN", dtype=np.uint8)
    with open(path, "wb") as f:
        for i, r in enumerate(reads):
            if lens is not None:
                r = r[:lens[i]]
            s = letters[np.minimum(r, 4)].tobytes()
            f.write(b"@r%d\n%s\n+\n%s\n" % (i, s, b"I" * len(s)))
