"""Host-side mirror of the reference's mate-rescue Smith-Waterman interface (seam S3).

`ksw_align2` keeps the reference's name and argument order (reference src/ksw.cpp:347-350);
`mem_sam_pe_batch` is the batched form the AVX512 build of the reference feeds (src/bwamem_pair.cpp:
880-979): jobs as SeqPair records whose h0 field carries `xtra`.  Both run on the CUDA kernels behind
`bwagpu_ksw_align2_batch` (include/bwagpu.h); there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import Context, KswParams
from .synth import SEQPAIR_DTYPE

# byte-for-byte mirror of kswr_t (reference src/ksw.h)
KSWR_DTYPE = np.dtype([(n, "<i4") for n in ("score", "te", "qe", "score2", "te2", "tb", "qb")])
assert KSWR_DTYPE.itemsize == 28
KSW_XBYTE, KSW_XSTOP, KSW_XSUBO, KSW_XSTART = 0x10000, 0x20000, 0x40000, 0x80000


def fill_scmat(a: int = 1, b: int = 4) -> np.ndarray:
    """bwa_fill_scmat (reference src/bwa.cpp:368-378): 5x5 matrix, ambiguous base scores -1."""
    m = np.full((5, 5), -1, np.int8)
    for i in range(4):
        for j in range(4):
            m[i, j] = a if i == j else -b
    return np.ascontiguousarray(m.reshape(-1))


class MateSW:
    def __init__(self, mat: np.ndarray | None = None, o_del=6, e_del=1, o_ins=6, e_ins=1,
                 ctx: Context | None = None, device: int = 0):
        self._own = ctx is None
        self.ctx = ctx if ctx is not None else Context(device)
        self.prm = KswParams()
        mat = fill_scmat() if mat is None else np.asarray(mat, np.int8).reshape(-1)
        if len(mat) != 25:
            raise ValueError("ksw: the score matrix must be 5x5 (m = 5)")
        for i in range(25):
            self.prm.mat[i] = int(mat[i])
        self.prm.o_del, self.prm.e_del, self.prm.o_ins, self.prm.e_ins = o_del, e_del, o_ins, e_ins

    def close(self):
        if self._own and self.ctx is not None:
            self.ctx.close()
        self.ctx = None

    def mem_sam_pe_batch(self, seqPairArray: np.ndarray, seqBufRef: np.ndarray, seqBufQer: np.ndarray,
                         pcnt: int | None = None) -> np.ndarray:
        """aln[i] = ksw_align2(len2, qer+idq, len1, ref+idr, 5, mat, ..., xtra = h0, 0)."""
        n = len(seqPairArray) if pcnt is None else pcnt
        if seqPairArray.dtype != SEQPAIR_DTYPE or not seqPairArray.flags.c_contiguous:
            raise ValueError("seqPairArray must be a contiguous SEQPAIR_DTYPE array")
        aln = np.zeros(n, KSWR_DTYPE)
        self.ctx.check(self.ctx.lib.bwagpu_ksw_align2_batch(
            self.ctx.h, C.byref(self.prm), seqPairArray.ctypes.data, seqBufRef.ctypes.data,
            seqBufQer.ctypes.data, aln.ctypes.data, n), "bwagpu_ksw_align2_batch")
        return aln

    def ksw_align2(self, qlen: int, query: np.ndarray, tlen: int, target: np.ndarray, xtra: int) -> np.void:
        """One job, argument order of the reference's ksw_align2 (m and mat come from the constructor)."""
        sp = np.zeros(1, SEQPAIR_DTYPE)
        sp["len1"], sp["len2"], sp["h0"] = tlen, qlen, xtra
        q = np.ascontiguousarray(query[:qlen], np.uint8)
        t = np.ascontiguousarray(target[:tlen], np.uint8)
        return self.mem_sam_pe_batch(sp, t if tlen else np.zeros(1, np.uint8), q if qlen else np.zeros(1, np.uint8))[0]

    def batch_device(self, d_pairs: int, d_ref: int, d_qer: int, d_aln: int, n: int) -> None:
        self.ctx.check(self.ctx.lib.bwagpu_ksw_align2_batch_device(
            self.ctx.h, C.byref(self.prm), C.c_void_p(d_pairs), C.c_void_p(d_ref), C.c_void_p(d_qer),
            C.c_void_p(d_aln), n), "bwagpu_ksw_align2_batch_device")
