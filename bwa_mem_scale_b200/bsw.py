"""Host-side mirror of the reference's `BandedPairWiseSW` class for the extension seam (S2).

Same constructor arguments and the same three batch calls as reference src/bandedSWA.h:113-172:
`getScores16`, `getScores8`, `scalarBandedSWAWrapper`, each taking
(pairArray, seqBufRef, seqBufQer, numPairs, numThreads, w) and writing results in place.  The work
is done by the CUDA kernels behind `bwagpu_bsw_get_scores` (include/bwagpu.h); `numThreads` is
accepted for signature parity and ignored.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import BswParams, Context
from .synth import SEQPAIR_DTYPE

KIND_SCALAR, KIND_8BIT, KIND_16BIT = 0, 8, 16


class BandedPairWiseSW:
    def __init__(self, o_del: int, e_del: int, o_ins: int, e_ins: int, zdrop: int, end_bonus: int,
                 mat=None, w_match: int = 1, w_mismatch: int = 4, numThreads: int = 1,
                 ctx: Context | None = None, device: int = 0):
        # `mat` is accepted like upstream; the kernels use (w_match, -w_mismatch, -1) exactly as the
        # reference's vector code does (src/bandedSWA.cpp:3366-3372)
        self.params = BswParams(o_del, e_del, o_ins, e_ins, zdrop, end_bonus, w_match, w_mismatch)
        self._own = ctx is None
        self.ctx = ctx if ctx is not None else Context(device)

    def close(self):
        if self._own and self.ctx is not None:
            self.ctx.close()
        self.ctx = None

    # -- host-buffer calls (the drop-in signatures) ---------------------------------------------
    def _run(self, kind: int, pairArray: np.ndarray, seqBufRef: np.ndarray, seqBufQer: np.ndarray,
             numPairs: int, w: int) -> None:
        assert pairArray.dtype == SEQPAIR_DTYPE and pairArray.flags.c_contiguous
        assert seqBufRef.dtype == np.uint8 and seqBufQer.dtype == np.uint8
        rc = self.ctx.lib.bwagpu_bsw_get_scores(
            self.ctx.h, C.byref(self.params), kind, pairArray.ctypes.data,
            seqBufRef.ctypes.data, seqBufQer.ctypes.data, int(numPairs), int(w))
        self.ctx.check(rc, f"bwagpu_bsw_get_scores(kind={kind})")

    def getScores16(self, pairArray, seqBufRef, seqBufQer, numPairs, numThreads, w):
        self._run(KIND_16BIT, pairArray, seqBufRef, seqBufQer, numPairs, w)

    def getScores8(self, pairArray, seqBufRef, seqBufQer, numPairs, numThreads, w):
        self._run(KIND_8BIT, pairArray, seqBufRef, seqBufQer, numPairs, w)

    def scalarBandedSWAWrapper(self, pairArray, seqBufRef, seqBufQer, numPairs, numThreads, w):
        self._run(KIND_SCALAR, pairArray, seqBufRef, seqBufQer, numPairs, w)

    # -- device-resident call (inputs already in HBM) --------------------------------------------
    def getScoresDevice(self, kind: int, d_pairs: int, d_ref: int, ref_bytes: int, d_qer: int,
                        qer_bytes: int, numPairs: int, w: int) -> None:
        rc = self.ctx.lib.bwagpu_bsw_get_scores_device(
            self.ctx.h, C.byref(self.params), kind, C.c_void_p(d_pairs), C.c_void_p(d_ref),
            int(ref_bytes), C.c_void_p(d_qer), int(qer_bytes), int(numPairs), int(w))
        self.ctx.check(rc, f"bwagpu_bsw_get_scores_device(kind={kind})")
