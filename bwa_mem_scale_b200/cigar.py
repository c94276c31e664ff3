"""Host-side mirror of the reference's CIGAR generation for final alignment regions (seam S4).

`RegToAln.mem_reg2aln` = the alignment part of mem_reg2aln (reference src/bwamem.cpp:2533-2600: band inference, up to
three bwa_gen_cigar2 tries, ksw_global2 with backtrack, NM / MD, position, deletion squeeze, soft clips) for a batch of
regions, behind `bwagpu_reg2aln_batch`.  The reference text must be loaded on the context (Extender)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import CigarOpt, CigarResult, Context

CIGAR_JOB_DTYPE = np.dtype([("rb", "<i8"), ("re", "<i8"), ("read", "<i4"), ("qb", "<i4"), ("qe", "<i4"), ("truesc", "<i4"),
                            ("w", "<i4"), ("pad_", "<i4")])
CIGAR_OUT_DTYPE = np.dtype([("pos", "<i8"), ("cigar_off", "<i8"), ("md_off", "<i8"), ("is_rev", "<i4"), ("n_cigar", "<i4"),
                            ("nm", "<i4"), ("l_md", "<i4"), ("score", "<i4"), ("ok", "<i4")])
assert CIGAR_JOB_DTYPE.itemsize == 40 and CIGAR_OUT_DTYPE.itemsize == 48


def cigar_opt(a=1, b=4, o_del=6, e_del=1, o_ins=6, e_ins=1, w=100) -> CigarOpt:
    """Scoring of mem_opt_init (reference src/bwamem.cpp:135-171); mat as bwa_fill_scmat builds it."""
    o = CigarOpt()
    k = 0
    for i in range(4):
        for j in range(4):
            o.mat[k] = a if i == j else -b
            k += 1
        o.mat[k] = -1
        k += 1
    for j in range(5):
        o.mat[k] = -1
        k += 1
    o.a, o.o_del, o.e_del, o.o_ins, o.e_ins, o.w = a, o_del, e_del, o_ins, e_ins, w
    return o


class RegToAln:
    def __init__(self, ctx: Context):
        self.ctx = ctx

    def mem_reg2aln(self, opt: CigarOpt, enc_reads, off, ln, jobs: np.ndarray):
        """(outs, cigar u32[], md bytes): outs[j].cigar_off / md_off index the two flat arrays."""
        jobs = np.ascontiguousarray(jobs, CIGAR_JOB_DTYPE)
        n = len(jobs)
        outs = np.zeros(n, CIGAR_OUT_DTYPE)
        ccap, mcap = 4 * n + 64, 16 * n + 64
        nc, nm = C.c_int64(0), C.c_int64(0)
        while True:
            cigar, md = np.zeros(ccap, np.uint32), np.zeros(mcap, np.uint8)
            rc = self.ctx.lib.bwagpu_reg2aln_batch(
                self.ctx.h, C.byref(opt), enc_reads.ctypes.data, off.ctypes.data, ln.ctypes.data, len(ln),
                jobs.ctypes.data, n, outs.ctypes.data, cigar.ctypes.data, ccap, md.ctypes.data, mcap, C.byref(nc), C.byref(nm))
            if rc == -5 and (nc.value > ccap or nm.value > mcap):
                ccap, mcap = max(ccap, nc.value), max(mcap, nm.value)
                continue
            self.ctx.check(rc, "bwagpu_reg2aln_batch")
            return outs, cigar[:nc.value], md[:nm.value]

    def mem_reg2aln_device(self, opt: CigarOpt, d_reads: int, d_off: int, d_len: int, nseq: int, d_jobs: int, n_jobs: int) -> CigarResult:
        """Same with reads and jobs resident in HBM; results stay on the device (context-owned)."""
        res = CigarResult()
        self.ctx.check(self.ctx.lib.bwagpu_reg2aln_device(
            self.ctx.h, C.byref(opt), C.c_void_p(d_reads), C.c_void_p(d_off), C.c_void_p(d_len), nseq, C.c_void_p(d_jobs),
            n_jobs, C.byref(res)), "bwagpu_reg2aln_device")
        return res

    def fetch(self, res: CigarResult):
        """Copy a device result back: (outs, cigar, md)."""
        outs, cigar, md = np.zeros(res.n_jobs, CIGAR_OUT_DTYPE), np.zeros(res.n_cigar, np.uint32), np.zeros(res.md_bytes, np.uint8)
        for arr, ptr in ((outs, res.d_outs), (cigar, res.d_cigar), (md, res.d_md)):
            if arr.nbytes:
                self.ctx.d2h(arr, ptr)
        return outs, cigar, md
