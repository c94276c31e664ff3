"""Host-side mirror of the worker-level entry (include/bwagpu.h, section W): one chunk of reads in, per-read
alignment regions out — what the two kt_for passes of the reference's mem_process_seqs (worker_bwt + worker_aln,
reference src/bwamem.cpp:1872-1876) leave in w.regs / seqs[i].perfect before mem_sort_dedup_patch.  A thin ctypes
pass-through: all composition (EMF -> seeding -> SA -> chaining -> extension -> compaction) happens inside
libbwagpu.so."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .emf import PERFECT_DTYPE
from .extend import REG_DTYPE, chain_opt, ext_opt
from .lib import Context, MemOpt, MemResult
from .seeding import seed_opt

BATCH_SIZE = 512            # reference src/macro.h:63


def mem_opt(seed=None, chain=None, ext=None, use_ert=False, use_emf=False, batch_size=BATCH_SIZE) -> MemOpt:
    return MemOpt(seed or seed_opt(), chain or chain_opt(), ext or ext_opt(), int(use_ert), int(use_emf), int(batch_size), 0)


class MemChunk:
    """bwagpu_mem_process_chunk + bwagpu_mem_fetch on one context (the index replicas are loaded through
    FMI_search / Extender / PerfectMatch / ErtSeeding on the same context)."""

    def __init__(self, ctx: Context):
        self.ctx = ctx

    def process(self, opt: MemOpt, enc_reads: np.ndarray, off: np.ndarray, ln: np.ndarray) -> MemResult:
        res = MemResult()
        self.ctx.check(self.ctx.lib.bwagpu_mem_process_chunk(self.ctx.h, C.byref(opt), enc_reads.ctypes.data,
                                                             off.ctypes.data, ln.ctypes.data, len(ln), C.byref(res)),
                       "bwagpu_mem_process_chunk")
        return res

    def process_device(self, opt: MemOpt, d_reads: int, d_off: int, d_len: int, nseq: int, max_len: int) -> MemResult:
        res = MemResult()
        self.ctx.check(self.ctx.lib.bwagpu_mem_process_chunk_device(self.ctx.h, C.byref(opt), C.c_void_p(d_reads),
                                                                    C.c_void_p(d_off), C.c_void_p(d_len), nseq, max_len,
                                                                    C.byref(res)), "bwagpu_mem_process_chunk_device")
        return res

    def fetch(self, res: MemResult, nseq: int, regs=None, reg_off=None, emf_ret=None, perfect=None):
        """Copies the results out; allocates the arrays the caller did not pass."""
        regs = np.zeros(res.n_regs, REG_DTYPE) if regs is None else regs
        reg_off = np.zeros(nseq + 1, np.int64) if reg_off is None else reg_off
        emf_ret = np.zeros(nseq, np.int32) if emf_ret is None else emf_ret
        perfect = np.zeros(nseq, PERFECT_DTYPE) if perfect is None else perfect
        assert len(regs) >= res.n_regs and len(reg_off) >= nseq + 1
        self.ctx.check(self.ctx.lib.bwagpu_mem_fetch(self.ctx.h, C.byref(res), nseq, regs.ctypes.data, reg_off.ctypes.data,
                                                     emf_ret.ctypes.data, perfect.ctypes.data), "bwagpu_mem_fetch")
        return regs[:res.n_regs], reg_off, emf_ret, perfect
