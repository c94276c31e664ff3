"""Host-side mirror of the reference's chain -> alignment-region stage.

`mem_chain2aln_across_reads_V2` keeps the reference's name (reference src/bwamem.cpp:2810-3745); chains
travel as the flattened form of `chain_ar` (CHAIN_DTYPE / SEED_DTYPE), regions come back as REG_DTYPE
records (the mem_alnreg_t fields that function sets).  All work happens in the CUDA kernels behind
`bwagpu_chain2aln` (include/bwagpu.h).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import C2aResult, ChainOpt, ChainResult, Context, ExtOpt

CHAIN_DTYPE = np.dtype([("seqid", "<i4"), ("rid", "<i4"), ("n", "<i4"), ("w", "<i4"), ("kept", "<i4"),
                        ("first", "<i4"), ("is_alt", "<i4"), ("frac_rep", "<f4"), ("pos", "<i8"),
                        ("seed_off", "<i8")])
SEED_DTYPE = np.dtype([("rbeg", "<i8"), ("qbeg", "<i4"), ("len", "<i4"), ("score", "<i4"), ("pad_", "<i4")])
REG_DTYPE = np.dtype([("rb", "<i8"), ("re", "<i8"), ("qb", "<i4"), ("qe", "<i4"), ("rid", "<i4"),
                      ("score", "<i4"), ("truesc", "<i4"), ("w", "<i4"), ("seedcov", "<i4"),
                      ("seedlen0", "<i4"), ("chain", "<i4"), ("frac_rep", "<f4")])
assert CHAIN_DTYPE.itemsize == 48 and SEED_DTYPE.itemsize == 24 and REG_DTYPE.itemsize == 56


def ext_opt(a=1, b=4, o_del=6, e_del=1, o_ins=6, e_ins=1, pen_clip5=5, pen_clip3=5, w=100, zdrop=100) -> ExtOpt:
    """Defaults of mem_opt_init (reference src/bwamem.cpp:135-171)."""
    return ExtOpt(a, b, o_del, e_del, o_ins, e_ins, pen_clip5, pen_clip3, w, zdrop)


def chain_opt(w=100, max_chain_gap=10000, max_occ=500, min_seed_len=19, min_chain_weight=0,
              max_chain_extend=1 << 30, mask_level=0.5, drop_ratio=0.5) -> ChainOpt:
    """Defaults of mem_opt_init (reference src/bwamem.cpp:135-171)."""
    return ChainOpt(w, max_chain_gap, max_occ, min_seed_len, min_chain_weight, max_chain_extend, mask_level, drop_ratio)


class Extender:
    """Owns the reference-text replica of one GPU context."""

    def __init__(self, ref_string: np.ndarray, l_pac: int, contig_off, contig_len, ctx: Context | None = None,
                 device: int = 0):
        self._own = ctx is None
        self.ctx = ctx if ctx is not None else Context(device)
        ref_string = np.ascontiguousarray(ref_string, np.uint8)
        if len(ref_string) != 2 * l_pac:
            raise ValueError("ref_string must hold 2*l_pac bases (forward + reverse complement)")
        co, cl = np.ascontiguousarray(contig_off, np.int64), np.ascontiguousarray(contig_len, np.int32)
        self.ctx.check(self.ctx.lib.bwagpu_index_load_ref(self.ctx.h, ref_string.ctypes.data, int(l_pac), len(co),
                                                          co.ctypes.data, cl.ctypes.data, None), "bwagpu_index_load_ref")

    def close(self):
        if self._own and self.ctx is not None:
            self.ctx.close()
        self.ctx = None

    def mem_chain_seeds(self, opt: ChainOpt, smems: np.ndarray, coords: np.ndarray, ln: np.ndarray):
        """mem_chain_seeds + mem_chain_flt (reference src/bwamem.cpp:789-936, 528-646) for a batch:
        sorted SMEMs + their SA coordinates -> (chains, seeds) in the layout chain2aln takes."""
        smems, coords = np.ascontiguousarray(smems), np.ascontiguousarray(coords, np.int64)
        cap_c, cap_s = len(coords) + 1, len(coords) + 1
        chains, seeds = np.zeros(cap_c, CHAIN_DTYPE), np.zeros(cap_s, SEED_DTYPE)
        nc, ns = C.c_int64(0), C.c_int64(0)
        self.ctx.check(self.ctx.lib.bwagpu_chain_seeds(
            self.ctx.h, C.byref(opt), smems.ctypes.data, len(smems), coords.ctypes.data, len(coords), ln.ctypes.data,
            len(ln), chains.ctypes.data, cap_c, seeds.ctypes.data, cap_s, C.byref(nc), C.byref(ns)), "bwagpu_chain_seeds")
        return chains[:nc.value], seeds[:ns.value]

    def chain_seeds_device(self, opt: ChainOpt, seed_res, d_len, nseq, max_len) -> ChainResult:
        res = ChainResult()
        self.ctx.check(self.ctx.lib.bwagpu_chain_seeds_device(
            self.ctx.h, C.byref(opt), C.c_void_p(seed_res.d_smems), seed_res.n_smem, C.c_void_p(seed_res.d_smem_off),
            C.c_void_p(seed_res.d_coords), seed_res.n_coords, C.c_void_p(seed_res.d_coord_off), C.c_void_p(d_len),
            nseq, max_len, C.byref(res)), "bwagpu_chain_seeds_device")
        return res

    def mem_chain2aln_across_reads_V2(self, opt: ExtOpt, enc_reads, off, ln, chains, seeds) -> np.ndarray:
        chains, seeds = np.ascontiguousarray(chains), np.ascontiguousarray(seeds)
        if chains.dtype != CHAIN_DTYPE or seeds.dtype != SEED_DTYPE:
            raise ValueError("chains / seeds must use CHAIN_DTYPE / SEED_DTYPE")
        regs = np.zeros(len(seeds), REG_DTYPE)
        self.ctx.check(self.ctx.lib.bwagpu_chain2aln(
            self.ctx.h, C.byref(opt), enc_reads.ctypes.data, off.ctypes.data, ln.ctypes.data, len(ln),
            chains.ctypes.data, len(chains), seeds.ctypes.data, len(seeds), regs.ctypes.data), "bwagpu_chain2aln")
        return regs

    def chain2aln_device(self, opt: ExtOpt, d_reads, d_off, d_len, nseq, d_chains, n_chains, d_seeds, n_seeds) -> C2aResult:
        res = C2aResult()
        self.ctx.check(self.ctx.lib.bwagpu_chain2aln_device(
            self.ctx.h, C.byref(opt), C.c_void_p(d_reads), C.c_void_p(d_off), C.c_void_p(d_len), nseq,
            C.c_void_p(d_chains), n_chains, C.c_void_p(d_seeds), n_seeds, C.byref(res)), "bwagpu_chain2aln_device")
        return res


def mem_core_device(fmi, ext: Extender, sopt, copt: ChainOpt, eopt: ExtOpt, d_reads, d_off, d_len, nseq, max_len,
                    emf=None, d_emf=None):
    """reads -> alignment regions entirely on the device: [EMF probe (bwagpu_emf_probe_device),] seeding
    (bwagpu_seed_fm_device), chaining (bwagpu_chain_seeds_device) and extension (bwagpu_chain2aln_device)
    back to back on one context — the compute of worker_bwt + worker_aln (reference src/bwamem.cpp:
    1489-1543) up to the regions mem_chain2aln_across_reads_V2 leaves.  With `emf` (a PerfectMatch) and
    `d_emf` = (d_ret, d_perfect, d_skip) device buffers of nseq records, reads the filter answers are
    left out of everything downstream (bwamem.cpp:1245-1272).  Returns (seed_result, chain_result,
    c2a_result)."""
    d_skip = None
    if emf is not None:
        d_ret, d_perfect, d_skip = d_emf
        emf.probe_device(d_reads, d_off, d_len, nseq, d_ret, d_perfect, d_skip)
    sres = fmi.seed_device(sopt, d_reads, d_off, d_len, d_skip, nseq, max_len)
    cres = ext.chain_seeds_device(copt, sres, d_len, nseq, max_len)
    ares = ext.chain2aln_device(eopt, d_reads, d_off, d_len, nseq, cres.d_chains, cres.n_chains, cres.d_seeds, cres.n_seeds)
    return sres, cres, ares
