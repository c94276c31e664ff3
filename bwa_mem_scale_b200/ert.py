"""Host-side mirror of the reference's ERT seeding interface (seam S1').

`ErtIndex` holds what `index -a ert` writes (reference src/ertindex.cpp, src/bwtindex.cpp:278-305):
`<prefix>.kmer_table` = fmi->kmer_offsets (2^30 x 8 B) and `<prefix>.mlt_table` = fmi->mlt_table.
`ErtSeeding.seed` = the per-read seeding body of mem_kernel1_core_ert (reference src/bwamem.cpp:1124-1193:
get_seeds[_prefix] -> reseed[_prefix] -> last -> sort) behind `bwagpu_seed_ert`; `seed_device` is the
HBM-resident form that chaining (`bwagpu_chain_mems_device` = mem_chain_new + mem_chain_flt) follows.
The reference text must be loaded on the same context (Extender / bwagpu_index_load_ref).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import ChainOpt, ChainResult, Context, ErtResult, SeedOpt

ERT_MEM_DTYPE = np.dtype([("rid", "<i4"), ("start", "<i4"), ("end", "<i4"), ("hitcount", "<i4"), ("hit_off", "<i8")])
assert ERT_MEM_DTYPE.itemsize == 24
N_KMERS = 1 << 30
MAX_READ_LEN = 151            # READ_LEN, reference src/macro.h:68


class ErtIndex:
    """k-mer table (dense, or the non-zero/needed entries as (idx, val)) + tree blob."""

    def __init__(self, mlt_table: np.ndarray, kmer_offsets: np.ndarray | None = None,
                 sparse_idx: np.ndarray | None = None, sparse_val: np.ndarray | None = None):
        self.mlt_table = np.ascontiguousarray(mlt_table, np.uint8)
        self.kmer_offsets = kmer_offsets
        if kmer_offsets is not None and len(kmer_offsets) != N_KMERS:
            raise ValueError("kmer_offsets must hold 4^15 entries")
        self.sparse_idx = None if sparse_idx is None else np.ascontiguousarray(sparse_idx, np.uint32)
        self.sparse_val = None if sparse_val is None else np.ascontiguousarray(sparse_val, np.uint64)

    @classmethod
    def from_files(cls, prefix: str) -> "ErtIndex":
        """Map the reference's index files (the 8 GiB table is memory-mapped, not read)."""
        return cls(np.fromfile(prefix + ".mlt_table", np.uint8),
                   kmer_offsets=np.memmap(prefix + ".kmer_table", dtype=np.uint64, mode="r"))


class ErtSeeding:
    def __init__(self, index: ErtIndex, ctx: Context):
        self.ctx = ctx
        k = index.kmer_offsets
        if k is not None:
            self.ctx.check(self.ctx.lib.bwagpu_index_load_ert(self.ctx.h, k.ctypes.data, None, None, 0,
                                                              index.mlt_table.ctypes.data, index.mlt_table.nbytes),
                           "bwagpu_index_load_ert")
        else:
            n = 0 if index.sparse_idx is None else len(index.sparse_idx)
            self.ctx.check(self.ctx.lib.bwagpu_index_load_ert(
                self.ctx.h, None, index.sparse_idx.ctypes.data if n else None,
                index.sparse_val.ctypes.data if n else None, n, index.mlt_table.ctypes.data, index.mlt_table.nbytes),
                "bwagpu_index_load_ert")

    def complete_from_fm(self) -> None:
        """Fill in the entries of absent k-mers from the FM-index resident on the same context."""
        self.ctx.check(self.ctx.lib.bwagpu_index_complete_ert(self.ctx.h), "bwagpu_index_complete_ert")

    def dense_table(self) -> np.ndarray:
        """The resident 2^30-entry k-mer table as a host array (8 GiB)."""
        out = np.empty(N_KMERS, np.uint64)
        self.ctx.check(self.ctx.lib.bwagpu_index_get_ert_table(self.ctx.h, out.ctypes.data), "bwagpu_index_get_ert_table")
        return out

    def kmer_entries(self, idx: np.ndarray) -> np.ndarray:
        idx = np.ascontiguousarray(idx, np.uint32)
        out = np.zeros(len(idx), np.uint64)
        self.ctx.check(self.ctx.lib.bwagpu_index_get_ert_kmers(self.ctx.h, idx.ctypes.data, len(idx), out.ctypes.data),
                       "bwagpu_index_get_ert_kmers")
        return out

    def seed(self, opt: SeedOpt, enc_reads: np.ndarray, off: np.ndarray, ln: np.ndarray, skip: np.ndarray | None = None):
        """(mems, hits, mem_off): sorted MEMs of every read + the coordinates of all their hits."""
        n = len(ln)
        mem_cap, hit_cap = n * 16 + 64, n * 256 + 1024
        nm, nh = C.c_int64(0), C.c_int64(0)
        while True:
            mems, hits, mem_off = np.zeros(mem_cap, ERT_MEM_DTYPE), np.zeros(hit_cap, np.int64), np.zeros(n + 1, np.int64)
            rc = self.ctx.lib.bwagpu_seed_ert(
                self.ctx.h, C.byref(opt), enc_reads.ctypes.data, off.ctypes.data, ln.ctypes.data,
                None if skip is None else skip.ctypes.data, n, mems.ctypes.data, mem_cap, hits.ctypes.data, hit_cap,
                mem_off.ctypes.data, C.byref(nm), C.byref(nh))
            if rc == -5 and (nm.value > mem_cap or nh.value > hit_cap):
                mem_cap, hit_cap = max(mem_cap, nm.value), max(hit_cap, nh.value)
                continue
            self.ctx.check(rc, "bwagpu_seed_ert")
            return mems[:nm.value], hits[:nh.value], mem_off

    def seed_device(self, opt: SeedOpt, d_reads: int, d_off: int, d_len: int, d_skip: int | None, nseq: int,
                    max_len: int) -> ErtResult:
        res = ErtResult()
        self.ctx.check(self.ctx.lib.bwagpu_seed_ert_device(
            self.ctx.h, C.byref(opt), C.c_void_p(d_reads), C.c_void_p(d_off), C.c_void_p(d_len),
            C.c_void_p(d_skip) if d_skip else None, nseq, max_len, C.byref(res)), "bwagpu_seed_ert_device")
        return res

    def chain_mems_device(self, opt: ChainOpt, seed_res: ErtResult, d_len: int, nseq: int, max_len: int) -> ChainResult:
        """mem_chain_new + mem_chain_flt over the device-resident MEMs."""
        res = ChainResult()
        self.ctx.check(self.ctx.lib.bwagpu_chain_mems_device(
            self.ctx.h, C.byref(opt), C.c_void_p(seed_res.d_smems), seed_res.n_smem, C.c_void_p(seed_res.d_smem_off),
            C.c_void_p(seed_res.d_coords), seed_res.n_coords, C.c_void_p(seed_res.d_coord_off), C.c_void_p(d_len),
            nseq, max_len, C.byref(res)), "bwagpu_chain_mems_device")
        return res


def mem_core_ert_device(ert: ErtSeeding, ext, sopt, copt, eopt, d_reads, d_off, d_len, nseq, max_len, emf=None,
                        d_emf=None):
    """reads -> alignment regions on the device with ERT seeding: the compute of mem_kernel1_core_ert
    (reference src/bwamem.cpp:1052-1207) + mem_chain2aln_across_reads_V2.  Mirrors extend.mem_core_device."""
    d_skip = None
    if emf is not None:
        d_ret, d_perfect, d_skip = d_emf
        emf.probe_device(d_reads, d_off, d_len, nseq, d_ret, d_perfect, d_skip)
    sres = ert.seed_device(sopt, d_reads, d_off, d_len, d_skip, nseq, max_len)
    cres = ert.chain_mems_device(copt, sres, d_len, nseq, max_len)
    ares = ext.chain2aln_device(eopt, d_reads, d_off, d_len, nseq, cres.d_chains, cres.n_chains, cres.d_seeds, cres.n_seeds)
    return sres, cres, ares
