"""Host-side mirror of the reference's FM-index seeding interface (seam S1).

`FMI_search` keeps the reference's names for the two calls the pipeline makes on this path:
`collect_smem` = mem_collect_smem (reference src/bwamem.cpp:648-786, which drives
FMI_search::getSMEMsAllPosOneThread / getSMEMsOnePosOneThread / bwtSeedStrategyAllPosOneThread /
sortSMEMs) and `get_sa_entries_prefetch` (reference src/FMI_search.cpp:2261-2379).  The work is done
by the CUDA kernels behind `bwagpu_collect_smem` / `bwagpu_get_sa_entries` (include/bwagpu.h); the
index replica lives in the context's HBM.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .fmindex import FMIndex
from .lib import Context, FmIndexDesc, SeedOpt, SeedResult

# byte-for-byte mirror of SMEM (reference src/FMI_search.h:85-93)
SMEM_DTYPE = np.dtype([("rid", "<u4"), ("m", "<u4"), ("n", "<u4"), ("pad_", "<u4"),
                       ("k", "<i8"), ("l", "<i8"), ("s", "<i8")])
assert SMEM_DTYPE.itemsize == 40


def seed_opt(min_seed_len=19, split_factor=1.5, split_width=10, max_occ=500, max_mem_intv=20) -> SeedOpt:
    """Defaults of mem_opt_init (reference src/bwamem.cpp:135-171)."""
    return SeedOpt(min_seed_len, split_factor, split_width, max_occ, max_mem_intv)


def pack_reads(reads):
    """(n, L) uint8 matrix or list of arrays -> (flat u8, off i64, len i32): the enc_qdb layout."""
    if isinstance(reads, np.ndarray) and reads.ndim == 2:
        n, L = reads.shape
        ln = np.full(n, L, np.int32)
        flat = np.ascontiguousarray(reads, dtype=np.uint8).reshape(-1)
    else:
        ln = np.array([len(r) for r in reads], np.int32)
        flat = (np.concatenate([np.asarray(r, np.uint8) for r in reads]) if len(reads)
                else np.zeros(0, np.uint8))
    off = np.zeros(len(ln), np.int64)
    if len(ln):
        off[1:] = np.cumsum(ln[:-1], dtype=np.int64)
    return np.ascontiguousarray(flat), off, ln


class FMI_search:
    def __init__(self, index: FMIndex | None, ctx: Context | None = None, device: int = 0):
        """`index` = an FMIndex to upload (FMI_search::load_index), or None = build it in HBM from the reference
        text the context already holds (bwagpu_index_build_fm; load the text with Extender first)."""
        self._own = ctx is None
        self.ctx = ctx if ctx is not None else Context(device)
        self.index = index
        if index is None:
            self.ctx.check(self.ctx.lib.bwagpu_index_build_fm(self.ctx.h), "bwagpu_index_build_fm")
            return
        d = FmIndexDesc()
        d.reference_seq_len = index.ref_seq_len
        for i, v in enumerate(index.count_loaded()):
            d.count[i] = int(v)
        self._keep = (np.ascontiguousarray(index.cp_occ), np.ascontiguousarray(index.sa_ms_byte),
                      np.ascontiguousarray(index.sa_ls_word))
        d.cp_occ, d.sa_ms_byte, d.sa_ls_word = (a.ctypes.data for a in self._keep)
        d.sentinel_index = index.sentinel_index
        self.ctx.check(self.ctx.lib.bwagpu_index_load_fm(self.ctx.h, C.byref(d)), "bwagpu_index_load_fm")

    def get_index(self) -> FMIndex:
        """The resident FM-index copied back as an FMIndex (count[] converted to the file form)."""
        n, sent = C.c_int64(0), C.c_int64(0)
        count = np.zeros(5, np.int64)
        lib, h = self.ctx.lib, self.ctx.h
        self.ctx.check(lib.bwagpu_index_get_fm(h, C.byref(n), count.ctypes.data, None, None, None, C.byref(sent)), "bwagpu_index_get_fm")
        n_cp, n_sa = (n.value >> 6) + 1, (n.value >> 3) + 1
        cp, ms, ls = np.empty((n_cp, 8), np.int64), np.empty(n_sa, np.int8), np.empty(n_sa, np.uint32)
        self.ctx.check(lib.bwagpu_index_get_fm(h, None, None, cp.ctypes.data, ms.ctypes.data, ls.ctypes.data, None), "bwagpu_index_get_fm")
        return FMIndex(int(n.value), count - 1, cp, ms, ls, int(sent.value))

    # -- FMA "smem tables" (reference `smem-table` sub-command / FMI_search::load_smem_table,
    #    src/FMI_search.cpp:229-277, 394-410) -------------------------------------------------
    FMA_ALL_WORDS = (1 << 22) * 32      # all_smem_t[4^11], 128 B each
    FMA_LAST_WORDS = (1 << 26) * 4      # last_smem_t[4^13], 16 B each

    def build_fma(self) -> None:
        """Build both tables in HBM from the resident FM-index and switch the kernels to them."""
        self.ctx.check(self.ctx.lib.bwagpu_index_build_fma(self.ctx.h), "bwagpu_index_build_fma")

    def load_fma(self, all_smem: np.ndarray | None, last_smem: np.ndarray | None) -> None:
        """Upload tables read from <prefix>.all_smem.11 / <prefix>.last_smem.13."""
        for a, n in ((all_smem, self.FMA_ALL_WORDS), (last_smem, self.FMA_LAST_WORDS)):
            if a is not None and a.nbytes != n * 4:
                raise ValueError("FMA table has the wrong size")
        self.ctx.check(self.ctx.lib.bwagpu_index_load_fma(
            self.ctx.h, None if all_smem is None else all_smem.ctypes.data,
            None if last_smem is None else last_smem.ctypes.data), "bwagpu_index_load_fma")

    def get_fma(self) -> tuple[np.ndarray, np.ndarray]:
        a = np.empty(self.FMA_ALL_WORDS, np.uint32)
        b = np.empty(self.FMA_LAST_WORDS, np.uint32)
        self.ctx.check(self.ctx.lib.bwagpu_index_get_fma(self.ctx.h, a.ctypes.data, b.ctypes.data),
                       "bwagpu_index_get_fma")
        return a, b

    def use_fma(self, on: bool) -> None:
        self.ctx.check(self.ctx.lib.bwagpu_index_use_fma(self.ctx.h, 1 if on else 0), "bwagpu_index_use_fma")

    def close(self):
        if self._own and self.ctx is not None:
            self.ctx.close()
        self.ctx = None

    def collect_smem(self, opt: SeedOpt, enc_qdb: np.ndarray, off: np.ndarray, ln: np.ndarray,
                     skip: np.ndarray | None = None) -> np.ndarray:
        """All SMEMs of the batch in the reference's final order (rid, m, n)."""
        n = len(ln)
        cap = max(int(ln.sum()) // 4, 1024)
        tot = C.c_int64(0)
        while True:
            out = np.zeros(cap, SMEM_DTYPE)
            rc = self.ctx.lib.bwagpu_collect_smem(
                self.ctx.h, C.byref(opt), enc_qdb.ctypes.data, off.ctypes.data, ln.ctypes.data,
                None if skip is None else skip.ctypes.data, n, out.ctypes.data, cap, C.byref(tot))
            if rc == -5 and tot.value > cap:          # BWAGPU_ERR_RANGE: grow like the reference's realloc
                cap = int(tot.value)
                continue
            self.ctx.check(rc, "bwagpu_collect_smem")
            return out[:tot.value]

    def get_sa_entries_prefetch(self, smems: np.ndarray, max_occ: int) -> np.ndarray:
        smems = np.ascontiguousarray(smems)
        cap = int(np.minimum(smems["s"], max_occ).sum()) + 1
        coords = np.zeros(cap, np.int64)
        n = C.c_int64(0)
        self.ctx.check(self.ctx.lib.bwagpu_get_sa_entries(
            self.ctx.h, smems.ctypes.data, len(smems), max_occ, coords.ctypes.data, cap, C.byref(n)),
            "bwagpu_get_sa_entries")
        return coords[:n.value]

    def seed_device(self, opt: SeedOpt, d_reads: int, d_off: int, d_len: int, d_skip: int | None,
                    nseq: int, max_len: int) -> SeedResult:
        res = SeedResult()
        self.ctx.check(self.ctx.lib.bwagpu_seed_fm_device(
            self.ctx.h, C.byref(opt), C.c_void_p(d_reads), C.c_void_p(d_off), C.c_void_p(d_len),
            C.c_void_p(d_skip) if d_skip else None, nseq, max_len, C.byref(res)), "bwagpu_seed_fm_device")
        return res
