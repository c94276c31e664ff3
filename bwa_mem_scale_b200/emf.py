"""Host-side mirror of the reference's exact-match filter (EMF).

`PerfectTable` is the `<prefix>.perfect.<seed_len>` file of the reference's `perfect-index`
sub-command (reference src/perfect_index.cpp:1418-1424: 64-byte header of perfect_table_t,
src/perfect.h:175-199, then loc_table, then seed_table); `PerfectTable.build` is that sub-command
(host code in libbwagpu.so, same table bytes).  `PerfectMatch.find_perfect_match_entry` keeps the
reference's name (src/perfect_map.cpp:638-659) and runs the probe kernel behind `bwagpu_emf_probe`
for a whole batch of reads; there is no CPU probe here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from .lib import Context, EmfDesc, load

SEED_ENTRY_DTYPE = np.dtype([("flags", "<u4"), ("location", "<u4"), ("left", "<u4"), ("right", "<u4")])
PERFECT_DTYPE = np.dtype([("flags", "<u4"), ("location", "<u4")])          # bseq1_perfect_t
HEADER_BYTES = 64
NO_ENTRY = 0xFFFFFFFF
(FIND_PERFECT_NO_TABLE, FIND_PERFECT_WITH_N, FIND_PERFECT_NOT_MATCHED, FIND_PERFECT_FW_MATCHED,
 FIND_PERFECT_RC_MATCHED, FIND_PERFECT_SEED_ONLY_MATCHED) = range(6)
# header: int seed_len; u32 num_loc_entry, num_seed_entry, num_seed_load; 3 pointers; u32 seq_len,
# num_seed_used, num_seed_key; pad to 64 (packed struct, perfect.h:175-199)
_HDR = np.dtype([("seed_len", "<i4"), ("num_loc_entry", "<u4"), ("num_seed_entry", "<u4"), ("num_seed_load", "<u4"),
                 ("ptrs", "<u8", 3), ("seq_len", "<u4"), ("num_seed_used", "<u4"), ("num_seed_key", "<u4"),
                 ("pad", "u1", 12)])
assert _HDR.itemsize == HEADER_BYTES


class PerfectTable:
    def __init__(self, seed_len, seq_len, loc_table, seed_table, num_seed_used=0, num_seed_key=0):
        self.seed_len, self.seq_len = int(seed_len), int(seq_len)
        self.loc_table = np.ascontiguousarray(loc_table, np.uint32)
        self.seed_table = np.ascontiguousarray(seed_table, SEED_ENTRY_DTYPE)
        self.num_seed_used, self.num_seed_key = int(num_seed_used), int(num_seed_key)

    @property
    def num_seed_entry(self):
        return len(self.seed_table)

    @staticmethod
    def load(path: str) -> "PerfectTable":
        """load_perfect_table (reference src/perfect_map.cpp:204-291) without the shm store."""
        hdr = np.fromfile(path, _HDR, 1)[0]
        nl, ns = int(hdr["num_loc_entry"]), int(hdr["num_seed_entry"])
        loc = np.fromfile(path, np.uint32, nl, offset=HEADER_BYTES)
        seed = np.fromfile(path, SEED_ENTRY_DTYPE, ns, offset=HEADER_BYTES + 4 * nl)
        if len(loc) != nl or len(seed) != ns:
            raise ValueError(f"{path}: truncated perfect table")
        return PerfectTable(hdr["seed_len"], hdr["seq_len"], loc, seed, hdr["num_seed_used"], hdr["num_seed_key"])

    def save(self, path: str) -> None:
        hdr = np.zeros(1, _HDR)
        hdr["seed_len"], hdr["num_loc_entry"], hdr["num_seed_entry"] = self.seed_len, len(self.loc_table), len(self.seed_table)
        hdr["seq_len"], hdr["num_seed_used"], hdr["num_seed_key"] = self.seq_len, self.num_seed_used, self.num_seed_key
        with open(path, "wb") as f:
            f.write(hdr.tobytes())
            f.write(self.loc_table.tobytes())
            f.write(self.seed_table.tobytes())

    @staticmethod
    def build(ref_fwd: np.ndarray, seed_len: int = 151, slack: float = 1.1, contig_off=None, contig_len=None,
              hole_off=None, hole_len=None, n_threads: int = 0) -> "PerfectTable":
        """`bwa-mem2 perfect-index -l seed_len -s slack` on the forward text (default: one contig, no holes)."""
        lib = load()
        ref_fwd = np.ascontiguousarray(ref_fwd, np.uint8)
        co = np.ascontiguousarray([0] if contig_off is None else contig_off, np.int64)
        cl = np.ascontiguousarray([len(ref_fwd)] if contig_len is None else contig_len, np.int32)
        ho = np.ascontiguousarray([] if hole_off is None else hole_off, np.int64)
        hl = np.ascontiguousarray([] if hole_len is None else hole_len, np.int32)
        desc, p_loc, p_seed = EmfDesc(), C.c_void_p(), C.c_void_p()
        rc = lib.bwagpu_emf_build(ref_fwd.ctypes.data, len(ref_fwd), len(co), co.ctypes.data, cl.ctypes.data, len(ho),
                                  ho.ctypes.data if len(ho) else None, hl.ctypes.data if len(hl) else None,
                                  int(seed_len), float(slack), int(n_threads), C.byref(desc), C.byref(p_loc),
                                  C.byref(p_seed))
        if rc != 0:
            raise RuntimeError(f"bwagpu_emf_build failed with status {rc} (table full: slack < 1?)")
        try:
            loc = np.ctypeslib.as_array(C.cast(p_loc, C.POINTER(C.c_uint32)), (desc.num_loc_entry,)).copy()
            raw = np.ctypeslib.as_array(C.cast(p_seed, C.POINTER(C.c_uint32)), (desc.num_seed_entry * 4,)).copy()
        finally:
            lib.bwagpu_emf_free_tables(p_loc, p_seed)
        return PerfectTable(desc.seed_len, desc.seq_len, loc, raw.view(SEED_ENTRY_DTYPE), desc.num_seed_used,
                            desc.num_seed_key)


class PerfectMatch:
    """The table replica of one GPU context; the reference text must already be resident there
    (Extender / bwagpu_index_load_ref)."""

    def __init__(self, table: PerfectTable | None, ctx: Context, num_seed_load: int = 0):
        self.ctx = ctx
        self.table = table
        if table is None:
            self.ctx.check(self.ctx.lib.bwagpu_emf_unload(self.ctx.h), "bwagpu_emf_unload")
            return
        d = EmfDesc(table.seed_len, len(table.loc_table), len(table.seed_table), int(num_seed_load), table.seq_len,
                    table.num_seed_used, table.num_seed_key)
        self.ctx.check(self.ctx.lib.bwagpu_emf_load(self.ctx.h, C.byref(d), table.loc_table.ctypes.data,
                                                    table.seed_table.ctypes.data), "bwagpu_emf_load")

    @staticmethod
    def build_on_device(ctx: Context, seed_len: int = 151, slack: float = 1.1) -> "PerfectMatch":
        """bwagpu_emf_build_device: the table is built in HBM from the text and contig table the context holds."""
        pm = PerfectMatch.__new__(PerfectMatch)
        pm.ctx, pm.table = ctx, None
        ctx.check(ctx.lib.bwagpu_emf_build_device(ctx.h, int(seed_len), float(slack)), "bwagpu_emf_build_device")
        return pm

    def get_table(self) -> PerfectTable:
        """The resident table copied back to the host (bwagpu_emf_get)."""
        d = EmfDesc()
        self.ctx.check(self.ctx.lib.bwagpu_emf_get(self.ctx.h, C.byref(d), None, None), "bwagpu_emf_get")
        loc, seed = np.empty(d.num_loc_entry, np.uint32), np.empty(d.num_seed_entry, SEED_ENTRY_DTYPE)
        self.ctx.check(self.ctx.lib.bwagpu_emf_get(self.ctx.h, None, loc.ctypes.data, seed.ctypes.data), "bwagpu_emf_get")
        return PerfectTable(d.seed_len, d.seq_len, loc, seed, d.num_seed_used, d.num_seed_key)

    def unload(self):
        self.ctx.check(self.ctx.lib.bwagpu_emf_unload(self.ctx.h), "bwagpu_emf_unload")

    def find_perfect_match_entry(self, enc_reads: np.ndarray, off: np.ndarray, ln: np.ndarray):
        """(ret[nseq] FIND_PERFECT_* codes, perfect[nseq] {flags, location}, skip[nseq])."""
        n = len(ln)
        ret, perfect, skip = np.zeros(n, np.int32), np.zeros(n, PERFECT_DTYPE), np.zeros(n, np.uint8)
        enc_reads = np.ascontiguousarray(enc_reads, np.uint8)
        off, ln = np.ascontiguousarray(off, np.int64), np.ascontiguousarray(ln, np.int32)
        self.ctx.check(self.ctx.lib.bwagpu_emf_probe(self.ctx.h, enc_reads.ctypes.data, off.ctypes.data, ln.ctypes.data,
                                                     n, ret.ctypes.data, perfect.ctypes.data, skip.ctypes.data),
                       "bwagpu_emf_probe")
        return ret, perfect, skip

    def probe_device(self, d_reads, d_off, d_len, nseq, d_ret, d_perfect, d_skip=None) -> None:
        self.ctx.check(self.ctx.lib.bwagpu_emf_probe_device(
            self.ctx.h, C.c_void_p(d_reads), C.c_void_p(d_off), C.c_void_p(d_len), nseq, C.c_void_p(d_ret),
            C.c_void_p(d_perfect), C.c_void_p(d_skip) if d_skip else None), "bwagpu_emf_probe_device")
