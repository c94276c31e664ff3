"""ctypes bindings to the CPU oracle (oracle/_build/liboracle.so) and, when present, to the
reference itself compiled here (oracle/_ref/*.so).  TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")


class BswParams(C.Structure):
    _fields_ = [(n, C.c_int32) for n in
                ("o_del", "e_del", "o_ins", "e_ins", "zdrop", "end_bonus", "w_match", "w_mismatch")]


def default_params(**kw) -> BswParams:
    d = dict(o_del=6, e_del=1, o_ins=6, e_ins=1, zdrop=100, end_bonus=5, w_match=1, w_mismatch=4)
    d.update(kw)
    return BswParams(**d)


def build_oracle() -> str:
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "oracle"], check=True)
    return os.path.join(ORACLE_DIR, "_build", "liboracle.so")


_oracle = None


def oracle() -> C.CDLL:
    global _oracle
    if _oracle is None:
        lib = C.CDLL(build_oracle())
        lib.orc_bsw_batch.restype = C.c_uint64
        lib.orc_bsw_batch.argtypes = [C.POINTER(BswParams), C.c_int, C.c_void_p, C.c_int32,
                                      C.c_void_p, C.c_void_p, C.c_int32]
        _oracle = lib
    return _oracle


def ref_bsw(isa: str = "sse41"):
    path = os.path.join(ORACLE_DIR, "_ref", f"libref_bsw_{isa}.so")
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    lib.ref_bsw_run.restype = C.c_int
    lib.ref_bsw_run.argtypes = [C.POINTER(BswParams), C.c_int, C.c_void_p, C.c_int32, C.c_void_p,
                                C.c_int64, C.c_void_p, C.c_int64, C.c_int32]
    return lib


def oracle_bsw(params: BswParams, kind: int, pairs: np.ndarray, ref: np.ndarray, qer: np.ndarray,
               w: int) -> tuple[np.ndarray, int]:
    out = pairs.copy()
    cells = oracle().orc_bsw_batch(C.byref(params), kind, out.ctypes.data, len(out),
                                   ref.ctypes.data, qer.ctypes.data, w)
    return out, int(cells)


def reference_bsw(lib, params: BswParams, kind: int, pairs: np.ndarray, ref: np.ndarray,
                  qer: np.ndarray, w: int) -> np.ndarray:
    out = pairs.copy()
    lib.ref_bsw_run(C.byref(params), kind, out.ctypes.data, len(out), ref.ctypes.data, len(ref),
                    qer.ctypes.data, len(qer), w)
    return out


# ---- FM-index seeding -------------------------------------------------------------------------
# byte-for-byte mirror of SMEM (reference src/FMI_search.h:85-93)
SMEM_DTYPE = np.dtype([("rid", "<u4"), ("m", "<u4"), ("n", "<u4"), ("pad_", "<u4"),
                       ("k", "<i8"), ("l", "<i8"), ("s", "<i8")])
assert SMEM_DTYPE.itemsize == 40


class OrcFmi(C.Structure):
    _fields_ = [("ref_seq_len", C.c_int64), ("count", C.c_int64 * 5), ("cp_occ", C.c_void_p),
                ("sa_ms_byte", C.c_void_p), ("sa_ls_word", C.c_void_p),
                ("sentinel_index", C.c_int64), ("fma_all", C.c_void_p), ("fma_last", C.c_void_p)]


class OrcSeedOpt(C.Structure):
    _fields_ = [("min_seed_len", C.c_int32), ("split_factor", C.c_float),
                ("split_width", C.c_int32), ("max_mem_intv", C.c_int64)]


def default_seed_opt(**kw) -> OrcSeedOpt:
    d = dict(min_seed_len=19, split_factor=1.5, split_width=10, max_mem_intv=20)
    d.update(kw)
    return OrcSeedOpt(**d)


def orc_fmi(idx, fma=None) -> OrcFmi:
    """idx: bwa_mem_scale_b200.fmindex.FMIndex (arrays must stay alive while the struct is used).
    fma: optional (all_smem u32[4^11*32], last_smem u32[4^13*4]) tables; either may be None."""
    f = OrcFmi()
    if fma is not None:
        f.fma_all = fma[0].ctypes.data if fma[0] is not None else None
        f.fma_last = fma[1].ctypes.data if fma[1] is not None else None
    f.ref_seq_len = idx.ref_seq_len
    for i, v in enumerate(idx.count_loaded()):
        f.count[i] = int(v)
    f.cp_occ = idx.cp_occ.ctypes.data
    f.sa_ms_byte = idx.sa_ms_byte.ctypes.data
    f.sa_ls_word = idx.sa_ls_word.ctypes.data
    f.sentinel_index = idx.sentinel_index
    return f


def pack_reads(reads, lens=None):
    """(n, L) matrix or list of arrays -> (flat u8, off i64, len i32)."""
    if isinstance(reads, np.ndarray) and reads.ndim == 2:
        n, L = reads.shape
        ln = np.full(n, L, np.int32) if lens is None else np.asarray(lens, np.int32)
        rows = [reads[i, :ln[i]] for i in range(n)] if lens is not None else None
        flat = np.ascontiguousarray(reads).reshape(-1) if rows is None else np.concatenate(rows)
    else:
        ln = np.array([len(r) for r in reads], np.int32)
        flat = np.concatenate([np.asarray(r, np.uint8) for r in reads]) if len(reads) else np.zeros(0, np.uint8)
    off = np.zeros(len(ln), np.int64)
    if len(ln):
        off[1:] = np.cumsum(ln[:-1])
    return np.ascontiguousarray(flat, dtype=np.uint8), off, ln


def oracle_fma_entries(idx, which: str, indices) -> np.ndarray:
    """Entries of the all (32 words) / last (4 words) FMA table for the given k-mer ranks."""
    lib = oracle()
    f = orc_fmi(idx)
    w = 32 if which == "all" else 4
    fn = lib.orc_fma_entry_all if which == "all" else lib.orc_fma_entry_last
    fn.restype = None
    fn.argtypes = [C.POINTER(OrcFmi), C.c_uint32, C.c_void_p]
    out = np.zeros((len(indices), w), np.uint32)
    for r, i in enumerate(indices):
        fn(C.byref(f), int(i), out[r].ctypes.data)
    return out


def oracle_fma_sparse(idx, reads):
    """(all, last) FMA tables holding only the entries `reads` (list of code arrays) can look up:
    every 11-mer / 13-mer window prefix the seeding code indexes with (N-free part, zero-padded, as
    src/FMI_search.cpp:1421-1427 builds the index).  Untouched entries stay zero (lazily-committed
    pages), so this is cheap; contents of the touched entries equal the full tables'."""
    lib = oracle()
    lib.orc_fma_fill.restype = None
    lib.orc_fma_fill.argtypes = [C.POINTER(OrcFmi), C.c_int, C.c_void_p, C.c_int64, C.c_void_p]
    f = orc_fmi(idx)
    ia, il = set(), set()
    for r in reads:
        r = np.asarray(r, np.int64)
        n = len(r)
        for x in range(n):
            if n - x >= 11 and r[x] < 4:
                v = 0
                for k in range(11):
                    if r[x + k] >= 4:
                        break
                    v |= int(r[x + k]) << ((10 - k) * 2)
                ia.add(v)
            if n - x >= 13 and r[x:x + 13].max() < 4:
                v = 0
                for k in range(13):
                    v |= int(r[x + k]) << ((12 - k) * 2)
                il.add(v)
    a = np.zeros((1 << 22) * 32, np.uint32)
    b = np.zeros((1 << 26) * 4, np.uint32)
    ia = np.array(sorted(ia), np.uint32)
    il = np.array(sorted(il), np.uint32)
    lib.orc_fma_fill(C.byref(f), 0, ia.ctypes.data, len(ia), a.ctypes.data)
    lib.orc_fma_fill(C.byref(f), 1, il.ctypes.data, len(il), b.ctypes.data)
    return a, b


def oracle_fma_build(idx, want_all=True, want_last=True):
    """Both FMA tables from the oracle's builder (slow: ~1 minute for the pair)."""
    lib = oracle()
    lib.orc_fma_build.restype = None
    lib.orc_fma_build.argtypes = [C.POINTER(OrcFmi), C.c_void_p, C.c_void_p]
    f = orc_fmi(idx)
    a = np.zeros((1 << 22) * 32, np.uint32) if want_all else None
    b = np.zeros((1 << 26) * 4, np.uint32) if want_last else None
    lib.orc_fma_build(C.byref(f), a.ctypes.data if want_all else None, b.ctypes.data if want_last else None)
    return a, b


def oracle_collect_smem(idx, opt: OrcSeedOpt, flat, off, ln, skip=None, fma=None):
    lib = oracle()
    lib.orc_collect_smem.restype = C.c_int64
    lib.orc_collect_smem.argtypes = [C.POINTER(OrcFmi), C.POINTER(OrcSeedOpt), C.c_void_p,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                     C.c_int64, C.POINTER(C.c_int64)]
    f = orc_fmi(idx, fma)
    cap = max(int(ln.sum()) * 4, 1024)
    n_ext = C.c_int64(0)
    while True:
        out = np.zeros(cap, SMEM_DTYPE)
        n_ext.value = 0
        tot = lib.orc_collect_smem(C.byref(f), C.byref(opt), flat.ctypes.data, off.ctypes.data,
                                   ln.ctypes.data, None if skip is None else skip.ctypes.data,
                                   len(ln), out.ctypes.data, cap, C.byref(n_ext))
        if tot <= cap:
            return out[:tot].copy(), int(n_ext.value)
        cap = tot


def oracle_sa_lookup(idx, smems, max_occ=500):
    lib = oracle()
    lib.orc_sa_lookup.restype = C.c_int64
    lib.orc_sa_lookup.argtypes = [C.POINTER(OrcFmi), C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                  C.POINTER(C.c_int64)]
    f = orc_fmi(idx)
    cap = int(np.minimum(smems["s"], max_occ).sum()) + 1
    coords = np.zeros(cap, np.int64)
    n_lf = C.c_int64(0)
    smems = np.ascontiguousarray(smems)
    n = lib.orc_sa_lookup(C.byref(f), smems.ctypes.data, len(smems), max_occ, coords.ctypes.data,
                          C.byref(n_lf))
    return coords[:n].copy(), int(n_lf.value)


def ref_binary(flavor="plain", isa="sse41"):
    p = os.path.join(ORACLE_DIR, "_ref", f"bwa-mem2.{flavor}.{isa}")
    return p if os.path.exists(p) else None


def ref_api(flavor="plain", isa="sse41"):
    path = os.path.join(ORACLE_DIR, "_ref", f"libref_api_{flavor}_{isa}.so")
    if not os.path.exists(path):
        return None
    lib = C.CDLL(path)
    lib.ref_fmi_open.restype = C.c_void_p
    lib.ref_fmi_open.argtypes = [C.c_char_p]
    lib.ref_fmi_close.argtypes = [C.c_void_p]
    lib.ref_collect_smem.restype = C.c_int64
    lib.ref_collect_smem.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_int64,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p,
                                     C.c_int64]
    lib.ref_sa_lookup.restype = C.c_int64
    lib.ref_sa_lookup.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p]
    return lib


def ref_index(prefix: str, fasta: str, flavor="plain", isa="sse41") -> None:
    """Run the reference's own `index` sub-command (writes <prefix>.bwt.2bit.64, .0123, .ann ...)."""
    subprocess.run([ref_binary(flavor, isa), "index", "-p", prefix, fasta], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


def reference_collect_smem(lib, h, opt: OrcSeedOpt, flat, off, ln):
    cap = max(int(ln.sum()) * 4, 1024)
    out = np.zeros(cap, SMEM_DTYPE)
    # the reference rewrites nothing in `flat` here (mem_collect_smem only reads seq)
    tot = lib.ref_collect_smem(h, opt.min_seed_len, opt.split_factor, opt.split_width,
                               opt.max_mem_intv, flat.ctypes.data, off.ctypes.data, ln.ctypes.data,
                               len(ln), out.ctypes.data, cap)
    assert tot <= cap
    return out[:tot].copy()


def reference_sa_lookup(lib, h, smems, max_occ=500):
    cap = int(np.minimum(smems["s"], max_occ).sum()) + 1
    coords = np.zeros(cap, np.int64)
    smems = np.ascontiguousarray(smems)
    n = lib.ref_sa_lookup(h, smems.ctypes.data, len(smems), max_occ, coords.ctypes.data)
    return coords[:n].copy()


# ---- mate-rescue local SW (ksw_align2) -----------------------------------------------------------
KSWR_DTYPE = np.dtype([(n, "<i4") for n in ("score", "te", "qe", "score2", "te2", "tb", "qb")])
KSW_XBYTE, KSW_XSTOP, KSW_XSUBO, KSW_XSTART = 0x10000, 0x20000, 0x40000, 0x80000


def score_matrix(a=1, b=4) -> np.ndarray:
    """bwa_fill_scmat (reference src/bwa.cpp:368-378): 5x5, ambiguous = -1."""
    m = np.full((5, 5), -1, np.int8)
    for i in range(4):
        for j in range(4):
            m[i, j] = a if i == j else -b
    return np.ascontiguousarray(m.reshape(-1))


def oracle_ksw(pairs, ref, qer, mat=None, o_del=6, e_del=1, o_ins=6, e_ins=1):
    lib = oracle()
    mat = score_matrix() if mat is None else mat
    aln = np.zeros(len(pairs), KSWR_DTYPE)
    lib.orc_ksw_batch.restype = None
    lib.orc_ksw_batch.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                  C.c_int, C.c_int, C.c_void_p]
    pairs = np.ascontiguousarray(pairs)
    lib.orc_ksw_batch(pairs.ctypes.data, len(pairs), ref.ctypes.data, qer.ctypes.data, mat.ctypes.data,
                      o_del, e_del, o_ins, e_ins, aln.ctypes.data)
    return aln


def ref_ksw():
    path = os.path.join(ORACLE_DIR, "_ref", "libref_ksw.so")
    return C.CDLL(path) if os.path.exists(path) else None


def reference_ksw(lib, pairs, ref, qer, mat=None, o_del=6, e_del=1, o_ins=6, e_ins=1):
    mat = score_matrix() if mat is None else mat
    aln = np.zeros(len(pairs), KSWR_DTYPE)
    lib.ref_ksw_batch.restype = None
    lib.ref_ksw_batch.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                  C.c_int, C.c_int, C.c_void_p]
    pairs = np.ascontiguousarray(pairs)
    lib.ref_ksw_batch(pairs.ctypes.data, len(pairs), ref.ctypes.data, qer.ctypes.data, mat.ctypes.data,
                      o_del, e_del, o_ins, e_ins, aln.ctypes.data)
    return aln


# ---- chaining + extension (mem_kernel1_core / mem_chain2aln_across_reads_V2 taps) ------------------
CHAIN_DTYPE = np.dtype([("seqid", "<i4"), ("rid", "<i4"), ("n", "<i4"), ("w", "<i4"), ("kept", "<i4"),
                        ("first", "<i4"), ("is_alt", "<i4"), ("frac_rep", "<f4"), ("pos", "<i8"),
                        ("seed_off", "<i8")])
SEED_DTYPE = np.dtype([("rbeg", "<i8"), ("qbeg", "<i4"), ("len", "<i4"), ("score", "<i4"), ("pad_", "<i4")])
REG_DTYPE = np.dtype([("rb", "<i8"), ("re", "<i8"), ("qb", "<i4"), ("qe", "<i4"), ("rid", "<i4"),
                      ("score", "<i4"), ("truesc", "<i4"), ("w", "<i4"), ("seedcov", "<i4"),
                      ("seedlen0", "<i4"), ("chain", "<i4"), ("frac_rep", "<f4")])
assert CHAIN_DTYPE.itemsize == 48 and SEED_DTYPE.itemsize == 24 and REG_DTYPE.itemsize == 56


def mem_opt(**kw):
    """Defaults of mem_opt_init (reference src/bwamem.cpp:135-171) for the fields the taps take."""
    d = dict(a=1, b=4, o_del=6, e_del=1, o_ins=6, e_ins=1, pen_clip5=5, pen_clip3=5, w=100, zdrop=100,
             min_seed_len=19, max_occ=500, max_chain_gap=10000, min_chain_weight=0, max_chain_extend=1 << 30,
             split_width=10, split_factor=1.5, mask_level=0.5, drop_ratio=0.5, max_mem_intv=20)
    d.update(kw)
    return d


def opt_arrays(o):
    oi = np.array([o[k] for k in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5", "pen_clip3", "w",
                                  "zdrop", "min_seed_len", "max_occ", "max_chain_gap", "min_chain_weight",
                                  "max_chain_extend", "split_width")], np.int32)
    of = np.array([o["split_factor"], o["mask_level"], o["drop_ratio"]], np.float32)
    return oi, of


class RefMem:
    """The reference's own seeding -> chaining -> extension core on an index written to disk."""

    def __init__(self, prefix: str, flavor="plain", isa="sse41", _parent=None):
        path = os.path.join(ORACLE_DIR, "_ref", f"libref_api_{flavor}_{isa}.so")
        self.lib = C.CDLL(path)
        self.lib.ref_mem_clone.restype = C.c_void_p
        self.lib.ref_mem_clone.argtypes = [C.c_void_p]
        self.lib.ref_mem_open.restype = C.c_void_p
        self.lib.ref_mem_open.argtypes = [C.c_char_p]
        self.lib.ref_mem_l_pac.restype = C.c_int64
        self.lib.ref_mem_l_pac.argtypes = [C.c_void_p]
        self.lib.ref_mem_n_seqs.argtypes = [C.c_void_p]
        self.lib.ref_mem_contigs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self.lib.ref_mem_core.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]
        self.h = self.lib.ref_mem_open(prefix.encode()) if _parent is None else self.lib.ref_mem_clone(_parent.h)
        self._args = (prefix, flavor, isa)
        self.l_pac = self.lib.ref_mem_l_pac(self.h)
        n = self.lib.ref_mem_n_seqs(self.h)
        self.contig_off, self.contig_len = np.zeros(n, np.int64), np.zeros(n, np.int32)
        self.lib.ref_mem_contigs(self.h, self.contig_off.ctypes.data, self.contig_len.ctypes.data)

    @classmethod
    def from_arrays(cls, idx, ref_string, contig_off, contig_len, flavor="plain", isa="sse41"):
        """A worker over an index already in memory (idx: bwa_mem_scale_b200.FMIndex with host arrays; ref_string: the
        .0123 text).  Nothing is written to disk; the arrays must stay alive as long as the worker (kept here)."""
        self = cls.__new__(cls)
        path = os.path.join(ORACLE_DIR, "_ref", f"libref_api_{flavor}_{isa}.so")
        self.lib = C.CDLL(path)
        self.lib.ref_mem_clone.restype = C.c_void_p
        self.lib.ref_mem_clone.argtypes = [C.c_void_p]
        self.lib.ref_mem_open_arrays.restype = C.c_void_p
        self.lib.ref_mem_open_arrays.argtypes = [C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                                 C.c_void_p, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
        self.lib.ref_mem_core.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_void_p]
        count = np.ascontiguousarray(idx.count_loaded(), np.int64)
        co, cl = np.ascontiguousarray(contig_off, np.int64), np.ascontiguousarray(contig_len, np.int32)
        cp, ms, ls = (np.ascontiguousarray(a) for a in (idx.cp_occ, idx.sa_ms_byte, idx.sa_ls_word))
        ref_string = np.ascontiguousarray(ref_string, np.uint8)
        self._keep = (count, co, cl, cp, ms, ls, ref_string)
        self.h = self.lib.ref_mem_open_arrays(idx.ref_seq_len, count.ctypes.data, cp.ctypes.data, ms.ctypes.data,
                                              ls.ctypes.data, idx.sentinel_index, ref_string.ctypes.data, idx.l_pac,
                                              len(co), co.ctypes.data, cl.ctypes.data)
        self._args = None
        self.l_pac, self.contig_off, self.contig_len = idx.l_pac, co, cl
        return self

    def set_fma(self, all_smem, last_smem):
        """FMA tables already in memory (accel / emf / scale flavours); shared by every clone of this worker."""
        self.lib.ref_mem_set_fma.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self._fma = (all_smem, last_smem)
        self.lib.ref_mem_set_fma(self.h, None if all_smem is None else all_smem.ctypes.data,
                                 None if last_smem is None else last_smem.ctypes.data)

    def set_perfect(self, table):
        """EMF table already in memory (emf / scale flavours): a bwa_mem_scale_b200.PerfectTable, or None = off."""
        self.lib.ref_mem_set_perfect.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        self._emf = table
        if table is None:
            self.lib.ref_mem_set_perfect(self.h, 0, 0, 0, 0, None, None)
        else:
            self.lib.ref_mem_set_perfect(self.h, table.seed_len, len(table.loc_table), len(table.seed_table), table.seq_len,
                                         table.loc_table.ctypes.data, table.seed_table.ctypes.data)

    def use_ert(self, ref_ert):
        """Switch seeding + chaining to mem_kernel1_core_ert over the tables of a RefErt (None: back to FM)."""
        self.lib.ref_mem_use_ert.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        if ref_ert is None:
            self.lib.ref_mem_use_ert(self.h, None, None)
        else:
            n = C.c_int64(0)
            self.lib.ref_mem_use_ert(self.h, ref_ert.lib.ref_ert_kmer(ref_ert.h), ref_ert.lib.ref_ert_mlt(ref_ert.h, C.byref(n)))
        self._ert = ref_ert

    def use_ert_tables(self, kmer: np.ndarray, mlt: np.ndarray):
        """Same with the tables given as host arrays (dense 2^30-entry k-mer table, tree blob)."""
        self.lib.ref_mem_use_ert.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        self._ert = (kmer, mlt)
        self.lib.ref_mem_use_ert(self.h, kmer.ctypes.data, mlt.ctypes.data)

    def clone(self):
        """Another worker on the same loaded index (own scratch arenas), for multi-threaded baselines."""
        if self._args is None:                      # opened from arrays
            w = RefMem.__new__(RefMem)
            w.lib, w._args, w._keep = self.lib, None, self._keep
            w.h = self.lib.ref_mem_clone(self.h)
            w.l_pac, w.contig_off, w.contig_len = self.l_pac, self.contig_off, self.contig_len
            return w
        return RefMem(*self._args, _parent=self)

    def core(self, opt, flat, off, ln):
        oi, of = opt_arrays(opt)
        cap = np.array([len(ln) * 64 + 1024, len(ln) * 256 + 4096, len(ln) * 256 + 4096], np.int64)
        while True:
            chains, seeds, regs = np.zeros(cap[0], CHAIN_DTYPE), np.zeros(cap[1], SEED_DTYPE), np.zeros(cap[2], REG_DTYPE)
            reg_off, n_out = np.zeros(len(ln) + 1, np.int64), np.zeros(3, np.int64)
            self.lib.ref_mem_core(self.h, oi.ctypes.data, of.ctypes.data, int(opt["max_mem_intv"]), flat.ctypes.data,
                                  off.ctypes.data, ln.ctypes.data, len(ln), chains.ctypes.data, seeds.ctypes.data,
                                  regs.ctypes.data, reg_off.ctypes.data, cap.ctypes.data, n_out.ctypes.data)
            if np.all(n_out <= cap):
                return chains[:n_out[0]].copy(), seeds[:n_out[1]].copy(), regs[:n_out[2]].copy(), reg_off
            cap = np.maximum(cap, n_out + 16)


class OrcExtOpt(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5", "pen_clip3",
                                         "w", "zdrop")]


def ext_opt(o) -> OrcExtOpt:
    return OrcExtOpt(*[o[k] for k in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5", "pen_clip3", "w", "zdrop")])


def oracle_chain2aln(opt, l_pac, contig_off, contig_len, ref_string, flat, off, ln, chains, seeds):
    """mem_chain2aln_across_reads_V2 restated (oracle/c2a_oracle.c): regs[len(seeds)] in creation order."""
    lib = oracle()
    lib.orc_chain2aln.restype = None
    lib.orc_chain2aln.argtypes = [C.POINTER(OrcExtOpt), C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p,
                                  C.c_void_p]
    regs = np.zeros(len(seeds), REG_DTYPE)
    eo = ext_opt(opt)
    chains, seeds = np.ascontiguousarray(chains), np.ascontiguousarray(seeds)
    contig_off, contig_len = np.ascontiguousarray(contig_off, np.int64), np.ascontiguousarray(contig_len, np.int32)
    lib.orc_chain2aln(C.byref(eo), int(l_pac), contig_off.ctypes.data, contig_len.ctypes.data, ref_string.ctypes.data,
                      flat.ctypes.data, off.ctypes.data, ln.ctypes.data, len(ln), chains.ctypes.data, len(chains),
                      seeds.ctypes.data, regs.ctypes.data, None)
    return regs


# ---- EMF (exact-match filter) ----------------------------------------------------------------
class OrcEmfTable(C.Structure):
    _fields_ = [("seed_len", C.c_int32), ("num_loc_entry", C.c_uint32), ("num_seed_entry", C.c_uint32),
                ("num_seed_load", C.c_uint32), ("seq_len", C.c_uint32), ("ref_string", C.c_void_p),
                ("loc_table", C.c_void_p), ("seed_table", C.c_void_p)]


def oracle_emf_probe(table, ref_string, flat, off, ln, num_seed_load=0):
    """find_perfect_match_entry restated (oracle/emf_oracle.c) over a bwa_mem_scale_b200.PerfectTable
    (or None = no table): (ret codes, perfect[n][2] = flags, location)."""
    lib = oracle()
    lib.orc_emf_probe.restype = None
    lib.orc_emf_probe.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]
    ret, perfect = np.zeros(len(ln), np.int32), np.zeros((len(ln), 2), np.uint32)
    flat = np.ascontiguousarray(flat, np.uint8)
    ref_string = np.ascontiguousarray(ref_string, np.uint8)
    if table is None:
        tp = None
    else:
        ns = len(table.seed_table)
        t = OrcEmfTable(table.seed_len, len(table.loc_table), ns, num_seed_load if 0 < num_seed_load < ns else ns,
                        table.seq_len, ref_string.ctypes.data, table.loc_table.ctypes.data, table.seed_table.ctypes.data)
        tp = C.addressof(t)
    lib.orc_emf_probe(tp, flat.ctypes.data, off.ctypes.data, ln.ctypes.data, len(ln), ret.ctypes.data, perfect.ctypes.data)
    return ret, perfect


def write_emf_inputs(prefix, ref_fwd, contig_off, contig_len, hole_off, hole_len):
    """The three files the reference's `perfect-index` reads (src/perfect_index.cpp:113-213, 1439-1462):
    <prefix>.ann, <prefix>.amb (text, bns_dump format src/bntseq.cpp:81-131) and <prefix>.0123."""
    n = len(ref_fwd)
    with open(prefix + ".ann", "w") as f:
        f.write(f"{n} {len(contig_off)} 11\n")
        for i, (o, l) in enumerate(zip(contig_off, contig_len)):
            na = sum(1 for h in hole_off if o <= h < o + l)
            f.write(f"0 chr{i + 1} (null)\n{int(o)} {int(l)} {na}\n")
    with open(prefix + ".amb", "w") as f:
        f.write(f"{n} {len(contig_off)} {len(hole_off)}\n")
        for o, l in zip(hole_off, hole_len):
            f.write(f"{int(o)} {int(l)} N\n")
    t = np.concatenate([ref_fwd, (3 - ref_fwd[::-1])]).astype(np.uint8)
    t.tofile(prefix + ".0123")
    return t


def ref_perfect_index(prefix, seed_len, slack=1.1, flavor="emf", isa="sse41"):
    """Run the reference's own `perfect-index` sub-command: writes <prefix>.perfect.<seed_len>."""
    subprocess.run([ref_binary(flavor, isa), "perfect-index", "-l", str(seed_len), "-s", str(slack), prefix],
                   check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return f"{prefix}.perfect.{seed_len}"


class RefEmf:
    """The reference's own find_perfect_match_entry over a table file (oracle/refwrap/ref_api_wrap.cpp)."""

    def __init__(self, table_path, ref0123_path, flavor="emf", isa="sse41", num_seed_load=0):
        self.lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", f"libref_api_{flavor}_{isa}.so"))
        self.lib.ref_emf_open.restype = C.c_void_p
        self.lib.ref_emf_open.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32]
        self.lib.ref_emf_close.argtypes = [C.c_void_p]
        self.lib.ref_emf_probe.argtypes = [C.c_void_p] * 4 + [C.c_int32, C.c_void_p, C.c_void_p]
        self.h = self.lib.ref_emf_open(table_path.encode(), ref0123_path.encode(), num_seed_load)
        if not self.h:
            raise OSError(f"cannot open {table_path}")

    def probe(self, flat, off, ln):
        ret, perfect = np.zeros(len(ln), np.int32), np.zeros((len(ln), 2), np.uint32)
        flat = np.ascontiguousarray(flat, np.uint8).copy()
        self.lib.ref_emf_probe(self.h, flat.ctypes.data, off.ctypes.data, ln.ctypes.data, len(ln), ret.ctypes.data,
                               perfect.ctypes.data)
        return ret, perfect

    def close(self):
        self.lib.ref_emf_close(self.h)
        self.h = None


# ---- ERT seeding (reference src/ertseeding.cpp through oracle/refwrap/ref_ert_wrap.cpp) ------------------
ERT_MEM_DTYPE = np.dtype([("rid", "<i4"), ("start", "<i4"), ("end", "<i4"), ("hitcount", "<i4"), ("hit_off", "<i8")])


class RefErt:
    """The reference's own ERT seeding on an index written by its `index -a ert` (files mmap'ed)."""

    def __init__(self, prefix, flavor="plain", isa="sse41"):
        self.lib = C.CDLL(os.path.join(ORACLE_DIR, "_ref", f"libref_api_{flavor}_{isa}.so"))
        self.lib.ref_ert_open.restype = C.c_void_p
        self.lib.ref_ert_open.argtypes = [C.c_char_p]
        self.lib.ref_ert_close.argtypes = [C.c_void_p]
        self.lib.ref_ert_kmer.restype = C.c_void_p
        self.lib.ref_ert_kmer.argtypes = [C.c_void_p]
        self.lib.ref_ert_mlt.restype = C.c_void_p
        self.lib.ref_ert_mlt.argtypes = [C.c_void_p, C.c_void_p]
        self.lib.ref_ert_tables.restype = C.c_void_p
        self.lib.ref_ert_seed.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                          C.c_void_p, C.c_void_p]
        self.lib.ref_ert_open_mem.restype = C.c_void_p
        self.lib.ref_ert_open_mem.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
        if isinstance(prefix, tuple):          # (dense k-mer table, tree blob, .0123 text) as host arrays
            self._keep = tuple(np.ascontiguousarray(a) for a in prefix)
            k, m, r = self._keep
            self.h = self.lib.ref_ert_open_mem(k.ctypes.data, m.ctypes.data, m.nbytes, r.ctypes.data, r.nbytes)
        else:
            self.h = self.lib.ref_ert_open(prefix.encode())
        assert self.h, f"cannot map the ERT index {prefix}"

    def kmer_table(self) -> np.ndarray:
        """The 2^30-entry k-mer table as a (lazy, mmap-backed) array."""
        p = self.lib.ref_ert_kmer(self.h)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint64)), shape=(1 << 30,))

    def mlt_table(self) -> np.ndarray:
        n = C.c_int64(0)
        p = self.lib.ref_ert_mlt(self.h, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value,))

    def tables(self):
        """(leaf_table, code_table) of src/ertseeding.cpp:25-430 as [3][256][4] arrays."""
        out = []
        for which in (0, 1):
            p = self.lib.ref_ert_tables(which)
            out.append(np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(3, 256, 4)).copy())
        return out

    def seed(self, flat, off, ln, min_seed_len=19, split_factor=1.5, split_width=10, max_mem_intv=20):
        mem_cap, hit_cap = len(ln) * 16 + 64, len(ln) * 256 + 1024
        while True:
            mems, hits = np.zeros(mem_cap, ERT_MEM_DTYPE), np.zeros(hit_cap, np.int64)
            mem_off, n_out = np.zeros(len(ln) + 1, np.int64), np.zeros(3, np.int64)
            self.lib.ref_ert_seed(self.h, min_seed_len, split_factor, split_width, max_mem_intv, flat.ctypes.data,
                                  off.ctypes.data, ln.ctypes.data, len(ln), mems.ctypes.data, mem_cap,
                                  hits.ctypes.data, hit_cap, mem_off.ctypes.data, n_out.ctypes.data)
            if n_out[0] <= mem_cap and n_out[1] <= hit_cap:
                return mems[:n_out[0]].copy(), hits[:n_out[1]].copy(), mem_off
            mem_cap, hit_cap = int(n_out[0]) + 16, int(n_out[1]) + 16

    def close(self):
        self.lib.ref_ert_close(self.h)
        self.h = None


# ---- CIGAR generation (reference mem_reg2aln through oracle/refwrap/ref_api_wrap.cpp) ---------------------------
CIGAR_JOB_DTYPE = np.dtype([("rb", "<i8"), ("re", "<i8"), ("read", "<i4"), ("qb", "<i4"), ("qe", "<i4"), ("truesc", "<i4"),
                            ("w", "<i4"), ("pad_", "<i4")])
CIGAR_OUT_DTYPE = np.dtype([("pos", "<i8"), ("cigar_off", "<i8"), ("md_off", "<i8"), ("is_rev", "<i4"), ("n_cigar", "<i4"),
                            ("nm", "<i4"), ("l_md", "<i4"), ("score", "<i4"), ("ok", "<i4")])


def reference_reg2aln(rm: "RefMem", opt, flat, off, ln, jobs):
    """The reference's own mem_reg2aln over `jobs` (CIGAR_JOB_DTYPE): (outs, cigar, md)."""
    lib = rm.lib
    lib.ref_reg2aln.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                                C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
    oi, _ = opt_arrays(opt)
    jobs = np.ascontiguousarray(jobs, CIGAR_JOB_DTYPE)
    outs = np.zeros(len(jobs), CIGAR_OUT_DTYPE)
    ccap, mcap = 8 * len(jobs) + 64, 32 * len(jobs) + 64
    while True:
        cigar, md, n_out = np.zeros(ccap, np.uint32), np.zeros(mcap, np.uint8), np.zeros(2, np.int64)
        lib.ref_reg2aln(rm.h, oi.ctypes.data, flat.ctypes.data, off.ctypes.data, ln.ctypes.data, jobs.ctypes.data, len(jobs),
                        outs.ctypes.data, cigar.ctypes.data, ccap, md.ctypes.data, mcap, n_out.ctypes.data)
        if n_out[0] <= ccap and n_out[1] <= mcap:
            return outs, cigar[:n_out[0]].copy(), md[:n_out[1]].copy()
        ccap, mcap = int(n_out[0]) + 16, int(n_out[1]) + 16
