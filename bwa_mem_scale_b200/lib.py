"""ctypes loader for libbwagpu.so — the C ABI declared in include/bwagpu.h.

There is no CPU path: if the library is missing it must be built (`__graft_entry__.build()` or
`make -C bwa_mem_scale_b200/csrc`); if no CUDA device is present `Context()` raises.
"""
from __future__ import annotations

import ctypes as C
import os

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
# BWAGPU_LIB selects a differently-tuned build of the same library (kernel A/B runs)
LIB_PATH = os.environ.get("BWAGPU_LIB") or os.path.join(PKG_DIR, "libbwagpu.so")


class BwaGpuError(RuntimeError):
    pass


class BswParams(C.Structure):
    """bwagpu_bsw_params — constructor arguments of the reference's BandedPairWiseSW
    (reference src/bandedSWA.cpp:51-98)."""
    _fields_ = [(n, C.c_int32) for n in
                ("o_del", "e_del", "o_ins", "e_ins", "zdrop", "end_bonus", "w_match", "w_mismatch")]


class KswParams(C.Structure):
    """bwagpu_ksw_params — the batch-constant arguments of ksw_align2 (reference src/ksw.cpp:347-350)."""
    _fields_ = [("mat", C.c_int8 * 25), ("pad_", C.c_int8 * 3), ("o_del", C.c_int32), ("e_del", C.c_int32),
                ("o_ins", C.c_int32), ("e_ins", C.c_int32)]


class EmfDesc(C.Structure):
    """bwagpu_emf_desc — header fields of perfect_table_t (reference src/perfect.h:175-199)."""
    _fields_ = [("seed_len", C.c_int32), ("num_loc_entry", C.c_uint32), ("num_seed_entry", C.c_uint32),
                ("num_seed_load", C.c_uint32), ("seq_len", C.c_uint32), ("num_seed_used", C.c_uint32),
                ("num_seed_key", C.c_uint32)]


class ExtOpt(C.Structure):
    """bwagpu_ext_opt — the mem_opt_t fields extension reads (reference src/bwamem.h:89-124)."""
    _fields_ = [(n, C.c_int32) for n in ("a", "b", "o_del", "e_del", "o_ins", "e_ins", "pen_clip5", "pen_clip3",
                                         "w", "zdrop")]


class C2aResult(C.Structure):
    _fields_ = [("d_regs", C.c_void_p)] + [(n, C.c_int64) for n in
                ("n_regs", "n_left", "n_right", "n_retry_left", "n_retry_right", "n_pairs8", "n_pairs16",
                 "n_pairs_scalar", "seq_bytes", "sw_cells")]


class SeedOpt(C.Structure):
    """bwagpu_seed_opt — the mem_opt_t fields seeding reads (reference src/bwamem.h:89-124)."""
    _fields_ = [("min_seed_len", C.c_int32), ("split_factor", C.c_float), ("split_width", C.c_int32),
                ("max_occ", C.c_int32), ("max_mem_intv", C.c_int64)]


class ChainOpt(C.Structure):
    """bwagpu_chain_opt — the mem_opt_t fields chaining reads (reference src/bwamem.h:89-124)."""
    _fields_ = [(n, C.c_int32) for n in ("w", "max_chain_gap", "max_occ", "min_seed_len", "min_chain_weight",
                                         "max_chain_extend")] + [("mask_level", C.c_float), ("drop_ratio", C.c_float)]


class MemOpt(C.Structure):
    """bwagpu_mem_opt — everything the worker-level entry needs of mem_opt_t / worker_t."""
    _fields_ = [("seed", SeedOpt), ("chain", ChainOpt), ("ext", ExtOpt), ("use_ert", C.c_int32), ("use_emf", C.c_int32),
                ("batch_size", C.c_int32), ("pad_", C.c_int32)]


class MemResult(C.Structure):
    """bwagpu_mem_result"""
    _fields_ = [("d_regs", C.c_void_p), ("n_regs", C.c_int64), ("d_reg_off", C.c_void_p), ("d_emf_ret", C.c_void_p),
                ("d_perfect", C.c_void_p)] + [(n, C.c_int64) for n in (
                    "n_emf_hits", "n_smem", "n_coords", "n_chains", "n_seeds", "n_regs_raw", "n_sw_left", "n_sw_right",
                    "n_sw_retry", "n_pairs8", "n_pairs16", "n_pairs_scalar", "sw_cells", "n_ext", "n_cp_records",
                    "n_lf_steps", "n_index_bytes")]


class FmIndexDesc(C.Structure):
    """bwagpu_fm_index_desc — the arrays FMI_search::load_index leaves in memory
    (reference src/FMI_search.cpp:1251-1351); count[] in loaded form (file value + 1)."""
    _fields_ = [("reference_seq_len", C.c_int64), ("count", C.c_int64 * 5), ("cp_occ", C.c_void_p),
                ("sa_ms_byte", C.c_void_p), ("sa_ls_word", C.c_void_p), ("sentinel_index", C.c_int64)]


class SeedResult(C.Structure):
    _fields_ = [("d_smems", C.c_void_p), ("n_smem", C.c_int64), ("d_coords", C.c_void_p),
                ("n_coords", C.c_int64), ("d_coord_off", C.c_void_p), ("n_ext", C.c_int64),
                ("n_cp_records", C.c_int64), ("n_lf_steps", C.c_int64), ("d_smem_off", C.c_void_p)]


class CigarOpt(C.Structure):
    """bwagpu_cigar_opt"""
    _fields_ = [("mat", C.c_int8 * 25), ("pad_", C.c_int8 * 3)] + [(n, C.c_int32) for n in ("a", "o_del", "e_del", "o_ins", "e_ins", "w")]


class CigarResult(C.Structure):
    """bwagpu_cigar_result"""
    _fields_ = [("d_outs", C.c_void_p), ("n_jobs", C.c_int64), ("d_cigar", C.c_void_p), ("n_cigar", C.c_int64),
                ("d_md", C.c_void_p), ("md_bytes", C.c_int64), ("n_gapfree", C.c_int64)]


class ErtResult(C.Structure):
    """bwagpu_ert_result"""
    _fields_ = [("d_smems", C.c_void_p), ("n_smem", C.c_int64), ("d_coords", C.c_void_p),
                ("n_coords", C.c_int64), ("d_coord_off", C.c_void_p), ("d_smem_off", C.c_void_p),
                ("n_index_bytes", C.c_int64), ("n_hits", C.c_int64), ("n_retry_reads", C.c_int64)]


class ChainResult(C.Structure):
    _fields_ = [("d_chains", C.c_void_p), ("n_chains", C.c_int64), ("d_seeds", C.c_void_p), ("n_seeds", C.c_int64)]


_lib = None

# name -> (restype, argtypes); every symbol include/bwagpu.h declares
_SIGS = {
    "bwagpu_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "bwagpu_destroy": (None, [C.c_void_p]),
    "bwagpu_ctx_share_index": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bwagpu_last_error": (C.c_char_p, [C.c_void_p]),
    "bwagpu_device_count": (C.c_int, []),
    "bwagpu_synchronize": (C.c_int, [C.c_void_p]),
    "bwagpu_launch_count": (C.c_int64, [C.c_void_p]),
    "bwagpu_set_profiling": (C.c_int, [C.c_void_p, C.c_int]),
    "bwagpu_kernel_time": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "bwagpu_timer_start": (C.c_int, [C.c_void_p]),
    "bwagpu_timer_stop": (C.c_int, [C.c_void_p, C.POINTER(C.c_float)]),
    "bwagpu_measure_int_peak": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_double)]),
    "bwagpu_probe_bext": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.c_int32, C.POINTER(C.c_double), C.POINTER(C.c_uint64),
                                    C.POINTER(C.c_float)]),
    "bwagpu_malloc": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]),
    "bwagpu_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bwagpu_h2d": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "bwagpu_d2h": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "bwagpu_memset": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int64]),
    "bwagpu_bsw_get_scores": (C.c_int, [C.c_void_p, C.POINTER(BswParams), C.c_int, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_int32, C.c_int32]),
    "bwagpu_bsw_get_scores_device": (C.c_int, [C.c_void_p, C.POINTER(BswParams), C.c_int,
                                               C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p,
                                               C.c_int64, C.c_int32, C.c_int32]),
    "bwagpu_bsw_get_scores_noalloc": (C.c_int, [C.c_void_p, C.POINTER(BswParams), C.c_int, C.c_void_p, C.c_void_p,
                                                C.c_void_p, C.c_int32, C.c_int32, C.c_int32]),
    "bwagpu_bsw_cells": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]),
    "bwagpu_mem_process_chunk": (C.c_int, [C.c_void_p, C.POINTER(MemOpt), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                           C.POINTER(MemResult)]),
    "bwagpu_mem_process_chunk_device": (C.c_int, [C.c_void_p, C.POINTER(MemOpt), C.c_void_p, C.c_void_p, C.c_void_p,
                                                  C.c_int32, C.c_int32, C.POINTER(MemResult)]),
    "bwagpu_mem_fetch": (C.c_int, [C.c_void_p, C.POINTER(MemResult), C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p,
                                   C.c_void_p]),
    "bwagpu_index_load_ref": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p,
                                        C.c_void_p]),
    "bwagpu_chain_seeds": (C.c_int, [C.c_void_p, C.POINTER(ChainOpt), C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                     C.c_void_p, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                     C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "bwagpu_chain_seeds_device": (C.c_int, [C.c_void_p, C.POINTER(ChainOpt), C.c_void_p, C.c_int64, C.c_void_p,
                                            C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                            C.POINTER(ChainResult)]),
    "bwagpu_chain_mems_device": (C.c_int, [C.c_void_p, C.POINTER(ChainOpt), C.c_void_p, C.c_int64, C.c_void_p,
                                           C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                           C.POINTER(ChainResult)]),
    "bwagpu_reg2aln_batch": (C.c_int, [C.c_void_p, C.POINTER(CigarOpt), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                       C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                       C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "bwagpu_reg2aln_device": (C.c_int, [C.c_void_p, C.POINTER(CigarOpt), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                        C.c_void_p, C.c_int64, C.POINTER(CigarResult)]),
    "bwagpu_index_load_ert": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]),
    "bwagpu_index_unload_ert": (C.c_int, [C.c_void_p]),
    "bwagpu_index_complete_ert": (C.c_int, [C.c_void_p]),
    "bwagpu_index_get_ert_table": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bwagpu_index_get_ert_kmers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "bwagpu_seed_ert": (C.c_int, [C.c_void_p, C.POINTER(SeedOpt), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                  C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p,
                                  C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "bwagpu_seed_ert_device": (C.c_int, [C.c_void_p, C.POINTER(SeedOpt), C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_int32, C.c_int32, C.POINTER(ErtResult)]),
    "bwagpu_chain2aln": (C.c_int, [C.c_void_p, C.POINTER(ExtOpt), C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32,
                                   C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]),
    "bwagpu_chain2aln_device": (C.c_int, [C.c_void_p, C.POINTER(ExtOpt), C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64,
                                          C.POINTER(C2aResult)]),
    "bwagpu_ksw_align2_batch": (C.c_int, [C.c_void_p, C.POINTER(KswParams), C.c_void_p, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_int32]),
    "bwagpu_ksw_align2_batch_device": (C.c_int, [C.c_void_p, C.POINTER(KswParams), C.c_void_p, C.c_void_p,
                                                 C.c_void_p, C.c_void_p, C.c_int32]),
    "bwagpu_host_alloc": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_void_p)]),
    "bwagpu_host_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bwagpu_h2d_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "bwagpu_d2h_async": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "bwagpu_h2d_overlap": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "bwagpu_h2d_fence": (C.c_int, [C.c_void_p]),
    "bwagpu_d2h_overlap": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64]),
    "bwagpu_copy_synchronize": (C.c_int, [C.c_void_p]),
    "bwagpu_emf_load": (C.c_int, [C.c_void_p, C.POINTER(EmfDesc), C.c_void_p, C.c_void_p]),
    "bwagpu_emf_unload": (C.c_int, [C.c_void_p]),
    "bwagpu_emf_probe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p,
                                   C.c_void_p]),
    "bwagpu_emf_probe_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                          C.c_void_p, C.c_void_p]),
    "bwagpu_emf_build": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p,
                                   C.c_void_p, C.c_int32, C.c_double, C.c_int32, C.POINTER(EmfDesc),
                                   C.POINTER(C.c_void_p), C.POINTER(C.c_void_p)]),
    "bwagpu_emf_build_device": (C.c_int, [C.c_void_p, C.c_int32, C.c_double]),
    "bwagpu_emf_get": (C.c_int, [C.c_void_p, C.POINTER(EmfDesc), C.c_void_p, C.c_void_p]),
    "bwagpu_emf_free_tables": (None, [C.c_void_p, C.c_void_p]),
    "bwagpu_index_load_fm": (C.c_int, [C.c_void_p, C.POINTER(FmIndexDesc)]),
    "bwagpu_index_build_fm": (C.c_int, [C.c_void_p]),
    "bwagpu_index_get_fm": (C.c_int, [C.c_void_p, C.POINTER(C.c_int64), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                      C.POINTER(C.c_int64)]),
    "bwagpu_index_build_fma": (C.c_int, [C.c_void_p]),
    "bwagpu_index_load_fma": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "bwagpu_index_get_fma": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "bwagpu_index_use_fma": (C.c_int, [C.c_void_p, C.c_int]),
    "bwagpu_collect_smem": (C.c_int, [C.c_void_p, C.POINTER(SeedOpt), C.c_void_p, C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_int32, C.c_void_p, C.c_int64,
                                      C.POINTER(C.c_int64)]),
    "bwagpu_get_sa_entries": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p,
                                        C.c_int64, C.POINTER(C.c_int64)]),
    "bwagpu_seed_fm_device": (C.c_int, [C.c_void_p, C.POINTER(SeedOpt), C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_int32, C.c_int32,
                                        C.POINTER(SeedResult)]),
}


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise BwaGpuError(
                f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(lib, name)          # AttributeError if the ABI lost a symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def exported_symbols() -> list[str]:
    return sorted(_SIGS)


class Context:
    """One GPU context (bwagpu_ctx): device, stream, scratch and (later) the index replica."""

    def __init__(self, device: int = 0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.bwagpu_create(device, C.byref(h))
        if rc != 0:
            raise BwaGpuError(f"bwagpu_create({device}) failed ({rc}): "
                              f"{self.lib.bwagpu_last_error(None).decode()}")
        self.h = h
        self.device = device

    def worker(self) -> "Context":
        """A second context on the same GPU (own stream and scratch) working on THIS context's index replica
        (bwagpu_ctx_share_index): one per extra host thread.  Close it before this one."""
        w = Context(self.device)
        w.check(self.lib.bwagpu_ctx_share_index(w.h, self.h), "bwagpu_ctx_share_index")
        w._owner = self                     # keeps the owner alive as long as the borrower
        return w

    def check(self, rc: int, what: str = "") -> None:
        if rc != 0:
            raise BwaGpuError(f"{what} failed ({rc}): {self.lib.bwagpu_last_error(self.h).decode()}")

    def close(self) -> None:
        if getattr(self, "h", None):
            for p in getattr(self, "_pinned", []):
                self.lib.bwagpu_host_free(self.h, p)
            self._pinned = []
            self.lib.bwagpu_destroy(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- helpers ---------------------------------------------------------------------------
    def synchronize(self) -> None:
        self.check(self.lib.bwagpu_synchronize(self.h), "synchronize")

    def launch_count(self) -> int:
        return int(self.lib.bwagpu_launch_count(self.h))

    def timer_start(self) -> None:
        self.check(self.lib.bwagpu_timer_start(self.h), "timer_start")

    def timer_stop(self) -> float:
        ms = C.c_float()
        self.check(self.lib.bwagpu_timer_stop(self.h, C.byref(ms)), "timer_stop")
        return float(ms.value)

    def set_profiling(self, on: bool) -> None:
        self.check(self.lib.bwagpu_set_profiling(self.h, 1 if on else 0), "set_profiling")

    def kernel_time(self, which: int) -> tuple[float, int]:
        ms, n = C.c_double(), C.c_int64()
        self.check(self.lib.bwagpu_kernel_time(self.h, which, C.byref(ms), C.byref(n)), "kernel_time")
        return float(ms.value), int(n.value)

    def int_peaks(self) -> dict:
        """Measured integer / DPX issue rates in G thread-ops/s (bwagpu_measure_int_peak)."""
        out = {}
        for which, name in enumerate(("iadd3", "vimax3_s32", "viaddmax_s32_relu", "viaddmax_s16x2_relu", "add_max_plain")):
            g = C.c_double()
            self.check(self.lib.bwagpu_measure_int_peak(self.h, which, C.byref(g)), "measure_int_peak")
            out[name] = float(g.value)
        return out

    def probe_bext(self, mode: int, n_walks: int, steps: int) -> dict:
        r, c, ms = C.c_double(), C.c_uint64(), C.c_float()
        self.check(self.lib.bwagpu_probe_bext(self.h, mode, n_walks, steps, C.byref(r), C.byref(c), C.byref(ms)), "probe_bext")
        return {"records_per_s": r.value, "gbs": r.value * 64 / 1e9, "checksum": int(c.value), "ms": ms.value}

    def malloc(self, nbytes: int) -> int:
        p = C.c_void_p()
        self.check(self.lib.bwagpu_malloc(self.h, int(nbytes), C.byref(p)), "malloc")
        return int(p.value)

    def free(self, dptr: int) -> None:
        self.check(self.lib.bwagpu_free(self.h, C.c_void_p(dptr)), "free")

    def h2d(self, dptr: int, arr) -> None:
        self.check(self.lib.bwagpu_h2d(self.h, C.c_void_p(dptr), arr.ctypes.data, arr.nbytes), "h2d")

    def d2h(self, arr, dptr: int) -> None:
        self.check(self.lib.bwagpu_d2h(self.h, arr.ctypes.data, C.c_void_p(dptr), arr.nbytes), "d2h")

    def h2d_async(self, dptr: int, arr) -> None:
        self.check(self.lib.bwagpu_h2d_async(self.h, C.c_void_p(dptr), arr.ctypes.data, arr.nbytes), "h2d_async")

    def d2h_async(self, arr, dptr: int) -> None:
        self.check(self.lib.bwagpu_d2h_async(self.h, arr.ctypes.data, C.c_void_p(dptr), arr.nbytes), "d2h_async")

    def h2d_overlap(self, dptr: int, arr) -> None:
        self.check(self.lib.bwagpu_h2d_overlap(self.h, C.c_void_p(dptr), arr.ctypes.data, arr.nbytes), "h2d_overlap")

    def h2d_fence(self) -> None:
        self.check(self.lib.bwagpu_h2d_fence(self.h), "h2d_fence")

    def d2h_overlap(self, arr, dptr: int) -> None:
        self.check(self.lib.bwagpu_d2h_overlap(self.h, arr.ctypes.data, C.c_void_p(dptr), arr.nbytes), "d2h_overlap")

    def copy_synchronize(self) -> None:
        self.check(self.lib.bwagpu_copy_synchronize(self.h), "copy_synchronize")

    def pinned(self, shape, dtype):
        """A numpy array over page-locked host memory (bwagpu_host_alloc); lives until the context closes."""
        import numpy as np
        dtype = np.dtype(dtype)
        n = int(np.prod(shape))
        p = C.c_void_p()
        self.check(self.lib.bwagpu_host_alloc(self.h, n * dtype.itemsize, C.byref(p)), "host_alloc")
        buf = (C.c_uint8 * max(n * dtype.itemsize, 1)).from_address(p.value)
        self._pinned = getattr(self, "_pinned", []) + [p]
        return np.frombuffer(buf, dtype=dtype, count=n).reshape(shape)

    def memset(self, dptr: int, value: int, nbytes: int) -> None:
        self.check(self.lib.bwagpu_memset(self.h, C.c_void_p(dptr), value, int(nbytes)), "memset")

    def upload(self, arr) -> int:
        import numpy as np
        arr = np.ascontiguousarray(arr)
        p = self.malloc(max(arr.nbytes, 1))
        if arr.nbytes:
            self.h2d(p, arr)
        return p
