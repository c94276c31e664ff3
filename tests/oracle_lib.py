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
