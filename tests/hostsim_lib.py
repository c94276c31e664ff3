"""Builds and binds tests/hostsim (the kernels' per-thread seeding code compiled as host C++).
TEST SCAFFOLDING: lets the CPU suite single-step the exact state machine the CUDA kernels run."""
import ctypes as C
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "hostsim", "fmi_hostsim.cpp")
SRC2 = os.path.join(ROOT, "tests", "hostsim", "chain_hostsim.cpp")
SRC3 = os.path.join(ROOT, "tests", "hostsim", "ert_hostsim.cpp")
OUT = os.path.join(ROOT, "tests", "hostsim", "_build", "libhostsim.so")


def load():
    deps = [SRC, SRC2, SRC3, os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "fmi_core.cuh"),
            os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "ert_core.cuh"),
            os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "chain_core.cuh"),
            os.path.join(ROOT, "include", "bwagpu.h")]
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-x", "c++", "-o", OUT, SRC, SRC2, SRC3], check=True)
    lib = C.CDLL(OUT)
    lib.hostsim_collect_smem.restype = C.c_int64
    lib.hostsim_sa_lookup.restype = C.c_int64
    lib.hostsim_chain_seeds.restype = C.c_int64
    lib.hostsim_ert_seed.restype = C.c_int
    lib.hostsim_ert_seed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p,
                                     C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    return lib
