"""Builds and binds tests/hostsim (the kernels' per-thread seeding code compiled as host C++).
TEST SCAFFOLDING: lets the CPU suite single-step the exact state machine the CUDA kernels run."""
import ctypes as C
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "hostsim", "fmi_hostsim.cpp")
SRC2 = os.path.join(ROOT, "tests", "hostsim", "chain_hostsim.cpp")
SRC3 = os.path.join(ROOT, "tests", "hostsim", "ert_hostsim.cpp")
SRC4 = os.path.join(ROOT, "tests", "hostsim", "cigar_hostsim.cpp")
OUT = os.path.join(ROOT, "tests", "hostsim", "_build", "libhostsim.so")


def load():
    deps = [SRC, SRC2, SRC3, SRC4, os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "cigar_core.cuh"), os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "fmi_core.cuh"),
            os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "ert_core.cuh"),
            os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "ert_build.cuh"),
            os.path.join(ROOT, "bwa_mem_scale_b200", "csrc", "chain_core.cuh"),
            os.path.join(ROOT, "include", "bwagpu.h")]
    if not os.path.exists(OUT) or any(os.path.getmtime(d) > os.path.getmtime(OUT) for d in deps):
        os.makedirs(os.path.dirname(OUT), exist_ok=True)
        subprocess.run(["g++", "-O2", "-fPIC", "-shared", "-x", "c++", "-o", OUT, SRC, SRC2, SRC3, SRC4], check=True)
    lib = C.CDLL(OUT)
    lib.hostsim_collect_smem.restype = C.c_int64
    lib.hostsim_sa_lookup.restype = C.c_int64
    lib.hostsim_chain_seeds.restype = C.c_int64
    lib.hostsim_ert_seed.restype = C.c_int
    lib.hostsim_ert_log_stop.restype = C.c_int64
    lib.hostsim_ert_log_stop.argtypes = [C.c_void_p, C.c_int64]
    lib.hostsim_ert_seed.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int, C.c_float, C.c_int, C.c_int,
                                     C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_void_p,
                                     C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
    return lib


def ert_seed(lib, kmer, mlt, ref_string, flat, off, ln, min_seed_len=19, split_factor=1.5, split_width=10,
             max_mem_intv=20, per_read_mems=16, per_read_hits=128):
    """The ERT walker of the CUDA kernels (csrc/ert_core.cuh) run read by read on the host: (mems, hits, mem_off)
    in the layout of oracle_lib.RefErt.seed.  `kmer` is the dense 2^30-entry table (may be a memory map)."""
    import numpy as np
    from oracle_lib import ERT_MEM_DTYPE
    mem_cap, hit_cap = len(ln) * 16 + 64, len(ln) * 256 + 1024
    while True:
        mems, hits = np.zeros(mem_cap, ERT_MEM_DTYPE), np.zeros(hit_cap, np.int64)
        mem_off, n_out = np.zeros(len(ln) + 1, np.int64), np.zeros(3, np.int64)
        lib.hostsim_ert_seed(kmer.ctypes.data, mlt.ctypes.data, ref_string.ctypes.data, len(ref_string) // 2,
                             min_seed_len, split_factor, split_width, max_mem_intv, flat.ctypes.data,
                             off.ctypes.data, ln.ctypes.data, len(ln), mems.ctypes.data, mem_cap, hits.ctypes.data,
                             hit_cap, mem_off.ctypes.data, n_out.ctypes.data, per_read_mems, per_read_hits)
        if n_out[0] <= mem_cap and n_out[1] <= hit_cap:
            ert_seed.last_bytes = int(n_out[2])
            return mems[:n_out[0]].copy(), hits[:n_out[1]].copy(), mem_off
        mem_cap, hit_cap = int(n_out[0]) + 16, int(n_out[1]) + 16


def reg2aln(lib, copt, l_pac, ref_string, flat, off, ln, jobs):
    """The CIGAR code of the CUDA kernel (csrc/cigar_core.cuh) run job by job on the host: (outs, cigar, md)."""
    import numpy as np
    from oracle_lib import CIGAR_JOB_DTYPE, CIGAR_OUT_DTYPE
    jobs = np.ascontiguousarray(jobs, CIGAR_JOB_DTYPE)
    outs = np.zeros(len(jobs), CIGAR_OUT_DTYPE)
    ccap, mcap = 8 * len(jobs) + 64, 32 * len(jobs) + 64
    lib.hostsim_reg2aln.restype = C.c_int
    lib.hostsim_reg2aln.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                    C.c_int64, C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p]
    while True:
        cigar, md, n_out = np.zeros(ccap, np.uint32), np.zeros(mcap, np.uint8), np.zeros(3, np.int64)
        lib.hostsim_reg2aln(C.byref(copt), l_pac, ref_string.ctypes.data, flat.ctypes.data, off.ctypes.data, ln.ctypes.data,
                            jobs.ctypes.data, len(jobs), outs.ctypes.data, cigar.ctypes.data, ccap, md.ctypes.data, mcap,
                            n_out.ctypes.data)
        if n_out[0] <= ccap and n_out[1] <= mcap:
            reg2aln.last_redo = int(n_out[2])
            return outs, cigar[:n_out[0]].copy(), md[:n_out[1]].copy()
        ccap, mcap = int(n_out[0]) + 16, int(n_out[1]) + 16
