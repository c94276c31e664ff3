"""Generates tests/golden/fmi_golden.npz from the REFERENCE ITSELF (oracle/_ref/libref_api_*.so and
the reference binary's `index` command, both built by oracle/Makefile from /root/reference/src):
the sorted SMEM list mem_collect_smem returns and the SA coordinates get_sa_entries_prefetch
returns, for the workloads in tests/fmi_cases.py.  Run in the build container:
    python tests/golden/make_golden_fmi.py
Inputs are regenerated from seeds at test time; only the reference's outputs are stored
(SMEMs as an int64 digest per field plus the full arrays of the smallest case)."""
import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import fmi_cases as fc  # noqa: E402
import oracle_lib as ol  # noqa: E402
from bwa_mem_scale_b200 import fmindex  # noqa: E402


def digest(a: np.ndarray) -> np.ndarray:
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def main():
    ref = fc.small_genome()
    reads = fc.small_reads(ref)
    flat, off, ln = ol.pack_reads(reads)
    out = {}
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_fasta(prefix + ".fa", ref)
        ol.ref_index(prefix, prefix + ".fa")
        out["index_sha256"] = digest(np.fromfile(prefix + ".bwt.2bit.64", np.uint8))
        res = {}
        for isa in ("sse41", "avx2"):
            lib = ol.ref_api("plain", isa)
            h = lib.ref_fmi_open(prefix.encode())
            for name, kw, max_occ in fc.OPT_CASES:
                sm = ol.reference_collect_smem(lib, h, ol.default_seed_opt(**kw), flat, off, ln)
                sa = ol.reference_sa_lookup(lib, h, sm, max_occ)
                res[(isa, name)] = (fc.smem_matrix(sm), sa)
            lib.ref_fmi_close(h)
        for name, _, _ in fc.OPT_CASES:
            a, b = res[("sse41", name)], res[("avx2", name)]
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), name
            out[f"{name}_smem_sha256"] = digest(a[0])
            out[f"{name}_sa_sha256"] = digest(a[1])
            out[f"{name}_counts"] = np.array([len(a[0]), len(a[1])], np.int64)
        # FMA ("accel") flavour: the reference's own smem-table output and its SMEM lists with the
        # tables in use (they differ from the plain lists on reads with N and for short seeds)
        import subprocess
        subprocess.run([ol.ref_binary("accel", "sse41"), "smem-table", prefix], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        all_t = np.fromfile(prefix + ".all_smem.11", np.uint32)
        last_t = np.fromfile(prefix + ".last_smem.13", np.uint32)
        out["fma_all_sha256"] = digest(all_t)
        out["fma_last_sha256"] = digest(last_t)
        rng = np.random.default_rng(11)
        ia = np.concatenate([np.arange(512), rng.integers(0, 1 << 22, 1536)])
        il = np.concatenate([np.arange(512), rng.integers(0, 1 << 26, 1536)])
        out["fma_all_sample_idx"], out["fma_all_sample"] = ia, all_t.reshape(-1, 32)[ia]
        out["fma_last_sample_idx"], out["fma_last_sample"] = il, last_t.reshape(-1, 4)[il]
        lib = ol.ref_api("accel", "sse41")
        h = lib.ref_fmi_open(prefix.encode())
        for name, kw, max_occ in fc.OPT_CASES:
            sm = ol.reference_collect_smem(lib, h, ol.default_seed_opt(**kw), flat, off, ln)
            out[f"fma_{name}_smem_sha256"] = digest(fc.smem_matrix(sm))
            out[f"fma_{name}_count"] = np.array([len(sm)], np.int64)
        lib.ref_fmi_close(h)
        # full vectors for the first 200 reads of the default case (small enough to commit)
        sm, sa = res[("sse41", "default")]
        keep = sm[:, 0] < 200
        out["default_head_smem"] = sm[keep].astype(np.int64)
        out["default_head_sa"] = sa[:int(np.minimum(sm[keep][:, 5], 500).sum())]
    np.savez_compressed(os.path.join(HERE, "fmi_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
