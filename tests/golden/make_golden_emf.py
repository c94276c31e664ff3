"""Generates tests/golden/emf_golden.npz from the REFERENCE ITSELF: the table its `perfect-index`
sub-command writes (oracle/_ref/bwa-mem2.emf.sse41) for the genome of tests/emf_cases.py, and the
return codes / bseq1_perfect_t its find_perfect_match_entry gives (oracle/_ref/libref_api_emf_sse41.so,
and libref_api_scale_sse41.so for the MEMSCALE partial load).  Run in the build container:
    python tests/golden/make_golden_emf.py
Inputs are regenerated from seeds at test time; only digests of the table and the probe outputs are stored."""
import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import emf_cases as ec  # noqa: E402
import oracle_lib as ol  # noqa: E402
from bwa_mem_scale_b200 import PerfectTable  # noqa: E402

CASES = [("l101", ec.SEED_LEN, 1.1), ("l151", 151, 1.02), ("l32", 32, 1.5)]


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def main():
    g = ec.genome()
    d = tempfile.mkdtemp()
    ol.write_emf_inputs(d + "/e", *g)
    flat, off, ln = ol.pack_reads(ec.reads(*g))
    out = {}
    for name, L, slack in CASES:
        path = ol.ref_perfect_index(d + "/e", L, slack)
        t = PerfectTable.load(path)
        out[f"{name}_hdr"] = np.array([t.seed_len, t.seq_len, t.num_seed_entry, len(t.loc_table), t.num_seed_used,
                                       t.num_seed_key], np.int64)
        out[f"{name}_loc_sha256"], out[f"{name}_seed_sha256"] = sha(t.loc_table), sha(t.seed_table)
        r = ol.RefEmf(path, d + "/e.0123")
        out[f"{name}_ret"], out[f"{name}_perfect"] = r.probe(flat, off, ln)
        r.close()
        r = ol.RefEmf(path, d + "/e.0123", flavor="scale", num_seed_load=t.num_seed_entry // 2)
        out[f"{name}_half_ret"], out[f"{name}_half_perfect"] = r.probe(flat, off, ln)
        r.close()
        print(name, out[f"{name}_hdr"], np.bincount(out[f"{name}_ret"], minlength=6))
    np.savez_compressed(os.path.join(HERE, "emf_golden.npz"), **out)


if __name__ == "__main__":
    main()
