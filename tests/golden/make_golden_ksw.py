"""Generates tests/golden/ksw_golden.npz from the REFERENCE ITSELF (oracle/_ref/libref_ksw.so =
reference src/ksw.cpp compiled by oracle/Makefile): the kswr_t records ksw_align2 returns for the
workloads in tests/ksw_cases.py.  Run in the build container:  python tests/golden/make_golden_ksw.py
Inputs are regenerated from seeds at test time; only the reference's outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import ksw_cases as kc  # noqa: E402
import oracle_lib as ol  # noqa: E402


def main():
    lib = ol.ref_ksw()
    out = {}
    for case in kc.CASES:
        name, _, (a, b, od, ed, oi, ei) = case
        pairs, ref, qer = kc.make(case)
        aln = ol.reference_ksw(lib, pairs, ref, qer, ol.score_matrix(a, b), od, ed, oi, ei)
        out[name] = np.stack([aln[f] for f in aln.dtype.names], axis=1).astype(np.int32)
    np.savez_compressed(os.path.join(HERE, "ksw_golden.npz"), **out)
    print({k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
