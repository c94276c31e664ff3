"""Generates tests/golden/bsw_golden.npz from the REFERENCE ITSELF (oracle/_ref/libref_bsw_*.so,
built by oracle/Makefile from /root/reference/src/bandedSWA.cpp).  Run in the build container:
    python tests/golden/make_golden.py
The reference ships no golden vectors for this path, so these are ours; inputs are
regenerated from the seeds at test time, only the reference's outputs are stored."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from bwa_mem_scale_b200 import synth  # noqa: E402

# (name, generator kwargs, params kwargs, kind, w)
CASES = [
    ("k16_default_w100", dict(n_pairs=3000, seed=101, n_rate=0.002, indel_rate=0.01, tiny_frac=0.05), {}, 16, 100),
    ("k16_default_w25", dict(n_pairs=3000, seed=102, n_rate=0.002, indel_rate=0.01, tiny_frac=0.05), {}, 16, 25),
    ("k16_zdrop15_e2", dict(n_pairs=3000, seed=103, n_rate=0.002, indel_rate=0.02), dict(e_del=2, e_ins=2, zdrop=15), 16, 100),
    ("k16_asym", dict(n_pairs=3000, seed=104, indel_rate=0.02), dict(o_del=4, o_ins=8, e_del=3, e_ins=1, zdrop=10, end_bonus=0), 16, 50),
    ("k8_default_w100", dict(n_pairs=3000, seed=105, n_rate=0.002, indel_rate=0.01, tiny_frac=0.05, eight_bit=True), {}, 8, 100),
    ("k8_asym", dict(n_pairs=3000, seed=106, indel_rate=0.02, eight_bit=True), dict(e_del=3, e_ins=1, zdrop=10), 8, 16),
    ("k0_default_w100", dict(n_pairs=2000, seed=107, n_rate=0.002, indel_rate=0.01, tiny_frac=0.05), {}, 0, 100),
    ("k0_e2", dict(n_pairs=2000, seed=108, indel_rate=0.02), dict(e_del=2, e_ins=3, zdrop=20), 0, 40),
]


def main():
    out = {}
    for name, gen, prm, kind, w in CASES:
        pairs, ref, qer = synth.extension_pairs(**gen)
        P = ol.default_params(**prm)
        res = {}
        for isa in ("sse41", "avx2"):
            lib = ol.ref_bsw(isa)
            assert lib is not None, "build oracle/_ref first (make -C oracle ref)"
            r = ol.reference_bsw(lib, P, kind, pairs, ref, qer, w)
            res[isa] = np.stack([r[f] for f in synth.RESULT_FIELDS], axis=1).astype(np.int16)
        assert np.array_equal(res["sse41"], res["avx2"]), name
        out[name] = res["sse41"]
    np.savez_compressed(os.path.join(HERE, "bsw_golden.npz"), **out)
    print("wrote", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
