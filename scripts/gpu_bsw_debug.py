import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import oracle_lib as ol
from bsw_cases import CASES
from bwa_mem_scale_b200 import BandedPairWiseSW, Context, synth
F = synth.RESULT_FIELDS
ctx = Context(0)
for name, gen, prm, kind, w in CASES:
    if kind == 0: continue
    pairs, ref, qer = synth.extension_pairs(**gen)
    P = ol.default_params(**prm)
    bsw = BandedPairWiseSW(P.o_del, P.e_del, P.o_ins, P.e_ins, P.zdrop, P.end_bonus, None, P.w_match, P.w_mismatch, 1, ctx=ctx)
    got = pairs.copy()
    (bsw.getScores16 if kind == 16 else bsw.getScores8)(got, ref, qer, len(got), 1, w)
    want, _ = ol.oracle_bsw(P, kind, pairs, ref, qer, w)
    bad = np.zeros(len(got), bool)
    for f in F: bad |= got[f] != want[f]
    print(name, "mismatches", int(bad.sum()), "of", len(got))
    for i in np.nonzero(bad)[0][:4]:
        print("  pair", i, {f: int(pairs[f][i]) for f in ("len1", "len2", "h0")}, "got", [int(got[f][i]) for f in F], "want", [int(want[f][i]) for f in F])
        q = qer[pairs["idq"][i]:pairs["idq"][i]+pairs["len2"][i]]; t = ref[pairs["idr"][i]:pairs["idr"][i]+pairs["len1"][i]]
        print("   q has N:", int((q>3).sum()), " t has N:", int((t>3).sum()))
