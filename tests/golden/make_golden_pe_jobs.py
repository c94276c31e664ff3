"""Generates tests/golden/pe_rescue_jobs.npz: mate-rescue jobs HARVESTED from a real paired-end run of the reference
(oracle/_ref/bwa-mem2.kswtap.sse41: the unmodified reference with every ksw_align2 call of src/bwamem_pair.cpp recorded,
oracle/refwrap/ref_ksw_tap.cpp) together with the kswr_t the reference's own ksw_align2 returned for each.
Workload: BASELINE config 4 shape at small scale — 2 Mbp repeat genome in two contigs, 2 x 151 bp pairs, insert N(400, 50),
FR; half of the second mates carry 6 % substitutions or a 12-30 bp indel so that their SE alignment is weak or missing and
mem_matesw has to rescue them.  Run in the build container:  python tests/golden/make_golden_pe_jobs.py"""
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle_lib as ol  # noqa: E402
from bwa_mem_scale_b200 import fmindex, synth  # noqa: E402

KEEP = 6000


def main():
    ref = synth.repeat_reference(2_000_000, 44, n_families=20, repeat_frac=0.15, divergence=0.02)
    r1, r2 = synth.simulate_pairs(ref, 12000, seed=5)
    rng = np.random.default_rng(6)
    r2 = [r.copy() for r in r2]
    for i in range(0, len(r2), 2):                       # damage every other second mate
        r = r2[i]
        if rng.random() < 0.5:
            sub = rng.random(len(r)) < 0.06
            r[sub] = (r[sub] + rng.integers(1, 4, int(sub.sum()))) & 3
        else:
            p, g = int(rng.integers(30, 100)), int(rng.integers(12, 31))
            r2[i] = np.concatenate([r[:p], rng.integers(0, 4, g).astype(np.uint8), r[p:]])[:151] if rng.random() < 0.5 \
                else np.concatenate([r[:p], r[p + g:], rng.integers(0, 4, g).astype(np.uint8)])
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_fasta_contigs(prefix + ".fa", ref, (1_200_000, 800_000))
        ol.ref_index(prefix, prefix + ".fa")
        f1, f2, dump = os.path.join(td, "r1.fq"), os.path.join(td, "r2.fq"), os.path.join(td, "jobs.bin")
        fmindex.write_fastq(f1, r1)
        fmindex.write_fastq(f2, r2)
        env = dict(os.environ, BWAGPU_KSW_TAP=dump)
        subprocess.run([ol.ref_binary("kswtap"), "mem", "-t", "4", prefix, f1, f2], check=True, env=env,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        raw = np.fromfile(dump, np.uint8)
    jobs, pos = [], 0
    while pos < len(raw):
        h = raw[pos:pos + 40].view(np.int32)
        ql, tl = int(h[0]), int(h[1])
        jobs.append((h.copy(), raw[pos + 40:pos + 40 + ql], raw[pos + 40 + ql:pos + 40 + ql + tl]))
        pos += 40 + ql + tl
    print(len(jobs), "jobs harvested")
    pick = np.sort(np.random.default_rng(7).choice(len(jobs), min(KEEP, len(jobs)), replace=False))
    pairs = np.zeros(len(pick), synth.SEQPAIR_DTYPE)
    want = np.zeros((len(pick), 7), np.int32)
    qs, ts, qo, to = [], [], 0, 0
    for k, j in enumerate(pick):
        h, q, t = jobs[j]
        pairs[k]["idq"], pairs[k]["idr"], pairs[k]["id"] = qo, to, k
        pairs[k]["len2"], pairs[k]["len1"], pairs[k]["h0"] = len(q), len(t), h[2]
        pairs[k]["seqid"] = pairs[k]["regid"] = k
        want[k] = h[3:10]
        qs.append(q); ts.append(t); qo += len(q); to += len(t)
    out = os.path.join(HERE, "pe_rescue_jobs.npz")
    np.savez_compressed(out, pairs=pairs, qer=np.concatenate(qs + [np.zeros(64, np.uint8)]),
                        ref=np.concatenate(ts + [np.zeros(64, np.uint8)]), kswr=want, n_harvested=np.array([len(jobs)]))
    x = pairs["h0"]
    print("wrote", out, os.path.getsize(out), "bytes;", len(pick), "jobs; 8-bit", int(((x & 0x10000) != 0).sum()),
          "with start", int(((x & 0x80000) != 0).sum()), "window", int(pairs["len1"].min()), int(pairs["len1"].max()),
          "hits (score >= 30):", int((want[:, 0] >= 30).sum()))


if __name__ == "__main__":
    main()
