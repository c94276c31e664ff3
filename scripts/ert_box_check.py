"""Bench-scale live parity of the ERT path on the GPU box: GPU (sparse index + k-mer table completion on the device)
against the UNMODIFIED reference (oracle/_ref taps) on fresh simulated reads over tests/golden/ert_small_genome.npz.
  seeding : bwagpu_seed_ert            vs get_seeds[_prefix] / reseed[_prefix] / last
  pipeline: reads -> regions on device vs mem_kernel1_core_ert + mem_chain2aln_across_reads_V2
Writes a one-line JSON verdict.  Test infrastructure (uses oracle/_ref); not a product path."""
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import c2a_cases as cc            # noqa: E402
import ert_cases as ec            # noqa: E402
import oracle_lib as ol           # noqa: E402
from bwa_mem_scale_b200 import (Context, ErtIndex, ErtSeeding, Extender, FMI_search, REG_DTYPE, chain_opt, ext_opt,  # noqa: E402
                                fmindex, mem_core_ert_device, pack_reads, seed_opt, synth)


def main():
    n_seed, n_pipe = int(sys.argv[1]) if len(sys.argv) > 1 else 200000, int(sys.argv[2]) if len(sys.argv) > 2 else 50000
    d = np.load(os.path.join(ROOT, "tests", "golden", "ert_small_genome.npz"))
    ref = np.ascontiguousarray(d["ref"])
    ref_string = fmindex.text_with_revcomp(ref)
    idx = fmindex.build_fm_index(ref)
    ctx = Context(0)
    FMI_search(idx, ctx=ctx)
    ext = Extender(ref_string, len(ref), [0], [len(ref)], ctx=ctx)
    ert = ErtSeeding(ErtIndex(d["mlt_table"], sparse_idx=d["kmer_idx"], sparse_val=d["kmer_val"]), ctx)
    t0 = time.perf_counter()
    ert.complete_from_fm()
    t_fill = time.perf_counter() - t0
    dense = ert.dense_table()
    mlt = np.ascontiguousarray(d["mlt_table"])
    reads, _, _ = synth.simulate_reads(ref, n_seed, seed=77, err=0.01)
    rng = np.random.default_rng(3)
    extra = ec._sim(rng, ref, 2000, 151, 0.02, 0.01) + ec._sim(rng, ref, 2000, 90, 0.03)
    extra += [ref[1000 + k:1151 + k].copy() for k in range(0, 300, 3)] + [ref[4900 + k:5051 + k].copy() for k in range(0, 700, 5)]
    flat, off, ln = pack_reads(list(reads) + extra)
    tap = ol.RefErt((dense, mlt, ref_string))
    t0 = time.perf_counter()
    want = tap.seed(flat, off, ln)
    t_ref = time.perf_counter() - t0
    t0 = time.perf_counter()
    got = ert.seed(seed_opt(), flat, off, ln)
    t_gpu = time.perf_counter() - t0
    seed_ok = bool(np.array_equal(ec.mem_matrix(got[0]), ec.mem_matrix(want[0])) and np.array_equal(got[1], want[1])
                   and np.array_equal(got[2], want[2]))
    # full pipeline on a subset
    sub = list(reads[:n_pipe]) + extra
    flat, off, ln = pack_reads(sub)
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_index_files(prefix, ref, idx)
        rm = ol.RefMem(prefix)
        rm.use_ert_tables(dense, mlt)
        chains, seeds, regs, _ = rm.core(ol.mem_opt(), flat, off, ln)
    d_reads, d_off, d_len = ctx.upload(flat), ctx.upload(off), ctx.upload(ln)
    sres, cres, ares = mem_core_ert_device(ert, ext, seed_opt(), chain_opt(), ext_opt(), d_reads, d_off, d_len, len(ln),
                                           int(ln.max()))
    g_chains, g_seeds = np.zeros(cres.n_chains, ol.CHAIN_DTYPE), np.zeros(cres.n_seeds, ol.SEED_DTYPE)
    g_regs = np.zeros(ares.n_regs, REG_DTYPE)
    ctx.d2h(g_chains, cres.d_chains); ctx.d2h(g_seeds, cres.d_seeds); ctx.d2h(g_regs, ares.d_regs)
    pipe_ok = bool(np.array_equal(g_chains, chains) and np.array_equal(g_seeds, seeds)
                   and np.array_equal(cc.reg_matrix(g_regs), cc.reg_matrix(regs)))
    print(json.dumps({"check": "ert_box_check", "reads_seeding": int(len(want[2]) - 1), "mems": int(len(want[0])),
                      "hits": int(len(want[1])), "seeding_bit_exact": seed_ok, "reads_pipeline": len(sub),
                      "chains": int(len(chains)), "regions": int(len(regs)), "pipeline_bit_exact": pipe_ok,
                      "kmer_table_completion_s": t_fill, "reference_seeding_1thread_s": t_ref,
                      "gpu_seeding_host_call_s": t_gpu, "retry_reads": int(sres.n_retry_reads)}))
    sys.exit(0 if seed_ok and pipe_ok else 1)


if __name__ == "__main__":
    main()
