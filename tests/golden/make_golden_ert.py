"""Generates tests/golden/ert_golden.npz from the UNMODIFIED reference (run in the build container only).

  1. `bwa-mem2 index -a ert` (oracle/_ref) over ert_cases.tiny_genome()  -> .kmer_table / .mlt_table / .0123
     (about 5 minutes; pass the prefix of an existing build as argv[1] to skip it)
  2. the reference's own ERT seeding (get_seeds[_prefix] / reseed[_prefix] / last, through
     oracle/refwrap/ref_ert_wrap.cpp) over ert_cases.golden_reads() for every option case  -> MEMs + hits
  3. the reference's mem_kernel1_core_ert + mem_chain2aln_across_reads_V2 (ref_api_wrap.cpp) -> chains, regions
  4. the k-mer table entries those reads touch (logged by the host build of ert_core.cuh, which test_ert_hostsim
     pins against the tap) -> sparse (idx, val); the tree blob and the text go in verbatim.
"""
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import c2a_cases as cc          # noqa: E402
import ert_cases as ec          # noqa: E402
import hostsim_lib as hl        # noqa: E402
import oracle_lib as ol         # noqa: E402
from bwa_mem_scale_b200 import fmindex  # noqa: E402


def main():
    ref = ec.tiny_genome()
    prefix = sys.argv[1] if len(sys.argv) > 1 else "/tmp/ert_golden/g"
    if not os.path.exists(prefix + ".kmer_table"):
        os.makedirs(os.path.dirname(prefix), exist_ok=True)
        fmindex.write_fasta(prefix + ".fa", ref)
        subprocess.run([ol.ref_binary("scale"), "index", "-a", "ert", "-t", str(os.cpu_count()), "-p", prefix,
                        prefix + ".fa"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not os.path.exists(prefix + ".bwt.2bit.64"):
        ol.ref_index(prefix, prefix + ".fa")
    ref_string = np.fromfile(prefix + ".0123", np.uint8)
    assert np.array_equal(ref_string, fmindex.text_with_revcomp(ref)), "index was built from another genome"
    tap = ol.RefErt(prefix)
    kmer, mlt = tap.kmer_table(), tap.mlt_table().copy()
    reads = ec.golden_reads(ref)
    flat, off, ln = ol.pack_reads(reads)
    hs = hl.load()
    out = dict(mlt_table=mlt, n_reads=np.int64(len(ln)))
    touched = []
    for name, kw in ec.OPT_CASES:
        mems, hits, mem_off = tap.seed(flat, off, ln, **kw)
        hs.hostsim_ert_log_start()
        m2, h2, _ = hl.ert_seed(hs, kmer, mlt, ref_string, flat, off, ln, **kw)
        buf = np.zeros(8_000_000, np.uint32)
        n = hs.hostsim_ert_log_stop(buf.ctypes.data, len(buf))
        assert n <= len(buf)
        touched.append(buf[:n].copy())
        assert np.array_equal(mems, m2) and np.array_equal(hits, h2), f"{name}: host build of the device code differs"
        out[f"{name}_mems"], out[f"{name}_hits"], out[f"{name}_mem_off"] = mems, hits, mem_off
        print(name, len(mems), "MEMs", len(hits), "hits")
    idx = np.unique(np.concatenate(touched))
    out["kmer_idx"], out["kmer_val"] = idx, np.asarray(kmer[idx.astype(np.int64)])
    print(len(idx), "k-mer entries touched")
    # chains + regions of the full ERT path
    rm = ol.RefMem(prefix)
    rm.use_ert(tap)
    for name, kw in ec.OPT_CASES[:2]:
        chains, seeds, regs, reg_off = rm.core(ol.mem_opt(**kw), flat, off, ln)
        out[f"{name}_chains"], out[f"{name}_seeds"], out[f"{name}_regs"] = chains, seeds, cc.reg_matrix(regs)
        print(name, len(chains), "chains", len(regs), "regions")
    out["contig_off"], out["contig_len"] = rm.contig_off, rm.contig_len
    np.savez_compressed(ec.GOLDEN, **out)
    print("wrote", ec.GOLDEN, os.path.getsize(ec.GOLDEN), "bytes")


if __name__ == "__main__":
    main()
