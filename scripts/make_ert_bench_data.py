"""Makes tests/golden/ert_small_genome.npz for `bench.py --ert 1` (run in the build container only).

The reference's `index -a ert` (oracle/_ref/bwa-mem2.scale.*) is run over tests/fmi_cases.small_genome() (300 kbp,
40 % repeats, a poly-A and a dinucleotide run) unless argv[1] names the prefix of an existing build.  Kept: the
genome, the k-mer table entries of the k-mers that OCCUR (type != INVALID) and the tree blob verbatim.  The entries
of absent k-mers hold nothing but a LEP mask, which bwagpu_index_complete_ert recomputes from the FM-index on the
device — so 8 GiB of table travel as a few MB.  the file is a tracked fixture (5 MB)."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import fmi_cases as fc              # noqa: E402
import oracle_lib as ol             # noqa: E402
from bwa_mem_scale_b200 import fmindex  # noqa: E402


def main():
    ref = fc.small_genome()
    prefix = sys.argv[1] if len(sys.argv) > 1 else "/tmp/ert_small/g"
    if not os.path.exists(prefix + ".kmer_table"):
        os.makedirs(os.path.dirname(prefix), exist_ok=True)
        fmindex.write_fasta(prefix + ".fa", ref)
        subprocess.run([ol.ref_binary("scale"), "index", "-a", "ert", "-t", str(os.cpu_count()), "-p", prefix,
                        prefix + ".fa"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    assert np.array_equal(np.fromfile(prefix + ".0123", np.uint8), fmindex.text_with_revcomp(ref))
    k = np.memmap(prefix + ".kmer_table", dtype=np.uint64, mode="r")
    idxs = []
    for s in range(0, 1 << 30, 1 << 26):
        blk = np.asarray(k[s:s + (1 << 26)])
        idxs.append((np.flatnonzero(blk & 3) + s).astype(np.uint32))
    idx = np.concatenate(idxs)
    out = os.path.join(ROOT, "tests", "golden", "ert_small_genome.npz")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    np.savez_compressed(out, ref=ref, kmer_idx=idx, kmer_val=np.asarray(k[idx.astype(np.int64)]),
                        mlt_table=np.fromfile(prefix + ".mlt_table", np.uint8))
    print("wrote", out, os.path.getsize(out), "bytes;", len(idx), "k-mers occur")


if __name__ == "__main__":
    main()
