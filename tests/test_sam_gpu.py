"""End-to-end drop-in check: the reference's own `mem` command with its seeding + extension passes routed into
libbwagpu.so (oracle/_ref/bwa-mem2.gpu_<flavour>, built from integration/bwagpu_shim.cpp + the unmodified reference
sources by oracle/Makefile) must write the SAME SAM, byte for byte, as the unmodified reference binary of the same
flavour — everything but the @PG line, which records the command line.  Index files are made on the box by the
reference's own `index` / `smem-table` / `perfect-index` sub-commands (oracle/_ref travels; /root/reference is not read)."""
import os
import subprocess

import numpy as np
import pytest

import c2a_cases as cc
import oracle_lib as ol
from bwa_mem_scale_b200 import fmindex, synth

pytestmark = pytest.mark.gpu


def need(flavour):
    cpu, gpu = ol.ref_binary(flavour), ol.ref_binary("gpu_" + flavour)
    assert cpu and gpu, f"oracle/_ref/bwa-mem2.{{,gpu_}}{flavour}.sse41 missing: run __graft_entry__.build() where /root/reference exists"
    return cpu, gpu


def run(cmd, out=None, log=None):
    with open(out or os.devnull, "wb") as fo, open(log or os.devnull, "wb") as fe:
        subprocess.run(cmd, check=True, stdout=fo, stderr=fe, timeout=900)


def sam_body(path):
    with open(path, "rb") as f:
        return [l for l in f if not l.startswith(b"@PG")]


def compare(cpu, gpu, prefix, fqs, args, tmp_path, tag):
    a, b = str(tmp_path / f"{tag}_cpu.sam"), str(tmp_path / f"{tag}_gpu.sam")
    run([cpu, "mem", "-t", "4", *args, prefix, *fqs], a, str(tmp_path / f"{tag}_cpu.log"))
    run([gpu, "mem", "-t", "4", *args, prefix, *fqs], b, str(tmp_path / f"{tag}_gpu.log"))
    glog = open(str(tmp_path / f"{tag}_gpu.log"), errors="ignore").read()
    assert "[bwagpu] chunk of" in glog, "the GPU path did not run"
    x, y = sam_body(a), sam_body(b)
    assert len(x) == len(y) and len(x) > 10
    bad = [i for i in range(len(x)) if x[i] != y[i]]
    assert not bad, f"{tag}: {len(bad)} of {len(x)} SAM lines differ, first:\n{x[bad[0]]!r}\n{y[bad[0]]!r}"
    return x, glog


def mapped(lines):
    recs = [l.split(b"\t") for l in lines if not l.startswith(b"@")]
    return sum(1 for r in recs if not int(r[1]) & 4), len(recs)


def test_sam_identical_config1_shape(tmp_path):
    """BASELINE config 1: 1 Mbp random reference, 10 k x 151 bp reads, 1 % substitutions, FM-index only."""
    cpu, gpu = need("plain")
    ref = synth.random_reference(1_000_000, 42)
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta(prefix + ".fa", ref)
    run([cpu, "index", "-p", prefix, prefix + ".fa"])
    reads, _, _ = synth.simulate_reads(ref, 10000, seed=1)
    fq = str(tmp_path / "reads.fq")
    fmindex.write_fastq(fq, reads)
    lines, _ = compare(cpu, gpu, prefix, [fq], [], tmp_path, "c1")
    m, n = mapped(lines)
    assert n >= 10000 and m > 0.99 * 10000
    # several chunks (-K fixes the chunk size in bases: 151 x 700 reads), i.e. ragged last 512-read batches
    compare(cpu, gpu, prefix, [fq], ["-K", str(151 * 700)], tmp_path, "c1k")


def test_sam_identical_gapped_multicontig(tmp_path):
    """Three contigs; reads with indels (CIGAR I/D, band retries), N runs, reads over contig ends and the strand
    boundary, ragged lengths including reads shorter than the minimum seed; default and narrow band."""
    cpu, gpu = need("plain")
    ref, _ = cc.hard_genome()
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta_contigs(prefix + ".fa", ref, cc.HARD_CONTIGS)
    run([cpu, "index", "-p", prefix, prefix + ".fa"])
    reads = cc.hard_reads(ref) + synth.gapped_reads(ref, 3000, seed=77) + synth.ragged_reads(ref, 500, 78)
    # 8 x 512 reads + one exact 19-base read alone in the last BATCH_SIZE slice: its single SMEM is the whole batch's,
    # which mem_chain_seeds' `pos < num_smem - 1` loop bound drops (src/bwamem.cpp:819) — the read comes out unmapped
    reads = reads[:4096] + [ref[30000:30019].copy()]
    fq = str(tmp_path / "reads.fq")
    fmindex.write_fastq(fq, reads)
    lines, glog = compare(cpu, gpu, prefix, [fq], [], tmp_path, "hard")
    body = [l.split(b"\t") for l in lines if not l.startswith(b"@")]
    assert sum(1 for r in body if b"I" in r[5] or b"D" in r[5]) > 300          # gapped alignments went through
    assert len({r[2] for r in body}) >= 4                                       # chr1-3 and '*'
    _, glog = compare(cpu, gpu, prefix, [fq], ["-w", "10", "-d", "20"], tmp_path, "hardw")
    assert "(0 retried)" not in glog                                            # band-doubling retries ran on the device
    lines, _ = compare(cpu, gpu, prefix, [fq], ["-T", "10"], tmp_path, "hardq")
    last = [l.split(b"\t") for l in lines if l.startswith(b"r4096\t")]
    assert len(last) == 1 and int(last[0][1]) & 4                               # the batch quirk was reproduced


def test_sam_identical_paired_end(tmp_path):
    """Paired-end (config 4 shape, small): seeding + extension of both ends on the device, the insert-size barrier
    (mem_pestat), mate rescue and pairing by the reference's own host code on the regions the device returned."""
    cpu, gpu = need("plain")
    ref = synth.repeat_reference(2_000_000, 44, n_families=20, repeat_frac=0.1, divergence=0.02)
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta_contigs(prefix + ".fa", ref, (1_200_000, 800_000))
    run([cpu, "index", "-p", prefix, prefix + ".fa"])
    r1, r2 = synth.simulate_pairs(ref, 6000, seed=5)
    f1, f2 = str(tmp_path / "r1.fq"), str(tmp_path / "r2.fq")
    fmindex.write_fastq(f1, r1)
    fmindex.write_fastq(f2, r2)
    lines, _ = compare(cpu, gpu, prefix, [f1, f2], ["-K", "1000000"], tmp_path, "pe")
    body = [l.split(b"\t") for l in lines if not l.startswith(b"@")]
    assert sum(1 for r in body if int(r[1]) & 2) > 0.9 * len(body)              # properly paired


def test_sam_identical_fma_emf(tmp_path):
    """FM + FMA smem tables + EMF in front (BASELINE config 2/3 index set without ERT): the `emf` feature set."""
    cpu, gpu = need("emf")
    ref = synth.repeat_reference(1_500_000, 45, n_families=20, repeat_frac=0.1, divergence=0.02)
    prefix = str(tmp_path / "ref")
    fmindex.write_fasta_contigs(prefix + ".fa", ref, (1_000_000, 500_000))
    run([cpu, "index", "-p", prefix, prefix + ".fa"])
    run([cpu, "smem-table", prefix])
    run([cpu, "perfect-index", "-l", "151", prefix])
    reads = list(synth.simulate_reads(ref, 8000, seed=6, err=0.005)[0]) + synth.gapped_reads(ref, 1500, seed=7)
    fq = str(tmp_path / "reads.fq")
    fmindex.write_fastq(fq, reads)
    lines, glog = compare(cpu, gpu, prefix, [fq], ["-l", "151"], tmp_path, "emf")
    hits = sum(int(l.split(" EMF hits")[0].split()[-1]) for l in glog.splitlines() if "[bwagpu] chunk of" in l)
    assert hits > 2500                                                          # (1 - 0.005)^151 = 47 % of 8000 reads are exact


def test_sam_identical_ert(gpu_ctx, tmp_path):
    """ERT seeding (the reference's `scale` feature set with its shm index store): the reference-built ERT index of the
    300 kbp repeat genome (fixture: tree blob + the k-mer entries that occur; the other entries completed on the device
    from the FM-index, bwagpu_index_complete_ert) is written out as <prefix>.kmer_table / .mlt_table; the unmodified
    reference and the shim binary both map it through the shm store and must write the same SAM."""
    cpu, gpu = need("scale")
    import fmi_cases as fc
    from bwa_mem_scale_b200 import ErtIndex, ErtSeeding, Extender, FMI_search
    d = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ert_small_genome.npz"))
    ref = np.ascontiguousarray(d["ref"])
    prefix = str(tmp_path / "g")
    text = fmindex.write_0123(prefix, ref)
    fmindex.write_bns_files(prefix, ref)
    Extender(text, len(ref), [0], [len(ref)], ctx=gpu_ctx)
    FMI_search(None, ctx=gpu_ctx)                                            # device-built FM-index (for the completion)
    ert = ErtSeeding(ErtIndex(d["mlt_table"], sparse_idx=d["kmer_idx"], sparse_val=d["kmer_val"]), gpu_ctx)
    ert.complete_from_fm()
    ert.dense_table().tofile(prefix + ".kmer_table")
    np.ascontiguousarray(d["mlt_table"]).tofile(prefix + ".mlt_table")
    gpu_ctx.check(gpu_ctx.lib.bwagpu_index_unload_ert(gpu_ctx.h), "bwagpu_index_unload_ert")
    reads = list(synth.simulate_reads(ref, 4000, seed=5, err=0.01)[0]) + synth.gapped_reads(ref, 1500, seed=6)
    fq = str(tmp_path / "reads.fq")
    fmindex.write_fastq(fq, reads)
    # the shm store is loaded once (load-shm: default budget -> k-mer table + tree), both runs then map it read-only;
    # without this the first `mem` leaves the store in its "manager is modifying" state and the second one waits forever
    run([cpu, "remove-shm"])
    llog = str(tmp_path / "load.log")
    subprocess.run([cpu, "load-shm", prefix], check=True, stdout=open(llog, "wb"), stderr=subprocess.STDOUT, timeout=600)
    try:
        a, b = str(tmp_path / "ert_cpu.sam"), str(tmp_path / "ert_gpu.sam")
        for binary, out, log in ((cpu, a, "ert_cpu.log"), (gpu, b, "ert_gpu.log")):
            with open(out, "wb") as fo, open(str(tmp_path / log), "wb") as fe:
                subprocess.run([binary, "mem", "-t", "4", prefix, fq], check=True, stdout=fo, stderr=fe, timeout=600)
    finally:
        subprocess.run([cpu, "remove-shm"], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=120)
    glog = open(str(tmp_path / "ert_gpu.log"), errors="ignore").read()
    assert "[bwagpu] chunk of" in glog and "useErt: 1" in open(llog, errors="ignore").read() + glog
    lines, y = sam_body(a), sam_body(b)
    bad = [i for i in range(min(len(lines), len(y))) if lines[i] != y[i]]
    assert len(lines) == len(y) and not bad, f"{len(bad)} SAM lines differ, first: {lines[bad[0]] if bad else None!r}"
    m, n = mapped(lines)
    assert n >= 5500 and m > 0.97 * 5500
