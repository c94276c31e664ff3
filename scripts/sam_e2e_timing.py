#!/usr/bin/env python
"""Whole-program drop-in timing on the GPU box: the reference's own `mem` command, unmodified (bwa-mem2.plain) against the
same program with its seeding + extension passes routed into libbwagpu.so (bwa-mem2.gpu_plain, integration/bwagpu_shim.cpp),
same index, same FASTQ, same thread count; SAM compared byte for byte (all but @PG) and both wall times reported.
Everything is made on the box (genome, index by the reference's `index`, reads); /root/reference is not read.
   python scripts/sam_e2e_timing.py [genome_bp] [n_reads] [threads] > gpurun_out/sam_e2e.json"""
import json, os, subprocess, sys, tempfile, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib as ol
from bwa_mem_scale_b200 import fmindex, synth

n_bp = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
n_reads = int(sys.argv[2]) if len(sys.argv) > 2 else 400_000
threads = int(sys.argv[3]) if len(sys.argv) > 3 else len(os.sched_getaffinity(0))
cpu, gpu = ol.ref_binary("plain"), ol.ref_binary("gpu_plain")
assert cpu and gpu
out = {"genome_bp": n_bp, "reads": n_reads, "threads": threads}
with tempfile.TemporaryDirectory() as td:
    ref = synth.repeat_reference(n_bp, 44, n_families=20, repeat_frac=0.10, divergence=0.02)
    prefix = os.path.join(td, "ref")
    fmindex.write_fasta_contigs(prefix + ".fa", ref, (n_bp // 2, n_bp - n_bp // 2))
    t0 = time.perf_counter()
    subprocess.run([cpu, "index", "-p", prefix, prefix + ".fa"], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    out["index_s"] = time.perf_counter() - t0
    reads, _, _ = synth.simulate_reads(ref, n_reads, seed=3)
    fq = os.path.join(td, "reads.fq")
    fmindex.write_fastq(fq, reads)
    sams = {}
    for name, binary in (("reference", cpu), ("gpu", gpu), ("gpu_again", gpu)):
        sam, log = os.path.join(td, name + ".sam"), os.path.join(td, name + ".log")
        t0 = time.perf_counter()
        with open(sam, "wb") as fo, open(log, "wb") as fe:
            subprocess.run([binary, "mem", "-t", str(threads), prefix, fq], check=True, stdout=fo, stderr=fe, timeout=1500)
        out[name + "_wall_s"] = time.perf_counter() - t0
        text = open(log, errors="ignore").read()
        keep = ("Overall time", "MEM_PROCESS_SEQ", "Total kernel", "SAM Processing", "Reading IO", "Index read", "bwagpu", "Writing IO", "index replicated")
        lines = text.splitlines()
        out[name + "_profile"] = [(l.strip() + " " + (lines[i + 1].strip() if l.rstrip().endswith(":") and i + 1 < len(lines) else "")).strip()
                                  for i, l in enumerate(lines) if any(k in l for k in keep)][-24:]
        sams[name] = [l for l in open(sam, "rb") if not l.startswith(b"@PG")]
    a, b = sams["reference"], sams["gpu"]
    out["sam_lines"] = len(a)
    out["sam_identical"] = bool(len(a) == len(b) and all(x == y for x, y in zip(a, b)))
    out["speedup_whole_program"] = out["reference_wall_s"] / out["gpu_again_wall_s"]
print(json.dumps(out))
