"""bench.py's default workload: BASELINE config 3 — seed + extend for 151-bp single-end reads on a 3.1 Gbp synthetic
reference (24 contigs, seed 44), FM-index + FMA smem tables (+ EMF with --emf 1) replicated per GPU, 10 M reads per GPU.

Everything goes through the worker-level C entry (bwagpu_mem_process_chunk{,_device} + bwagpu_mem_fetch): one call per
chunk of reads, as the reference's mem_process_seqs is called once per chunk (src/fastmap.cpp:392-424).  The index is
built in HBM by bwagpu_index_build_fm / bwagpu_index_build_fma from the synthetic text (there is no 3.1 Gbp index to
ship and the reference's own builder needs the better part of an hour); for the CPU arms it is copied back and
handed to the unmodified reference in memory (oracle/refwrap ref_mem_open_arrays).
"""
from __future__ import annotations

import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
METRIC = "aligned reads/s (151bp SE, seed+extend)"
READ_LEN = 151
GEN_CHUNK = 500_000          # reads per generator task (fixes the read set independently of thread count)


def log(*a):
    print("[bench]", *a, file=sys.stderr, flush=True)


def contig_table(n_bp: int, n_contigs: int = 24):
    """Contig lengths falling linearly from ~2x to ~1x of the smallest, like chr1..chr22,X,Y; offsets in the
    concatenated forward text."""
    w = np.arange(2 * n_contigs, n_contigs, -1, dtype=np.float64)
    ln = np.floor(w / w.sum() * n_bp).astype(np.int64)
    ln[0] += n_bp - ln.sum()
    off = np.concatenate([[0], np.cumsum(ln)[:-1]]).astype(np.int64)
    return off, ln.astype(np.int32)


def make_reference(n_bp: int, seed: int = 44, repeat_frac: float = 0.10, n_families: int = 50, divergence: float = 0.02):
    """Same generator as synth.repeat_reference (i.i.d. bases + mutated copies of 50 repeat families, either strand),
    with the random background drawn in parallel blocks so that 3.1 G bases take seconds."""
    n_bp = int(n_bp)
    ref = np.empty(n_bp, np.uint8)
    blk = 1 << 27
    starts = list(range(0, n_bp, blk))

    def fill(i):
        s = starts[i]
        e = min(s + blk, n_bp)
        ref[s:e] = np.random.default_rng([seed, i]).integers(0, 4, size=e - s, dtype=np.uint8)

    with ThreadPoolExecutor(min(16, os.cpu_count() or 1)) as ex:
        list(ex.map(fill, range(len(starts))))
    rng = np.random.default_rng([seed, 1 << 20])
    fams = [rng.integers(0, 4, size=int(rng.integers(200, 2001)), dtype=np.uint8) for _ in range(n_families)]
    target, placed = int(n_bp * repeat_frac), 0
    while placed < target:
        f = fams[int(rng.integers(0, n_families))]
        if len(f) >= n_bp:
            break
        cp = f.copy()
        mut = rng.random(len(cp)) < divergence
        cp[mut] = (cp[mut] + rng.integers(1, 4, size=int(mut.sum()))) & 3
        if rng.random() < 0.5:
            cp = (3 - cp[::-1]).astype(np.uint8)
        p = int(rng.integers(0, n_bp - len(cp)))
        ref[p:p + len(cp)] = cp
        placed += len(cp)
    return ref


def make_reads(ref: np.ndarray, first: int, count: int, seed: int, err: float = 0.01, out: np.ndarray | None = None):
    """Reads [first, first + count) of the infinite read set `seed`: generated in fixed blocks of GEN_CHUNK reads, each
    from its own stream (so any slice is reproducible), 151 bp, uniform start, 1 % substitutions, random strand."""
    from bwa_mem_scale_b200 import synth
    if out is None:
        out = np.empty((count, READ_LEN), np.uint8)
    b0, b1 = first // GEN_CHUNK, (first + count + GEN_CHUNK - 1) // GEN_CHUNK

    def gen(b):
        r, _, _ = synth.simulate_reads(ref, GEN_CHUNK, READ_LEN, err, seed=seed * 100003 + b)
        lo, hi = max(first, b * GEN_CHUNK), min(first + count, (b + 1) * GEN_CHUNK)
        out[lo - first:hi - first] = r[lo - b * GEN_CHUNK:hi - b * GEN_CHUNK]

    with ThreadPoolExecutor(min(8, os.cpu_count() or 1)) as ex:
        list(ex.map(gen, range(b0, b1)))
    return out


def workload_name(args, n, world, chunk):
    idx = "FM-index + FMA smem tables" if args.fma else "FM-index"
    if args.emf:
        idx += " + EMF (l=151) in front"
    return (f"BASELINE config 3 shape: {args.ref_mbp / 1e3:g} Gbp synthetic reference (24 contigs, 10% repeats, seed 44), "
            f"{n} x 151 bp SE reads per GPU (1% substitutions), {idx}, index replicated per GPU; one "
            f"bwagpu_mem_process_chunk call per {chunk} reads: reads -> [EMF] -> SMEMs -> SA -> chains -> banded SW "
            "(both band tries) -> compacted per-read alignment regions.  ERT seeding is not part of this line: the ERT tree "
            "builder is the reference's offline tool (hours at this size); --workload mem2 --ert 1 times ERT seeding on the "
            "small reference-built index")


def cpu_reference_arm(idx_host, fma_tabs, text, coff, clen, reads, threads, emf_table=None):
    """The reference's own mem_kernel1_core + mem_chain2aln_across_reads_V2 over `reads` (oracle/_ref tap library, AVX2
    build; the `accel` feature set when the FMA tables are on, `emf` with the exact-match filter), one worker per host thread on one shared in-memory index.
    Returns (seconds, description)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    flavor = "emf" if emf_table is not None else "accel" if fma_tabs is not None else "plain"
    isa = "avx2" if os.path.exists(os.path.join(ol.ORACLE_DIR, "_ref", f"libref_api_{flavor}_avx2.so")) else "sse41"
    rm0 = ol.RefMem.from_arrays(idx_host, text, coff, clen, flavor, isa)
    if fma_tabs is not None:
        rm0.set_fma(*fma_tabs)
    if emf_table is not None:
        rm0.set_perfect(emf_table)
    workers = [rm0] + [rm0.clone() for _ in range(threads - 1)]
    chunks = [c for c in np.array_split(np.arange(len(reads)), threads) if len(c)]
    packed = [ol.pack_reads(reads[c]) for c in chunks]            # outside the timed region
    o = ol.mem_opt()

    def work(a):
        w, (flat, off, ln) = a
        w.core(o, flat, off, ln)

    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, zip(workers, packed)))
    return time.perf_counter() - t0, f"reference ({flavor} feature set, {isa} build)"


def bench_mem3(args, rank, world, H):
    """H: helpers of bench.py (barrier, allreduce_max, ClockSampler, peaks, cpu_threads)."""
    from bwa_mem_scale_b200 import (Context, Extender, FMI_search, MemChunk, REG_DTYPE, fmindex, mem_opt)
    local = int(os.environ.get("LOCAL_RANK", rank))
    n_bp = int(args.ref_mbp * 1e6)
    strong = getattr(args, "scaling", "weak") == "strong"
    n = args.reads // world if strong else args.reads         # strong: --reads is the whole job, split evenly over the ranks
    chunk = min(args.chunk, n)
    thr = H.cpu_threads()
    K = max(1, int(args.contexts))
    cfg = {"workload": workload_name(args, n, world, chunk), "reads_per_gpu": n, "ref_mbp": args.ref_mbp,
           "chunk_reads": chunk, "contexts_per_gpu": K, "scaling": "strong" if strong else "weak", "fma": int(args.fma), "emf": int(args.emf),
           "l2_policy": "index and reads larger than L2 (CP_OCC %.1f GB, reads %.2f GB per step)" % (2 * n_bp / 1e9, n * READ_LEN / 1e9)}
    t_all = time.perf_counter()
    ref = make_reference(n_bp, 44)
    coff, clen = contig_table(n_bp)
    text = fmindex.text_with_revcomp(ref)
    log(f"reference {n_bp / 1e6:g} Mbp generated in {time.perf_counter() - t_all:.1f}s")
    ctx = Context(local)
    t0 = time.perf_counter()
    Extender(text, n_bp, coff, clen, ctx=ctx)
    t1 = time.perf_counter()
    fmi = FMI_search(None, ctx=ctx)                       # bwagpu_index_build_fm: the index is built in HBM
    ctx.synchronize()
    t2 = time.perf_counter()
    if args.fma:
        fmi.build_fma()
        ctx.synchronize()
    t3 = time.perf_counter()
    setup = {"text_upload_s": t1 - t0, "fm_index_build_s": t2 - t1, "fma_build_s": t3 - t2}
    log("index:", setup)
    emf = None
    if args.emf:
        if not args.fma:
            raise SystemExit("--emf 1 needs --fma 1 (the reference's EMF feature set includes the FMA tables)")
        from bwa_mem_scale_b200 import PerfectMatch
        t0 = time.perf_counter()
        emf = PerfectMatch.build_on_device(ctx, READ_LEN, 1.1)        # bwagpu_emf_build_device: the table is built in HBM
        ctx.synchronize()
        setup["emf_build_s"] = time.perf_counter() - t0
        log("EMF table built in", setup["emf_build_s"], "s")
    # reference arm / CPU baseline need the index on the host: copied back once, outside every timed region
    reference_arm = args.impl == "reference"
    want_cpu = reference_arm or (rank == 0 and not args.no_cpu)
    idx_host = fma_tabs = emf_table = None
    if want_cpu:
        t0 = time.perf_counter()
        idx_host = fmi.get_index()
        fma_tabs = fmi.get_fma() if args.fma else None
        emf_table = emf.get_table() if emf is not None else None
        setup["index_copy_back_s"] = time.perf_counter() - t0
    if reference_arm:
        ctx.close()
        per_step = min(n, args.cpu_reads_per_thread * thr)
        reads = make_reads(ref, 0, per_step, seed=500)
        cpu_reference_arm(idx_host, fma_tabs, text, coff, clen, reads[: 200 * thr], thr, emf_table)        # warm-up
        t, kind = 0.0, ""
        for _ in range(args.steps):
            dt, kind = cpu_reference_arm(idx_host, fma_tabs, text, coff, clen, reads, thr, emf_table)
            t += dt
        v = per_step * args.steps / t
        return {"impl": "reference", "metric": METRIC, "value": v, "unit": "reads/s", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
                "scaling": cfg["scaling"], "vs_baseline": None, "dtype": "int64/int16", "data": "synthetic", "config": cfg,
                "cpu_baseline": {"value": v, "unit": "reads/s", "cores": thr, "kind": "reference",
                                 "sample": f"the first {per_step} reads of rank 0's read set per step, {kind}, mem_kernel1_core + "
                                           "mem_chain2aln_across_reads_V2 on one shared index (built on the GPU, handed over in "
                                           "memory)"},
                "e2e": {"value": v, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "setup": setup}

    # ---- the GPU arm -------------------------------------------------------------------------------------------
    t0 = time.perf_counter()
    flat = ctx.pinned((n, READ_LEN), np.uint8)
    make_reads(ref, 0, n, seed=500 + rank, out=flat)
    setup["reads_generate_s"] = time.perf_counter() - t0
    flat1 = flat.reshape(-1)
    off = np.arange(n, dtype=np.int64) * READ_LEN
    ln = np.full(n, READ_LEN, np.int32)
    d_reads, d_off, d_len = ctx.upload(flat1), ctx.upload(off), ctx.upload(ln)
    # K worker contexts on this GPU over the one index replica (bwagpu_ctx_share_index), one host thread each: chunk j goes
    # to context j % K, so one context's host-side waits and copies overlap another's kernels
    ctxs = [ctx] + [ctx.worker() for _ in range(K - 1)]
    mcs = [MemChunk(c) for c in ctxs]
    mc = mcs[0]
    pool = ThreadPoolExecutor(K) if K > 1 else None
    opt = mem_opt(use_emf=bool(args.emf))
    bounds = [(c0, min(c0 + chunk, n)) for c0 in range(0, n, chunk)]
    STAT = ("n_emf_hits", "n_smem", "n_coords", "n_chains", "n_seeds", "n_regs_raw", "n_regs", "n_sw_left", "n_sw_right",
            "n_sw_retry", "n_pairs8", "n_pairs16", "n_pairs_scalar", "sw_cells", "n_ext", "n_cp_records", "n_lf_steps")

    def step_device():
        def run(k):
            part = dict.fromkeys(STAT, 0)
            for c0, c1 in bounds[k::K]:
                r = mcs[k].process_device(opt, d_reads, d_off + 8 * c0, d_len + 4 * c0, c1 - c0, READ_LEN)
                for key in STAT:
                    part[key] += getattr(r, key)
            ctxs[k].synchronize()
            return part
        parts = [run(0)] if K == 1 else list(pool.map(run, range(K)))
        return {key: sum(p[key] for p in parts) for key in STAT}

    for _ in range(args.warmup):
        tot = step_device()
    ctx.synchronize()
    H.barrier()
    sampler = H.ClockSampler(local)
    sampler.start()
    l0 = sum(c.launch_count() for c in ctxs)
    ms = 0.0
    for _ in range(args.steps):
        ctx.timer_start()                   # events on context 0's stream; every context is idle here and again
        tot = step_device()                 # when step_device returns (each worker synchronises its own stream)
        ms += ctx.timer_stop()
    ms = H.allreduce_max(ms)
    launches = sum(c.launch_count() for c in ctxs) - l0

    # e2e: HOST buffers through bwagpu_mem_process_chunk + bwagpu_mem_fetch, every chunk's reads copied in and its regions,
    # offsets and EMF answers copied out inside the timed region
    cap = int(tot["n_regs"] / len(bounds) * 1.3) + 4096
    PERF = np.dtype([("flags", "<u4"), ("location", "<u4")])
    pin = [dict(regs=c.pinned(cap, REG_DTYPE), off=c.pinned(chunk + 1, np.int64), ret=c.pinned(chunk, np.int32),
                perf=c.pinned(chunk, PERF)) for c in ctxs]
    off_rel = np.arange(chunk, dtype=np.int64) * READ_LEN
    d2h = 0

    def step_host():
        nonlocal d2h
        def run(k):
            out = 0
            for c0, c1 in bounds[k::K]:
                nc = c1 - c0
                r = mcs[k].process(opt, flat1[c0 * READ_LEN:c1 * READ_LEN], off_rel[:nc], ln[c0:c1])
                regs = pin[k]["regs"] if r.n_regs <= cap else np.zeros(r.n_regs, REG_DTYPE)
                mcs[k].fetch(r, nc, regs, pin[k]["off"], pin[k]["ret"], pin[k]["perf"])
                out += r.n_regs * REG_DTYPE.itemsize + (nc + 1) * 8 + nc * 12
            return out
        d2h = run(0) if K == 1 else sum(pool.map(run, range(K)))

    step_host()
    H.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_host()
    e2e_s = H.allreduce_max(time.perf_counter() - t0)
    clocks = sampler.stop()
    # per-kernel device times: one more step with every launch bracketed by events (serialises the stream)
    ctx.set_profiling(True)
    for c0, c1 in bounds:                   # all chunks on context 0, one after the other
        mc.process_device(opt, d_reads, d_off + 8 * c0, d_len + 4 * c0, c1 - c0, READ_LEN)
    KERN = {0: "seed_collect", 1: "seed_sort", 2: "sa_walk", 3: "bsw", 4: "bsw_generic", 7: "chain2aln", 8: "chain_seeds", 9: "emf_probe"}
    kt = {name: ctx.kernel_time(k) for k, name in KERN.items()}
    ctx.set_profiling(False)
    int_peaks = ctx.int_peaks()
    if rank != 0:
        return None
    return report(args, world, H, ctx, cfg, setup, n, ms, e2e_s, launches, clocks, tot, kt, int_peaks, flat1.nbytes + 12 * n, d2h,
                  dict(ref=ref, text=text, coff=coff, clen=clen, idx_host=idx_host, fma_tabs=fma_tabs, emf_table=emf_table, flat=flat, thr=thr,
                       t_all=t_all))


# SASS instructions per DP cell of bsw_fast_kernel<16>'s whole-word loop (118 instructions per 8 cells in cuobjdump -sass of
# the shipped build, loop overhead and arg-max bookkeeping included)
BSW_SASS_PER_CELL = 14.75
# integer operations a cell needs at best with two int16 cells per DPX operation (~7 instructions / 2 cells)
BSW_IDEAL_OPS_PER_2CELLS = 7.0


def report(args, world, H, ctx, cfg, setup, n, ms, e2e_s, launches, clocks, tot, kt, int_peaks, h2d, d2h, S):
    pk, pk_kind = H.peaks()
    steps = args.steps
    kms = {name: t for name, (t, c) in kt.items()}                       # one profiled step
    kernels = []
    # banded SW: integer / DPX issue bound.  GCUPS = DP cells computed (counted on the device, one per execution of the
    # inner-loop body, the reference's SW_cells) / kernel time
    bsw_ms = kms["bsw"] + kms["bsw_generic"]
    gcups = tot["sw_cells"] / (bsw_ms * 1e-3) / 1e9 if bsw_ms > 0 else 0.0
    peak_gcups = int_peaks["viaddmax_s16x2_relu"] * 2.0 / BSW_IDEAL_OPS_PER_2CELLS
    kernels.append({"kernel": "bsw_fast_kernel<16|8>", "ms": bsw_ms, "bound": "int_issue", "achieved": gcups, "unit": "GCUPS",
                    "peak": peak_gcups, "frac": gcups / peak_gcups if peak_gcups else None,
                    "peak_def": "measured __viaddmax_s16x2_relu issue rate x 2 cells / 7 operations per 2 cells",
                    "sass_per_cell": BSW_SASS_PER_CELL,
                    "frac_of_issue": gcups * BSW_SASS_PER_CELL / int_peaks["iadd3"] if int_peaks["iadd3"] else None,
                    "cells": int(tot["sw_cells"])})
    # seeding: algorithmic bytes = SURVEY.md 8(d)'s figure, 64 B per CP_OCC record a backward extension touches (the reference
    # reads the whole record).  With the paired record layout this kernel asks HBM for one 32-byte half per touch, so the
    # bytes it really requests are half of that: `requested_*` says so next to the algorithmic figure.
    sc_ms = kms["seed_collect"]
    sc_gbs = tot["n_cp_records"] * 64 / (sc_ms * 1e-3) / 1e9 if sc_ms > 0 else 0.0
    kernels.append({"kernel": "seed_collect_kernel", "ms": sc_ms, "bound": "hbm", "achieved": sc_gbs, "unit": "GB/s",
                    "peak": pk["hbm_gbs"], "frac": sc_gbs / pk["hbm_gbs"], "algorithmic_bytes": int(tot["n_cp_records"] * 64),
                    "requested_bytes": int(tot["n_cp_records"] * 32), "requested_gbs": sc_gbs / 2,
                    "requested_frac": sc_gbs / 2 / pk["hbm_gbs"]})
    sa_ms = kms["sa_walk"]
    sa_gbs = (tot["n_lf_steps"] * 64 + tot["n_coords"] * 5) / (sa_ms * 1e-3) / 1e9 if sa_ms > 0 else 0.0
    kernels.append({"kernel": "sa_walk_kernel", "ms": sa_ms, "bound": "hbm", "achieved": sa_gbs, "unit": "GB/s",
                    "peak": pk["hbm_gbs"], "frac": sa_gbs / pk["hbm_gbs"]})
    for name in ("chain_seeds", "chain2aln", "seed_sort", "emf_probe"):
        kernels.append({"kernel": name, "ms": kms[name], "bound": "latency"})
    kernels.sort(key=lambda k: -k["ms"])
    top = kernels[0]
    roof = {"bound": "hbm" if top["bound"] == "hbm" else "int_issue", "kernel": top["kernel"], "achieved": top.get("achieved"),
            "peak": top.get("peak"), "unit": top.get("unit"), "frac": top.get("frac"), "traffic": None,
            "traffic_note": "no in-run DRAM counter; ncu captures of the same command are summarised under profiles/",
            "peak_kind": pk_kind, "kernel_ms": top["ms"], "profiled_step_ms": sum(kms.values()), "kernels": kernels,
            "int_peaks_gops": int_peaks, "gcups": gcups}
    nsw = tot["n_sw_left"] + tot["n_sw_right"]
    res = {"metric": METRIC, "value": n * world * steps / (ms * 1e-3), "unit": "reads/s", "n_gpus": world, "steps": steps,
           "warmup": args.warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": cfg["scaling"], "vs_baseline": None,
           "dtype": "int64/int16", "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": int(launches),
           "e2e": {"value": n * world * steps / e2e_s, "unit": "reads/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
           "gcups": gcups, "roofline": roof, "setup": setup,
           "work": {"emf_answered_frac": tot["n_emf_hits"] / n, "smem_per_read": tot["n_smem"] / n,
                    "sa_rows_per_read": tot["n_coords"] / n, "bext_per_read": tot["n_ext"] / n,
                    "cp_records_per_read": tot["n_cp_records"] / n, "chains_per_read": tot["n_chains"] / n,
                    "regions_per_read": tot["n_regs"] / n, "regions_purged_per_read": (tot["n_regs_raw"] - tot["n_regs"]) / n,
                    "sw_jobs_per_read": nsw / n, "sw_jobs_retried_per_read": tot["n_sw_retry"] / n,
                    "sw_jobs_8bit_frac": tot["n_pairs8"] / max(nsw, 1), "sw_cells_per_job": tot["sw_cells"] / max(nsw + tot["n_sw_retry"], 1)}}
    if not args.no_cpu:
        thr = S["thr"]
        m = min(n, args.cpu_reads_per_thread * thr)
        dt, kind = cpu_reference_arm(S["idx_host"], S["fma_tabs"], S["text"], S["coff"], S["clen"], np.asarray(S["flat"][:m]), thr,
                                     S["emf_table"])
        res["cpu_baseline"] = {"value": m / dt, "unit": "reads/s", "cores": thr, "kind": "reference",
                               "sample": f"the first {m} reads of rank 0's read set, {kind}, mem_kernel1_core + "
                                         f"mem_chain2aln_across_reads_V2 on one shared in-memory index, {dt:.1f}s"}
    res["setup"]["total_s"] = time.perf_counter() - S["t_all"]
    return res
