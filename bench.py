#!/usr/bin/env python
"""bench.py — throughput of the seeding + extension hot path on B200 (contract in the task spec).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload bsw] [--impl reference]

One "step" = one pass of the hot path over one batch of synthetic input.  `value` is measured with
inputs resident in HBM; `e2e` goes through the host-buffer C-ABI call (H2D + D2H inside the timed
region).  `--impl reference` times the reference's own CPU kernels (oracle/_ref) on host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


# ------------------------------------------------------------------------------------------------
def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler:
    """Samples SM clocks / throttle reasons of this rank's GPU during the timed region: NVML in-process (the counters
    nvidia-smi prints), one light query every 0.25 s.  Spawning nvidia-smi itself five times a second from every rank was
    measured to cost the 8-GPU run 10 % — each start initialises NVML over all GPUs and holds driver locks the other
    ranks' kernel launches wait for — so the subprocess form is only the fallback, and then at one query per second."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, device: int):
        self.device, self.rows, self.stop_flag, self.t = device, [], False, None
        self.nvml = self.handle = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nvml, self.handle = pynvml, pynvml.nvmlDeviceGetHandleByIndex(device)
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n, h = self.nvml, self.handle
        sm = n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)
        mx = n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)
        get = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        r = int(get(h))
        bits = [n.nvmlClocksThrottleReasonHwSlowdown, n.nvmlClocksThrottleReasonHwThermalSlowdown,
                n.nvmlClocksThrottleReasonSwThermalSlowdown, n.nvmlClocksThrottleReasonSwPowerCap]
        self.rows.append([str(sm), str(mx)] + ["Active" if r & b else "Not Active" for b in bits])

    def _run(self):
        while not self.stop_flag:
            try:
                if self.nvml is not None:
                    self._sample_nvml()
                else:
                    out = subprocess.run(["nvidia-smi", f"--id={self.device}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits"], capture_output=True,
                                         text=True, timeout=5).stdout.strip()
                    if out:
                        self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.25 if self.nvml is not None else 1.0)

    def start(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def stop(self):
        self.stop_flag = True
        if self.t:
            self.t.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        reasons = sorted({n for r in self.rows for n, v in zip(self.NAMES, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


# ------------------------------------------------------------------------------------------------
def cpu_threads() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def ref_bsw_cpu(pairs, ref, qer, w, kind, threads, isa="avx2"):
    """The reference's own getScores16/8 (oracle/_ref) over `pairs`, chunked across host threads
    (ctypes releases the GIL).  Returns seconds."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    lib = ol.ref_bsw(isa) or ol.ref_bsw("sse41")
    P = ol.default_params()
    chunks = np.array_split(np.arange(len(pairs)), max(threads * 4, 1))
    outs = [pairs[c].copy() for c in chunks if len(c)]
    kind_used = "reference" if lib is not None else "port"

    def work(i):
        o = outs[i]
        if lib is not None:
            import ctypes as C
            lib.ref_bsw_run(C.byref(P), kind, o.ctypes.data, len(o), ref.ctypes.data, len(ref),
                            qer.ctypes.data, len(qer), w)
        else:
            ol.oracle().orc_bsw_batch(__import__("ctypes").byref(P), kind, o.ctypes.data, len(o),
                                      ref.ctypes.data, qer.ctypes.data, w)

    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, range(len(outs))))
    return time.perf_counter() - t0, kind_used


def scalar_cells(pairs, ref, qer, w, sample=20000):
    """Cell updates as the reference's scalar kernel executes them (SW_cells, bandedSWA.cpp:199),
    counted by the oracle on a prefix sample and scaled to the batch."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    n = min(sample, len(pairs))
    _, cells = ol.oracle_bsw(ol.default_params(), 0, pairs[:n], ref, qer, w)
    return cells * (len(pairs) / n)


# ------------------------------------------------------------------------------------------------
def bench_bsw(args, rank, world):
    from bwa_mem_scale_b200 import BandedPairWiseSW, Context, synth
    n_pairs, w, kind = args.pairs, args.band, args.kind
    pairs, ref, qer = synth.extension_pairs_fast(n_pairs, seed=1000 + rank, eight_bit=(kind == 8))
    cells = scalar_cells(pairs, ref, qer, w)
    cfg = {"workload": f"banded-SW extension sweep (BASELINE config 5): {n_pairs} pairs/GPU, "
                       f"len2 U[1,150], {kind}-bit kernel, w={w}",
           "pairs_per_gpu": n_pairs, "band": w, "kind": kind,
           "l2_policy": "inputs larger than L2 (pairs+sequences > 126 MB)" if
           pairs.nbytes + ref.nbytes + qer.nbytes > 126e6 else "flush between steps"}

    if args.impl == "reference":
        if rank != 0:
            return None
        thr = cpu_threads()
        sample = pairs[: min(len(pairs), 100000 * max(thr // 4, 1))]
        scells = scalar_cells(sample, ref, qer, w)
        for _ in range(args.warmup):
            ref_bsw_cpu(sample[:2000], ref, qer, w, kind, thr)
        t = 0.0
        for _ in range(args.steps):
            dt, k = ref_bsw_cpu(sample, ref, qer, w, kind, thr)
            t += dt
        v = scells * args.steps / t / 1e9
        return {"impl": "reference", "metric": "banded-SW GCUPS", "value": v, "unit": "GCUPS",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int16" if kind == 16 else "u8", "data": "synthetic",
                "config": cfg,
                "cpu_baseline": {"value": v, "unit": "GCUPS", "cores": thr, "kind": k,
                                 "sample": f"{len(sample)} pairs of the same set, avx2 build"},
                "e2e": {"value": v, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}

    local = int(os.environ.get("LOCAL_RANK", rank))
    ctx = Context(local)
    bsw = BandedPairWiseSW(6, 1, 6, 1, 100, 5, None, 1, 4, 1, ctx=ctx)
    dp, dr, dq = ctx.upload(pairs), ctx.upload(ref), ctx.upload(qer)
    flush = ctx.malloc(256 << 20) if "flush" in cfg["l2_policy"] else None

    def step_dev():
        bsw.getScoresDevice(kind, dp, dr, ref.nbytes, dq, qer.nbytes, n_pairs, w)

    for _ in range(args.warmup):
        step_dev()
    ctx.synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = ctx.launch_count()
    ms = 0.0
    for _ in range(args.steps):
        if flush:
            ctx.memset(flush, 1, 256 << 20)
        ctx.timer_start()
        step_dev()
        ms += ctx.timer_stop()
    launches = ctx.launch_count() - l0
    ms = allreduce_max(ms)
    # e2e: host buffers through the drop-in call
    host = pairs.copy()
    bsw.getScores16(host, ref, qer, n_pairs, 1, w) if kind == 16 else bsw.getScores8(host, ref, qer, n_pairs, 1, w)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        (bsw.getScores16 if kind == 16 else bsw.getScores8)(host, ref, qer, n_pairs, 1, w)
    e2e_s = allreduce_max(time.perf_counter() - t0)
    clocks = sampler.stop()
    if rank != 0:
        return None
    pk, pk_kind = peaks()
    gcups = cells * world * args.steps / (ms * 1e-3) / 1e9
    e2e = cells * world * args.steps / e2e_s / 1e9
    # algorithmic HBM bytes of the kernel: pairs in+out, both sequence arenas once
    alg_bytes = pairs.nbytes * 2 + ref.nbytes + qer.nbytes
    ach = alg_bytes * args.steps / (ms * 1e-3) / 1e9
    out = {"metric": "banded-SW GCUPS", "value": gcups, "unit": "GCUPS", "n_gpus": world,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "int16" if kind == 16 else "u8", "data": "synthetic", "config": cfg,
           "clocks": clocks, "gpu_launches": int(launches),
           "e2e": {"value": e2e, "unit": "GCUPS", "h2d_bytes_per_step": int(alg_bytes - pairs.nbytes),
                   "d2h_bytes_per_step": int(pairs.nbytes)},
           "roofline": {"bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s",
                        "frac": ach / pk["hbm_gbs"], "traffic": None, "peak_kind": pk_kind,
                        "note": "SW is integer-issue bound, not HBM bound; see DESIGN.md"}}
    if not args.no_cpu:
        thr = cpu_threads()
        sample = pairs[: min(len(pairs), 50000 * max(thr // 4, 1))]
        scells = scalar_cells(sample, ref, qer, w)
        dt, k = ref_bsw_cpu(sample, ref, qer, w, kind, thr)
        out["cpu_baseline"] = {"value": scells / dt / 1e9, "unit": "GCUPS", "cores": thr, "kind": k,
                               "sample": f"{len(sample)} pairs of the same set, avx2 build, {dt:.1f}s"}
    return out



# ------------------------------------------------------------------------------------------------
def seed_inputs(args, rank):
    """Synthetic reference + FM-index + reads of BASELINE config 2's shape (per GPU)."""
    from bwa_mem_scale_b200 import fmindex, synth
    n_bp = int(args.ref_mbp * 1e6)
    ref = synth.repeat_reference(n_bp, 43, n_families=50, repeat_frac=0.10, divergence=0.02)
    idx = fmindex.build_fm_index(ref)
    err = 0.01
    if getattr(args, "exact_frac", None) is not None:      # P(no substitution in 151 bp) = (1-err)^151
        err = 1.0 - float(args.exact_frac) ** (1.0 / 151)
    reads, _, _ = synth.simulate_reads(ref, args.reads, seed=500 + rank, err=err)
    return ref, idx, reads


def cpu_seed(idx, reads, threads, max_occ=500):
    """CPU baseline for seeding: the oracle port (oracle/fmi_oracle.c) over `reads`, chunked over
    host threads.  Returns seconds."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    chunks = [c for c in np.array_split(np.arange(len(reads)), threads) if len(c)]

    def work(c):
        flat, off, ln = ol.pack_reads(reads[c])
        sm, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(), flat, off, ln)
        ol.oracle_sa_lookup(idx, sm, max_occ)

    from concurrent.futures import ThreadPoolExecutor
    ol.oracle()
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, chunks))
    return time.perf_counter() - t0


def bench_seed(args, rank, world):
    from bwa_mem_scale_b200 import Context, FMI_search, seed_opt, pack_reads
    from bwa_mem_scale_b200.seeding import SMEM_DTYPE  # noqa: F401
    ref, idx, reads = seed_inputs(args, rank)
    n = len(reads)
    cfg = {"workload": f"FM-index seeding (SMEM rounds 1-3 + SA look-up), BASELINE config 2 shape: "
                       f"{args.ref_mbp:g} Mbp synthetic reference (10% repeats), {n} x 151 bp SE reads/GPU, "
                       f"1% substitutions, FM-index" + (" + FMA smem tables" if args.fma else " only (no FMA tables)"),
           "reads_per_gpu": n, "ref_mbp": args.ref_mbp,
           "l2_policy": "index (%.0f MB) + reads larger than L2" % (idx.nbytes() / 1e6)
           if idx.nbytes() > 126e6 else "flush between steps"}
    if args.impl == "reference":
        if rank != 0:
            return None
        thr = cpu_threads()
        sample = reads[: min(n, 32000 * thr)]
        for _ in range(args.warmup):
            cpu_seed(idx, sample[:200 * thr], thr)
        t = sum(cpu_seed(idx, sample, thr) for _ in range(args.steps))
        v = len(sample) * args.steps / t
        return {"impl": "reference", "metric": "reads/s (151bp SE, seeding)", "value": v, "unit": "reads/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int64", "data": "synthetic", "config": cfg,
                "cpu_baseline": {"value": v, "unit": "reads/s", "cores": thr, "kind": "port",
                                 "sample": f"{len(sample)} reads of the same set"},
                "e2e": {"value": v, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    local = int(os.environ.get("LOCAL_RANK", rank))
    ctx = Context(local)
    fmi = FMI_search(idx, ctx=ctx)
    if args.fma:
        fmi.build_fma()
    opt = seed_opt()
    flat, off, ln = pack_reads(reads)
    d_reads, d_off, d_len = ctx.upload(flat), ctx.upload(off), ctx.upload(ln)
    flush = ctx.malloc(256 << 20) if "flush" in cfg["l2_policy"] else None

    def step():
        return fmi.seed_device(opt, d_reads, d_off, d_len, None, n, 151)

    for _ in range(args.warmup):
        res = step()
    ctx.synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = ctx.launch_count()
    ms = 0.0
    for _ in range(args.steps):
        if flush:
            ctx.memset(flush, 1, 256 << 20)
        ctx.timer_start()
        res = step()
        ms += ctx.timer_stop()
    ms = allreduce_max(ms)
    launches = ctx.launch_count() - l0
    # e2e: host reads in, SMEMs + coordinates back out through the drop-in calls
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        sm = fmi.collect_smem(opt, flat, off, ln)
        sa = fmi.get_sa_entries_prefetch(sm, opt.max_occ)
    e2e_s = allreduce_max(time.perf_counter() - t0)
    clocks = sampler.stop()
    # per-kernel device time of the dominant kernel (profiling serialises: separate pass)
    ctx.set_profiling(True)
    for _ in range(args.steps):
        res = step()
    k_ms, k_n = ctx.kernel_time(0)
    sa_ms, sa_n = ctx.kernel_time(2)
    ctx.set_profiling(False)
    if rank != 0:
        return None
    pk, pk_kind = peaks()
    alg_bytes = res.n_cp_records * 64          # per launch: 64 B per CP_OCC record fetched
    ach = alg_bytes / (k_ms / k_n * 1e-3) / 1e9
    out = {"metric": "reads/s (151bp SE, seeding)", "value": n * world * args.steps / (ms * 1e-3), "unit": "reads/s",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
           "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": int(launches),
           "e2e": {"value": n * world * args.steps / e2e_s, "unit": "reads/s",
                   "h2d_bytes_per_step": int(flat.nbytes + off.nbytes + ln.nbytes + sm.nbytes),
                   "d2h_bytes_per_step": int(sm.nbytes + sa.nbytes)},
           "roofline": {"bound": "hbm", "kernel": "seed_collect_kernel", "achieved": ach, "peak": pk["hbm_gbs"],
                        "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": None, "peak_kind": pk_kind,
                        "kernel_ms": k_ms / k_n, "bext_per_read": res.n_ext / n,
                        "cp_records_per_read": res.n_cp_records / n, "smem_per_read": res.n_smem / n,
                        "sa_rows_per_read": res.n_coords / n, "lf_steps_per_row": res.n_lf_steps / max(res.n_coords, 1),
                        "sa_walk_ms": sa_ms / max(sa_n, 1),
                        "sa_walk_gbs": (res.n_lf_steps * 64 + res.n_coords * 5) / (sa_ms / max(sa_n, 1) * 1e-3) / 1e9}}
    if not args.no_cpu:
        thr = cpu_threads()
        sample = reads[: min(n, 3000 * thr)]
        dt = cpu_seed(idx, sample, thr)
        out["cpu_baseline"] = {"value": len(sample) / dt, "unit": "reads/s", "cores": thr, "kind": "port",
                               "sample": f"{len(sample)} reads of the same set, oracle port, {dt:.1f}s"}
    return out

# ------------------------------------------------------------------------------------------------
# dram__bytes_read.sum + dram__bytes_write.sum of seed_collect_kernel<true>, one launch of the default
# configuration (100 Mbp, 1 M reads, FMA on): profiles/r01_ncu_mem_kernels_summary.txt
NCU_SEED_COLLECT_TRAFFIC = int(61.099131e9 + 10.094755e9)
# same for the first (all reads) launch of ert_seed_kernel, bench.py --ert 1, 1 M reads: profiles/r01_ncu_ert_seed_v2_summary.txt
NCU_ERT_SEED_TRAFFIC = int(4.793306e9 + 1.737873e9)
KERNELS = {0: "seed_collect", 1: "seed_sort", 2: "sa_walk", 3: "bsw", 4: "bsw_generic", 7: "chain2aln", 8: "chain_seeds",
           9: "emf_probe", 10: "ert_seed", 11: "ert_emit"}


def ert_bench_data():
    """Reference-built ERT index of the small bench genome (scripts/make_ert_bench_data.py): the genome, the k-mer
    table entries of the k-mers that occur and the tree blob.  The entries of absent k-mers are computed on the
    device (bwagpu_index_complete_ert)."""
    p = os.path.join(ROOT, "tests", "golden", "ert_small_genome.npz")
    if not os.path.exists(p):
        raise SystemExit("bench.py --ert 1 needs tests/golden/ert_small_genome.npz (scripts/make_ert_bench_data.py)")
    return np.load(p)


def cpu_mem_core(prefix, reads, threads, ert=None):
    """CPU baseline for seed+extend: the REFERENCE's own mem_kernel1_core + mem_chain2aln_across_reads_V2
    (oracle/_ref tap library) over `reads`, one worker per host thread on a shared index.  Returns
    (seconds, kind) — timing excludes loading the index."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    isa = "avx2" if os.path.exists(os.path.join(ol.ORACLE_DIR, "_ref", "libref_api_plain_avx2.so")) else "sse41"
    rm0 = ol.RefMem(prefix, "plain", isa)
    workers = [rm0] + [rm0.clone() for _ in range(threads - 1)]
    if ert is not None:                      # (dense k-mer table, tree blob): mem_kernel1_core_ert instead
        for w in workers:
            w.use_ert_tables(*ert)
    chunks = [c for c in np.array_split(np.arange(len(reads)), threads) if len(c)]
    o = ol.mem_opt()

    def work(a):
        w, c = a
        flat, off, ln = ol.pack_reads(reads[c])
        w.core(o, flat, off, ln)

    from concurrent.futures import ThreadPoolExecutor
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        list(ex.map(work, zip(workers, chunks)))
    return time.perf_counter() - t0, f"reference ({isa} build)"


def bench_mem_multi(args, rank, world, local, ref, idx, reads, cfg):
    """Same step, reads split over K contexts (streams) of the same GPU, each driven by its own host
    thread, so that one sub-batch's latency-bound seeding/chaining overlaps another's ALU-bound SW."""
    from concurrent.futures import ThreadPoolExecutor
    from bwa_mem_scale_b200 import (Context, Extender, FMI_search, REG_DTYPE, chain_opt, ext_opt, fmindex,
                                    mem_core_device, pack_reads, seed_opt)
    K, n = args.contexts, len(reads)
    ref_string = fmindex.text_with_revcomp(ref)
    so, co, eo = seed_opt(), chain_opt(), ext_opt()
    parts = np.array_split(np.arange(n), K)
    W = []
    for k in range(K):
        ctx = Context(local)
        fmi = FMI_search(idx, ctx=ctx)
        if args.fma:
            fmi.build_fma()
        ext = Extender(ref_string, len(ref), [0], [len(ref)], ctx=ctx)
        flat, off, ln = pack_reads(reads[parts[k]])
        W.append(dict(ctx=ctx, fmi=fmi, ext=ext, flat=flat, off=off, ln=ln, d=(ctx.upload(flat), ctx.upload(off), ctx.upload(ln))))

    def step(w, timed=False, copies=False):
        c = w["ctx"]
        if copies:
            c.h2d(w["d"][0], w["flat"]); c.h2d(w["d"][1], w["off"]); c.h2d(w["d"][2], w["ln"])
        if timed:
            c.timer_start()
        r = mem_core_device(w["fmi"], w["ext"], so, co, eo, w["d"][0], w["d"][1], w["d"][2], len(w["ln"]), 151)
        ms = c.timer_stop() if timed else 0.0
        if copies:
            out = np.zeros(r[2].n_regs, REG_DTYPE)
            c.d2h(out, r[2].d_regs)
        return ms, r

    ex = ThreadPoolExecutor(K)
    for _ in range(args.warmup):
        list(ex.map(step, W))
    for w in W:
        w["ctx"].synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = sum(w["ctx"].launch_count() for w in W)
    ms, wall = 0.0, 0.0
    for _ in range(args.steps):
        t0 = time.perf_counter()
        res = list(ex.map(lambda w: step(w, True), W))
        wall += time.perf_counter() - t0
        ms += max(r[0] for r in res)               # contexts start together: the step ends with the slowest
    ms = allreduce_max(max(ms, 0.0))
    wall = allreduce_max(wall)
    launches = sum(w["ctx"].launch_count() for w in W) - l0
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        res = list(ex.map(lambda w: step(w, False, True), W))
    e2e_s = allreduce_max(time.perf_counter() - t0)
    clocks = sampler.stop()
    if rank != 0:
        return None
    n_regs = sum(r[1][2].n_regs for r in res)
    cfg = dict(cfg, contexts=K)
    # the step is bracketed by host wall clock around K concurrently-driven streams; per-context CUDA-event
    # times are reported next to it
    return {"metric": "aligned reads/s (151bp SE, seed+extend)", "value": n * world * args.steps / wall,
            "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * wall / args.steps, "ms_per_step_events_max": ms / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64/int16",
            "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": n * world * args.steps / e2e_s, "unit": "reads/s",
                    "h2d_bytes_per_step": int(sum(w["flat"].nbytes + w["off"].nbytes + w["ln"].nbytes for w in W)),
                    "d2h_bytes_per_step": int(n_regs * REG_DTYPE.itemsize)}}


def bench_mem(args, rank, world):
    """Seed + extend for 151-bp SE reads (BASELINE config 2 shape): reads -> SMEMs -> SA -> chains ->
    banded-SW extension -> alignment regions, everything on the device."""
    from bwa_mem_scale_b200 import (Context, Extender, FMI_search, REG_DTYPE, chain_opt, ext_opt, fmindex,
                                    mem_core_device, pack_reads, seed_opt)
    ert_data = None
    if args.ert:
        from bwa_mem_scale_b200 import ErtIndex, ErtSeeding, mem_core_ert_device, synth
        ert_data = ert_bench_data()
        ref = np.ascontiguousarray(ert_data["ref"])
        idx = fmindex.build_fm_index(ref)
        reads, _, _ = synth.simulate_reads(ref, args.reads, seed=500 + rank, err=0.01)
        args.ref_mbp = len(ref) / 1e6
    else:
        ref, idx, reads = seed_inputs(args, rank)
    n = len(reads)
    cfg = {"workload": (f"seed+extend with ERT seeding: {args.ref_mbp:g} Mbp synthetic reference (40% repeats; the ERT index is "
                        f"the reference's own `index -a ert` build), {n} x 151 bp SE reads/GPU, 1% substitutions; reads -> "
                        "ERT MEMs + hits -> chains -> banded SW (both band tries) -> alignment regions" if args.ert else
                        f"seed+extend, BASELINE config 2 shape: {args.ref_mbp:g} Mbp synthetic reference (10% repeats), "
                        f"{n} x 151 bp SE reads/GPU, 1% substitutions, FM-index" + (" + FMA smem tables" if args.fma else "")
                        + "; reads -> SMEMs -> SA -> chains -> banded SW (both band tries) -> alignment regions"),
           "reads_per_gpu": n, "ref_mbp": args.ref_mbp,
           "l2_policy": "index (%.0f MB) + reference text + reads larger than L2" % (idx.nbytes() / 1e6)
           if idx.nbytes() > 126e6 else "flush between steps"}
    if args.impl == "reference":
        if rank != 0:
            return None
        import tempfile
        thr = cpu_threads()
        sample = reads[: min(n, 32000 * thr)]
        tables = None
        if args.ert:
            # input preparation only (not timed): the dense k-mer table is completed on the device and read back
            from bwa_mem_scale_b200 import Context, Extender, FMI_search
            c0 = Context(0)
            FMI_search(idx, ctx=c0)
            Extender(fmindex.text_with_revcomp(ref), len(ref), [0], [len(ref)], ctx=c0)
            e0 = ErtSeeding(ErtIndex(ert_data["mlt_table"], sparse_idx=ert_data["kmer_idx"], sparse_val=ert_data["kmer_val"]), c0)
            e0.complete_from_fm()
            tables = (e0.dense_table(), np.ascontiguousarray(ert_data["mlt_table"]))
            c0.close()
        with tempfile.TemporaryDirectory() as td:
            prefix = os.path.join(td, "ref")
            fmindex.write_index_files(prefix, ref, idx)
            for _ in range(min(args.warmup, 1)):
                cpu_mem_core(prefix, sample[: 200 * thr], thr, tables)
            t, kind = 0.0, ""
            for _ in range(args.steps):
                dt, kind = cpu_mem_core(prefix, sample, thr, tables)
                t += dt
        v = len(sample) * args.steps / t
        return {"impl": "reference", "metric": "aligned reads/s (151bp SE, seed+extend)", "value": v, "unit": "reads/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64/int16", "data": "synthetic",
                "config": cfg,
                "cpu_baseline": {"value": v, "unit": "reads/s", "cores": thr, "kind": "reference",
                                 "sample": f"{len(sample)} reads of the same set per step, {kind}, "
                                           + ("mem_kernel1_core_ert" if args.ert else "mem_kernel1_core")
                                           + " + mem_chain2aln_across_reads_V2"},
                "e2e": {"value": v, "unit": "reads/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    local = int(os.environ.get("LOCAL_RANK", rank))
    if args.contexts > 1:
        return bench_mem_multi(args, rank, world, local, ref, idx, reads, cfg)
    ctx = Context(local)
    fmi = FMI_search(idx, ctx=ctx)
    if args.fma:
        fmi.build_fma()
    ext = Extender(fmindex.text_with_revcomp(ref), len(ref), [0], [len(ref)], ctx=ctx)
    so, co, eo = seed_opt(), chain_opt(), ext_opt()
    ert = None
    if args.ert:
        t0 = time.perf_counter()
        ert = ErtSeeding(ErtIndex(ert_data["mlt_table"], sparse_idx=ert_data["kmer_idx"], sparse_val=ert_data["kmer_val"]), ctx)
        ert.complete_from_fm()
        cfg["ert_index"] = {"kmer_table_gb": 8.59, "mlt_table_mb": ert_data["mlt_table"].nbytes / 1e6,
                            "kmers_present": int(len(ert_data["kmer_idx"])), "load_and_complete_s": time.perf_counter() - t0}

    def core(dr, do, dl):
        if ert is not None:
            return mem_core_ert_device(ert, ext, so, co, eo, dr, do, dl, n, 151, emf, d_emf)
        return mem_core_device(fmi, ext, so, co, eo, dr, do, dl, n, 151, emf, d_emf)

    flat, off, ln = pack_reads(reads)
    d_reads, d_off, d_len = ctx.upload(flat), ctx.upload(off), ctx.upload(ln)
    flush = ctx.malloc(256 << 20) if "flush" in cfg["l2_policy"] else None
    emf = d_emf = None
    if args.emf:
        from bwa_mem_scale_b200 import PerfectMatch, PerfectTable
        t0 = time.perf_counter()
        table = PerfectTable.build(ref, 151, 1.1)
        cfg["emf"] = {"seed_len": 151, "slack": 1.1, "table_mb": table.seed_table.nbytes / 1e6,
                      "host_build_s": time.perf_counter() - t0}
        emf = PerfectMatch(table, ctx)
        d_emf = (ctx.malloc(4 * n), ctx.malloc(8 * n), ctx.malloc(n))

    def step():
        return core(d_reads, d_off, d_len)

    for _ in range(args.warmup):
        sres, cres, ares = step()
    ctx.synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = ctx.launch_count()
    ms = 0.0
    for _ in range(args.steps):
        if flush:
            ctx.memset(flush, 1, 256 << 20)
        ctx.timer_start()
        sres, cres, ares = step()
        ms += ctx.timer_stop()
    ms = allreduce_max(ms)
    launches = ctx.launch_count() - l0
    # e2e: every step copies its reads in from pinned host memory and its regions out to pinned host
    # memory; the copies run on the context's second stream, double-buffered, so the next batch's H2D and
    # the previous batch's D2H overlap the kernels (the reference's kt_pipeline overlaps its IO the same way)
    p_flat, p_off, p_ln = ctx.pinned(flat.shape, flat.dtype), ctx.pinned(off.shape, off.dtype), ctx.pinned(ln.shape, ln.dtype)
    p_flat[:], p_off[:], p_ln[:] = flat, off, ln
    cap = int(ares.n_regs * 1.25) + 1024
    outs = [ctx.pinned(cap, REG_DTYPE), ctx.pinned(cap, REG_DTYPE)]
    d_in = [(d_reads, d_off, d_len), (ctx.malloc(flat.nbytes), ctx.malloc(off.nbytes), ctx.malloc(ln.nbytes))]

    def copy_in(buf):
        ctx.h2d_overlap(buf[0], p_flat); ctx.h2d_overlap(buf[1], p_off); ctx.h2d_overlap(buf[2], p_ln)

    barrier()
    t0 = time.perf_counter()
    copy_in(d_in[0])
    for k in range(args.steps):
        cur = d_in[k % 2]
        ctx.h2d_fence()
        if k + 1 < args.steps:
            copy_in(d_in[(k + 1) % 2])
        sres, cres, ares = core(cur[0], cur[1], cur[2])
        out = outs[k % 2][:ares.n_regs] if ares.n_regs <= cap else np.zeros(ares.n_regs, REG_DTYPE)
        ctx.d2h_overlap(out, ares.d_regs)
    ctx.copy_synchronize()
    ctx.synchronize()
    e2e_s = allreduce_max(time.perf_counter() - t0)
    cfg["e2e_pipeline"] = "double-buffered: H2D of batch k+1 and D2H of batch k-1 on a second stream under the kernels of batch k"
    clocks = sampler.stop()
    ctx.set_profiling(True)
    for _ in range(min(args.steps, 2)):
        sres, cres, ares = step()
    ktimes = {name: ctx.kernel_time(k) for k, name in KERNELS.items()}
    ctx.set_profiling(False)
    if rank != 0:
        return None
    pk, pk_kind = peaks()
    if args.ert:
        # ert_seed_kernel: algorithmic bytes = k-mer entries (8 B), tree / hit records and text bytes the walks read,
        # counted by the kernel itself.  The walk is a chain of dependent 1-8 byte loads, so the HBM fraction is tiny by
        # construction; the figure that matters is reads in flight x load latency
        k_ms, k_n = ktimes["ert_seed"]
        alg_bytes = sres.n_index_bytes
    else:
        k_ms, k_n = ktimes["seed_collect"]
        alg_bytes = sres.n_cp_records * 64
    ach = alg_bytes / (k_ms / max(k_n, 1) * 1e-3) / 1e9
    nprof = max(min(args.steps, 2), 1)
    traffic = None
    if args.ert and n == 1_000_000 and not args.emf:
        traffic = NCU_ERT_SEED_TRAFFIC         # same configuration as the committed capture (first of the two launches)
    elif args.ref_mbp == 100.0 and n == 1_000_000 and args.fma and not args.emf and args.exact_frac is None:
        traffic = NCU_SEED_COLLECT_TRAFFIC     # same configuration as the committed capture
    purged = int(((out["qb"] == -1) & (out["qe"] == -1)).sum())
    res = {"metric": "aligned reads/s (151bp SE, seed+extend)", "value": n * world * args.steps / (ms * 1e-3),
           "unit": "reads/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "int64/int16", "data": "synthetic", "config": cfg, "clocks": clocks,
           "gpu_launches": int(launches),
           "e2e": {"value": n * world * args.steps / e2e_s, "unit": "reads/s",
                   "h2d_bytes_per_step": int(flat.nbytes + off.nbytes + ln.nbytes),
                   "d2h_bytes_per_step": int(ares.n_regs * REG_DTYPE.itemsize)},
           "roofline": dict({"bound": "hbm", "kernel": "ert_seed_kernel" if args.ert else "seed_collect_kernel",
                             "achieved": ach, "peak": pk["hbm_gbs"],
                             "unit": "GB/s", "frac": ach / pk["hbm_gbs"], "traffic": traffic, "algorithmic_bytes": int(alg_bytes),
                             "peak_kind": pk_kind, "kernel_ms": k_ms / max(k_n, 1),
                             "kernel_ms_per_step": {name: t / nprof for name, (t, c) in ktimes.items()}},
                            **({"index_bytes_per_read": sres.n_index_bytes / n, "hits_per_read": sres.n_hits / n,
                                "reads_rerun_with_larger_scratch": int(sres.n_retry_reads)} if args.ert else
                               {"bext_per_read": sres.n_ext / n, "cp_records_per_read": sres.n_cp_records / n})),
           "work": {"smem_per_read": sres.n_smem / n, "sa_rows_per_read": sres.n_coords / n,
                    "chains_per_read": cres.n_chains / n, "regions_per_read": ares.n_regs / n,
                    "regions_purged_per_read": purged / n, "sw_jobs_per_read": (ares.n_left + ares.n_right) / n,
                    "sw_jobs_retried_per_read": (ares.n_retry_left + ares.n_retry_right) / n,
                    "sw_jobs_8bit_frac": ares.n_pairs8 / max(ares.n_left + ares.n_right, 1)}}
    if emf is not None:
        skip = np.zeros(n, np.uint8)
        ctx.d2h(skip, d_emf[2])
        res["work"]["emf_answered_frac"] = float(skip.mean())
    if not args.no_cpu:
        import tempfile
        thr = cpu_threads()
        sample = reads[: min(n, 32000 * thr)]
        with tempfile.TemporaryDirectory() as td:
            prefix = os.path.join(td, "ref")
            fmindex.write_index_files(prefix, ref, idx)
            tables = (ert.dense_table(), np.ascontiguousarray(ert_data["mlt_table"])) if args.ert else None
            dt, kind = cpu_mem_core(prefix, sample, thr, tables)
        res["cpu_baseline"] = {"value": len(sample) / dt, "unit": "reads/s", "cores": thr, "kind": "reference",
                               "sample": f"{len(sample)} reads of the same set, {kind}, "
                                         + ("mem_kernel1_core_ert" if args.ert else "mem_kernel1_core")
                                         + f" + mem_chain2aln_across_reads_V2, {dt:.1f}s"}
    return res

# ------------------------------------------------------------------------------------------------
def ref_ksw_cpu(pairs, ref, qer, threads):
    """The reference's own ksw_align2 (oracle/_ref/libref_ksw.so = src/ksw.cpp, SSE) over `pairs`,
    chunked across host threads.  Returns seconds."""
    from concurrent.futures import ThreadPoolExecutor
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    lib = ol.ref_ksw()
    if lib is None:
        raise RuntimeError("oracle/_ref/libref_ksw.so is missing")
    chunks = [c for c in np.array_split(np.arange(len(pairs)), max(threads * 4, 1)) if len(c)]
    t0 = time.perf_counter()
    with ThreadPoolExecutor(max_workers=threads) as ex:
        list(ex.map(lambda c: ol.reference_ksw(lib, pairs[c], ref, qer), chunks))
    return time.perf_counter() - t0


def bench_ksw(args, rank, world):
    """Mate-rescue SW (BASELINE config 4's kernel): ksw_align2 jobs shaped like mem_matesw's, 151-bp query
    against an insert-size window, xtra = SUBO | START | BYTE | min_seed_len."""
    from bwa_mem_scale_b200 import Context, KSWR_DTYPE, MateSW, synth
    if args.harvested:
        # the ksw_align2 jobs mem_matesw issued in a real paired-end run of the reference (tests/golden/pe_rescue_jobs.npz,
        # generated by tests/golden/make_golden_pe_jobs.py): real windows, real mismatch structure, real hit/no-hit mix
        g = np.load(os.path.join(ROOT, "tests", "golden", "pe_rescue_jobs.npz"))
        pairs, ref, qer = np.ascontiguousarray(g["pairs"]), g["ref"], g["qer"]
        uniq = len(pairs)
    else:
        uniq = min(args.pairs, 20000)
        pairs, ref, qer = synth.rescue_jobs(uniq, seed=2000 + rank)
    reps = max(args.pairs // uniq, 1)
    pairs = np.tile(pairs, reps)                     # the same windows again: the kernel is compute bound
    pairs["id"] = np.arange(len(pairs))
    n = len(pairs)
    cells = float((pairs["len1"].astype(np.int64) * pairs["len2"]).sum())
    shape = ("harvested from a paired-end run of the reference (2 x 151 bp, insert N(400,50), 2 Mbp genome): 151 bp query x "
             f"{int(pairs['len1'].min())}-{int(pairs['len1'].max())} bp window") if args.harvested else "151 bp query x 200-900 bp window"
    cfg = {"workload": f"mate-rescue SW (BASELINE config 4 kernel): {n} ksw_align2 jobs/GPU ({uniq} distinct), {shape}, "
                       "xtra = SUBO|START|BYTE|19; cell updates = len1 x len2 of the forward pass",
           "jobs_per_gpu": n, "l2_policy": "flush between steps"}
    if args.impl == "reference":
        if rank != 0:
            return None
        thr = cpu_threads()
        sample = pairs[: min(n, 4000 * thr)]
        sc = float((sample["len1"].astype(np.int64) * sample["len2"]).sum())
        ref_ksw_cpu(sample[:200 * thr], ref, qer, thr)
        t = sum(ref_ksw_cpu(sample, ref, qer, thr) for _ in range(args.steps))
        v = sc * args.steps / t / 1e9
        return {"impl": "reference", "metric": "mate-rescue SW GCUPS", "value": v, "unit": "GCUPS", "n_gpus": args.gpus,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "u8/int16", "data": "synthetic", "config": cfg,
                "cpu_baseline": {"value": v, "unit": "GCUPS", "cores": thr, "kind": "reference",
                                 "sample": f"{len(sample)} jobs of the same set, ksw_align2 (SSE build)"},
                "e2e": {"value": v, "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    local = int(os.environ.get("LOCAL_RANK", rank))
    ctx = Context(local)
    sw = MateSW(ctx=ctx)
    dp, dr, dq, da = ctx.upload(pairs), ctx.upload(ref), ctx.upload(qer), ctx.malloc(n * KSWR_DTYPE.itemsize)
    flush = ctx.malloc(256 << 20)
    for _ in range(args.warmup):
        sw.batch_device(dp, dr, dq, da, n)
    ctx.synchronize()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = ctx.launch_count()
    ms = 0.0
    for _ in range(args.steps):
        ctx.memset(flush, 1, 256 << 20)
        ctx.timer_start()
        sw.batch_device(dp, dr, dq, da, n)
        ms += ctx.timer_stop()
    launches = ctx.launch_count() - l0
    ms = allreduce_max(ms)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        aln = sw.mem_sam_pe_batch(pairs, ref, qer)
    e2e_s = allreduce_max(time.perf_counter() - t0)
    clocks = sampler.stop()
    if rank != 0:
        return None
    pk, pk_kind = peaks()
    alg_bytes = pairs.nbytes + float((pairs["len1"] + pairs["len2"]).sum()) + n * KSWR_DTYPE.itemsize
    ach = alg_bytes * args.steps / (ms * 1e-3) / 1e9
    out = {"metric": "mate-rescue SW GCUPS", "value": cells * world * args.steps / (ms * 1e-3) / 1e9, "unit": "GCUPS",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/int16", "data": "synthetic",
           "config": cfg, "clocks": clocks, "gpu_launches": int(launches),
           "e2e": {"value": cells * world * args.steps / e2e_s / 1e9, "unit": "GCUPS",
                   "h2d_bytes_per_step": int(pairs.nbytes + ref.nbytes + qer.nbytes), "d2h_bytes_per_step": int(aln.nbytes)},
           "roofline": {"bound": "hbm", "achieved": ach, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": ach / pk["hbm_gbs"],
                        "traffic": None, "peak_kind": pk_kind, "note": "SW is integer-issue bound, not HBM bound; see DESIGN.md"},
           "work": {"u8_jobs_frac": float(((pairs["h0"] & 0x10000) != 0).mean()),
                    "jobs_with_hit_frac": float((aln["score"] >= 19).mean())}}
    if not args.no_cpu:
        thr = cpu_threads()
        sample = pairs[: min(n, 4000 * thr)]
        sc = float((sample["len1"].astype(np.int64) * sample["len2"]).sum())
        dt = ref_ksw_cpu(sample, ref, qer, thr)
        out["cpu_baseline"] = {"value": sc / dt / 1e9, "unit": "GCUPS", "cores": thr, "kind": "reference",
                               "sample": f"{len(sample)} jobs of the same set, ksw_align2 (SSE build), {dt:.1f}s"}
    return out


# ------------------------------------------------------------------------------------------------
_dist = None


def bench_cigar(args, rank, world):
    """CIGAR / NM / MD of final alignment regions: jobs = the regions the device pipeline keeps
    for the default workload's reads, with 0.2 % indels added so that the banded global alignment runs; one step =
    bwagpu_reg2aln_batch over all of them (host buffers in and out: this seam has no device-resident form yet)."""
    from bwa_mem_scale_b200 import (CHAIN_DTYPE, CIGAR_JOB_DTYPE, Context, Extender, FMI_search, REG_DTYPE, RegToAln, chain_opt,
                                    cigar_opt, ext_opt, fmindex, mem_core_device, pack_reads, seed_opt)
    ref, idx, reads = seed_inputs(args, rank)
    rng = np.random.default_rng(900 + rank)
    reads = np.ascontiguousarray(reads).copy()
    for i in np.flatnonzero(rng.random(len(reads)) < 0.25):          # one short deletion-like shift in a quarter of the reads
        p, k = int(rng.integers(20, 120)), int(rng.integers(1, 4))
        reads[i, p:-k] = reads[i, p + k:]
        reads[i, -k:] = rng.integers(0, 4, k)
    n = len(reads)
    cfg = {"workload": f"CIGAR/NM/MD generation (mem_reg2aln): regions kept by seed+extend for {n} x 151 bp reads/GPU on the "
                       f"{args.ref_mbp:g} Mbp synthetic reference, 25 % of the reads with a 1-3 base indel", "reads_per_gpu": n,
           "l2_policy": "inputs and backtrack slabs larger than L2"}
    local = int(os.environ.get("LOCAL_RANK", rank))
    ctx = Context(local)
    fmi = FMI_search(idx, ctx=ctx)
    ref_string = fmindex.text_with_revcomp(ref)
    ext = Extender(ref_string, len(ref), [0], [len(ref)], ctx=ctx)
    flat, off, ln = pack_reads(reads)
    d = (ctx.upload(flat), ctx.upload(off), ctx.upload(ln))
    sres, cres, ares = mem_core_device(fmi, ext, seed_opt(), chain_opt(), ext_opt(), d[0], d[1], d[2], n, 151)
    regs, chains = np.zeros(ares.n_regs, REG_DTYPE), np.zeros(cres.n_chains, CHAIN_DTYPE)
    ctx.d2h(regs, ares.d_regs); ctx.d2h(chains, cres.d_chains)
    keep = ~((regs["qb"] == -1) & (regs["qe"] == -1))
    r = regs[keep]
    jobs = np.zeros(len(r), CIGAR_JOB_DTYPE)
    for f in ("rb", "re", "qb", "qe", "truesc", "w"):
        jobs[f] = r[f]
    jobs["read"] = chains["seqid"][r["chain"]]
    cfg["jobs_per_gpu"] = int(len(jobs))
    if args.impl == "reference" or not args.no_cpu:
        # CPU arm only: the reference's own mem_reg2aln through the test-side tap library
        import tempfile
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as ol
        thr = cpu_threads()
        sample = jobs[: min(len(jobs), 40000 * thr)]
        with tempfile.TemporaryDirectory() as td:
            prefix = os.path.join(td, "ref")
            fmindex.write_index_files(prefix, ref, idx)
            rm0 = ol.RefMem(prefix)
            workers = [rm0] + [rm0.clone() for _ in range(thr - 1)]
            parts = [c for c in np.array_split(np.arange(len(sample)), thr) if len(c)]
            from concurrent.futures import ThreadPoolExecutor
            def work(a):
                ol.reference_reg2aln(a[0], ol.mem_opt(), flat, off, ln, sample[a[1]])
            t0 = time.perf_counter()
            reps = args.steps if args.impl == "reference" else 1
            for _ in range(reps):
                with ThreadPoolExecutor(thr) as ex:
                    list(ex.map(work, zip(workers, parts)))
            dt = (time.perf_counter() - t0) / reps
        cpu = {"value": len(sample) / dt, "unit": "regions/s", "cores": thr, "kind": "reference",
               "sample": f"{len(sample)} regions of the same set, the reference's own mem_reg2aln, {dt:.1f}s"}
        if args.impl == "reference":
            if rank != 0:
                return None
            return {"impl": "reference", "metric": "regions/s (CIGAR/NM/MD generation)", "value": cpu["value"], "unit": "regions/s",
                    "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt, "higher_is_better": True,
                    "scaling": "weak", "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": cfg, "cpu_baseline": cpu,
                    "e2e": {"value": cpu["value"], "unit": "regions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    r2a, co = RegToAln(ctx), cigar_opt()
    for _ in range(args.warmup):
        outs, cigar, md = r2a.mem_reg2aln(co, flat, off, ln, jobs)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    l0 = ctx.launch_count()
    ctx.set_profiling(True)
    for _ in range(args.steps):
        outs, cigar, md = r2a.mem_reg2aln(co, flat, off, ln, jobs)
    k_ms, k_n = ctx.kernel_time(12)
    ctx.set_profiling(False)
    # e2e: jobs in from the host, device-resident call (planning, sort, kernels), results back to the host
    t0 = time.perf_counter()
    for _ in range(args.steps):
        d_jobs = ctx.upload(jobs)
        res_d = r2a.mem_reg2aln_device(co, d[0], d[1], d[2], n, d_jobs, len(jobs))
        outs2, cigar2, md2 = r2a.fetch(res_d)
        ctx.free(d_jobs)
    e2e_s = allreduce_max(time.perf_counter() - t0)
    assert np.array_equal(cigar2, cigar) and np.array_equal(md2, md)
    launches = ctx.launch_count() - l0
    clocks = sampler.stop()
    if rank != 0:
        return None
    pk, pk_kind = peaks()
    dp = outs["n_cigar"] > 1
    alg = float(((jobs["qe"] - jobs["qb"]).astype(np.int64) + (jobs["re"] - jobs["rb"])).sum() + cigar.nbytes + md.nbytes)
    ms = allreduce_max(k_ms) / args.steps
    res = {"metric": "regions/s (CIGAR/NM/MD generation)", "value": len(jobs) * world / (ms * 1e-3), "unit": "regions/s",
           "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": cfg, "clocks": clocks, "gpu_launches": int(launches),
           "e2e": {"value": len(jobs) * world * args.steps / e2e_s, "unit": "regions/s",
                   "h2d_bytes_per_step": int(jobs.nbytes),
                   "d2h_bytes_per_step": int(outs.nbytes + cigar.nbytes + md.nbytes)},
           "roofline": {"bound": "hbm", "kernel": "cigar_kernel + cigar_emit_kernel", "achieved": alg / (ms * 1e-3) / 1e9, "peak": pk["hbm_gbs"],
                        "unit": "GB/s", "frac": alg / (ms * 1e-3) / 1e9 / pk["hbm_gbs"], "traffic": None, "peak_kind": pk_kind,
                        "kernel_ms": ms, "algorithmic_bytes": int(alg),
                        "note": "value = device time of the CIGAR kernels; e2e = jobs uploaded, bwagpu_reg2aln_device (planning + sort + kernels), results copied back; reads already resident"},
           "work": {"regions_per_read": len(jobs) / n, "multi_op_frac": float(dp.mean()), "cigar_ops_per_region": len(cigar) / len(jobs),
                    "md_bytes_per_region": len(md) / len(jobs)}}
    if not args.no_cpu:
        res["cpu_baseline"] = cpu
    return res


def bench_peaks(args, rank, world):
    """Integer / DPX issue-rate micro-benchmarks (the SW roofline's denominator), written to profiles/ by the caller."""
    from bwa_mem_scale_b200 import Context
    ctx = Context(int(os.environ.get("LOCAL_RANK", rank)))
    p = ctx.int_peaks()
    pk, kind = peaks()
    probe = None
    if args.ref_mbp:
        # backwardExt design probe on an index of the asked size: one walk per thread vs per 8-lane group
        import bench_config3 as b3
        from bwa_mem_scale_b200 import Extender, FMI_search, fmindex
        n_bp = int(args.ref_mbp * 1e6)
        ref = b3.make_reference(n_bp, 44)
        coff, clen = b3.contig_table(n_bp)
        Extender(fmindex.text_with_revcomp(ref), n_bp, coff, clen, ctx=ctx)
        FMI_search(None, ctx=ctx)
        probe = {"ref_mbp": args.ref_mbp, "walks": 8_000_000, "steps": 120,
                 "thread_per_walk": ctx.probe_bext(0, 8_000_000, 120), "group8_per_walk": ctx.probe_bext(1, 8_000_000, 120),
                 "group4_per_walk": ctx.probe_bext(2, 8_000_000, 120), "group2_per_walk": ctx.probe_bext(3, 8_000_000, 120)}
        assert len({v["checksum"] for v in probe.values() if isinstance(v, dict)}) == 1, "the backwardExt designs disagree"
    return {"metric": "integer issue rate", "unit": "G thread-ops/s", "int_peaks_gops": p, "hbm_gbs": pk["hbm_gbs"], "bext_probe": probe,
            "peak_kind": kind, "sm_count": 148, "how": "bwagpu_measure_int_peak: 8 independent chains per thread, 2^15 iterations, "
            "8 blocks of 256 threads per SM, best of 3 after a warm-up, CUDA events"}


def bench_mem3(args, rank, world):
    """Default workload: BASELINE config 3 shape through the worker-level C entry (bench_config3.py)."""
    import types
    import bench_config3
    H = types.SimpleNamespace(barrier=barrier, allreduce_max=allreduce_max, ClockSampler=ClockSampler, peaks=peaks,
                              cpu_threads=cpu_threads)
    return bench_config3.bench_mem3(args, rank, world, H)


def init_dist(world):
    global _dist
    if world > 1:
        import torch
        import torch.distributed as dist
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line
        if torch.cuda.is_available():
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        dist.init_process_group("nccl" if torch.cuda.is_available() else "gloo")
        _dist = dist


def barrier():
    if _dist is not None:
        _dist.barrier()


def allreduce_max(x: float) -> float:
    if _dist is None:
        return x
    import torch
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    _dist.all_reduce(t, op=_dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="mem", choices=["mem", "mem2", "bsw", "seed", "ksw", "cigar", "peaks"],
                    help="mem = BASELINE config 3 shape (default); mem2 = config 2 shape / --ert 1 (round-1 workload)")
    ap.add_argument("--ref-mbp", type=float, default=None, help="reference size in Mbp (mem: 3100, mem2: 100)")
    ap.add_argument("--reads", type=int, default=None, help="reads per GPU (mem: 10 M, mem2: 1 M)")
    ap.add_argument("--chunk", type=int, default=2_500_000, help="mem: reads per bwagpu_mem_process_chunk call")
    ap.add_argument("--cpu-reads-per-thread", type=int, default=150_000, help="mem: size of the CPU arm's bounded sample")
    ap.add_argument("--pairs", type=int, default=1_000_000)
    ap.add_argument("--harvested", type=int, default=0, help="ksw workload: use the jobs recorded from a real PE run of the reference")
    ap.add_argument("--band", type=int, default=100)
    ap.add_argument("--kind", type=int, default=16, choices=[8, 16])
    ap.add_argument("--fma", type=int, default=1, help="use the FMA smem tables (config 2)")
    ap.add_argument("--ert", type=int, default=0, help="mem workload: ERT seeding on the small reference-built index in tests/golden/")
    ap.add_argument("--emf", type=int, default=None, help="exact-match filter in front (mem: default 1, table built in HBM; mem2: default 0, host-built)")
    ap.add_argument("--exact-frac", type=float, default=None, help="fraction of error-free reads (default: what 1%% substitutions give)")
    ap.add_argument("--contexts", type=int, default=None, help="worker contexts (host thread + stream + scratch each) per GPU over one index replica (mem: default 2; mem2: 1)")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="mem workload: weak = --reads per GPU (default); strong = --reads in total, split evenly over the ranks")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    # stdout carries exactly one JSON line: everything else that writes to fd 1 (NCCL's version banner, the reference's
    # own printf chatter in the CPU arm) is sent to stderr; the line itself goes to a private copy of the original fd
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    big = args.workload == "mem"
    if args.contexts is None:
        args.contexts = 2 if big else 1
    if args.ref_mbp is None:
        args.ref_mbp = 3100.0 if big or args.workload == "peaks" else 100.0
    if args.reads is None:
        args.reads = 10_000_000 if big else 1_000_000
    if args.emf is None:
        args.emf = 1 if big else 0
    fn = {"mem": bench_mem3, "mem2": bench_mem, "peaks": bench_peaks, "bsw": bench_bsw, "seed": bench_seed, "ksw": bench_ksw, "cigar": bench_cigar}[args.workload]
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(fn(args, 0, world)), file=real_stdout, flush=True)
        return
    init_dist(world)
    out = fn(args, rank, world)
    if rank == 0 and out is not None:
        print(json.dumps(out), file=real_stdout, flush=True)
    if _dist is not None:
        _dist.destroy_process_group()


if __name__ == "__main__":
    main()
