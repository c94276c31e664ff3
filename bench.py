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
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device, self.rows, self.stop_flag, self.t = device, [], False, None

    def _run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.device}", f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True,
                                     text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def start(self):
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def stop(self):
        self.stop_flag = True
        if self.t:
            self.t.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None,
                "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


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
    ctx.timer_start()
    for _ in range(args.steps):
        step_dev()
    ms = ctx.timer_stop()
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
_dist = None


def init_dist(world):
    global _dist
    if world > 1:
        import torch
        import torch.distributed as dist
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
    ap.add_argument("--workload", default="bsw", choices=["bsw"])
    ap.add_argument("--pairs", type=int, default=1_000_000)
    ap.add_argument("--band", type=int, default=100)
    ap.add_argument("--kind", type=int, default=16, choices=[8, 16])
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(bench_bsw(args, 0, world)))
        return
    init_dist(world)
    out = bench_bsw(args, rank, world)
    if rank == 0 and out is not None:
        print(json.dumps(out))
    if _dist is not None:
        _dist.destroy_process_group()


if __name__ == "__main__":
    main()
