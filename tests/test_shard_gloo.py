"""CPU test of the multi-GPU path's host logic with world_size 2 over gloo: every
rank takes its slice of the chunk (bwa_mem_scale_b200.shard), runs seeding -> chaining -> extension on
it (here the CPU stand-ins the other tests pin to the reference: the oracle and the host build of
the device chaining code), rank 0 stitches the per-rank results back in input order and they must
equal the single-batch golden records; bench.py's barrier / max-over-ranks timing helpers run on the
same process group."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    sys.path.insert(0, HERE)
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import bench
    import c2a_cases as cc
    import fmi_cases as fc
    import oracle_lib as ol
    from bwa_mem_scale_b200 import shard
    from test_chain_hostsim import GOLD, hostsim_chains, seeds_of

    bench.init_dist(world)
    assert dist.get_backend() == "gloo" and dist.get_world_size() == world
    ref, idx = fc.small_index()
    flat, off, ln = ol.pack_reads(cc.golden_reads(ref))
    o = ol.mem_opt()
    f_r, off_r, ln_r, first = shard.shard_reads(flat, off, ln, world, rank)
    sm, co = seeds_of(idx, o, f_r, off_r, ln_r)
    chains, seeds = hostsim_chains(o, GOLD["l_pac"][0], GOLD["contig_off"], GOLD["contig_len"], sm, co, ln_r)
    text = np.concatenate([ref, 3 - ref[::-1]]).astype(np.uint8)
    regs = ol.oracle_chain2aln(o, len(ref), GOLD["contig_off"], GOLD["contig_len"], text, f_r, off_r, ln_r, chains, seeds)
    bench.barrier()
    t = bench.allreduce_max(10.0 + rank)                 # max over ranks, as the timed region is reduced
    parts = [None] * world if rank == 0 else None
    dist.gather_object((first, sm, chains, seeds, regs, t), parts, dst=0)
    if rank == 0:
        firsts = [p[0] for p in parts]
        all_sm = shard.merge_smems([p[1] for p in parts], firsts)
        c, s, r = shard.merge_regions([p[2] for p in parts], [p[3] for p in parts], [p[4] for p in parts], firsts)
        np.savez(out_path, sm=fc.smem_matrix(all_sm), chains=c, seeds=s, regs=cc.reg_matrix(r),
                 t=np.array([p[5] for p in parts]), firsts=np.array(firsts))
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_two_rank_shards_stitch_back_to_the_single_batch_result(tmp_path):
    import torch.multiprocessing as mp
    sys.path.insert(0, HERE)
    import c2a_cases as cc
    import fmi_cases as fc
    import oracle_lib as ol
    from test_chain_hostsim import GOLD, seeds_of
    world, port = 2, 29500 + os.getpid() % 2000
    out = str(tmp_path / "merged.npz")
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    got = np.load(out)
    assert list(got["t"]) == [11.0, 11.0]                       # both ranks see the max
    assert got["firsts"][0] == 0 and 0 < got["firsts"][1] < len(cc.golden_reads(fc.small_index()[0]))
    ref, idx = fc.small_index()
    flat, off, ln = ol.pack_reads(cc.golden_reads(ref))
    sm, _ = seeds_of(idx, ol.mem_opt(), flat, off, ln)
    assert np.array_equal(got["sm"], fc.smem_matrix(sm))
    assert np.array_equal(got["chains"], GOLD["default_chains"])
    assert np.array_equal(got["seeds"], GOLD["default_seeds"])
    assert np.array_equal(got["regs"], GOLD["default_regs"])


def test_shard_bounds_cover_every_read_once():
    from bwa_mem_scale_b200 import shard
    for n in (0, 1, 7, 512, 1000003):
        for w in (1, 2, 3, 8):
            b = shard.shard_bounds(n, w)
            assert b[0] == 0 and b[-1] == n and np.all(np.diff(b) >= 0) and np.diff(b).max() - np.diff(b).min() <= 1
    flat = np.arange(20, dtype=np.uint8)
    off, ln = np.array([0, 5, 9, 15], np.int64), np.array([5, 4, 6, 5], np.int32)
    f1, o1, l1, first = shard.shard_reads(flat, off, ln, 2, 1)
    assert first == 2 and list(o1) == [0, 6] and list(l1) == [6, 5] and list(f1) == list(range(9, 20))
    f0, o0, l0, first = shard.shard_reads(flat, off, ln, 8, 7)
    assert first == 3 and list(f0) == list(range(15, 20))
    assert len(shard.shard_reads(flat, off, ln, 8, 0)[2]) == 0
