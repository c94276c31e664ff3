"""GPU parity tests (through the C ABI): the EMF probe kernel vs golden data from the reference's
find_perfect_match_entry and vs the oracle."""
import os

import numpy as np
import pytest

import emf_cases as ec
import oracle_lib as ol
from bwa_mem_scale_b200 import Extender, PerfectMatch, PerfectTable, fmindex, synth
from test_oracle_emf import CASES, GOLD, table, workload

pytestmark = pytest.mark.gpu


def resident_text(gpu_ctx, g, text):
    return Extender(text, len(g[0]), g[1], g[2], ctx=gpu_ctx)


@pytest.mark.parametrize("case", CASES, ids=[c[0] for c in CASES])
def test_gpu_probe_matches_golden_and_oracle(gpu_ctx, case):
    name, L, slack = case
    g, text, (flat, off, ln) = workload()
    resident_text(gpu_ctx, g, text)
    t = table(*case)
    pm = PerfectMatch(t, gpu_ctx)
    ret, perfect, skip = pm.find_perfect_match_entry(flat, off, ln)
    pf = np.stack([perfect["flags"], perfect["location"]], axis=1)
    assert np.array_equal(ret, GOLD[f"{name}_ret"]) and np.array_equal(pf, GOLD[f"{name}_perfect"])
    assert np.array_equal(skip, ((ret == 3) | (ret == 4)).astype(np.uint8))
    want = ol.oracle_emf_probe(t, text, flat, off, ln)
    assert np.array_equal(ret, want[0]) and np.array_equal(pf, want[1])
    # MEMSCALE partial load (perfect.h:752-758): entries past num_seed_load are absent
    pm = PerfectMatch(t, gpu_ctx, num_seed_load=t.num_seed_entry // 2)
    ret, perfect, _ = pm.find_perfect_match_entry(flat, off, ln)
    pf = np.stack([perfect["flags"], perfect["location"]], axis=1)
    assert np.array_equal(ret, GOLD[f"{name}_half_ret"]) and np.array_equal(pf, GOLD[f"{name}_half_perfect"])
    pm.unload()


def test_gpu_probe_without_table_and_empty_batch(gpu_ctx):
    g, text, (flat, off, ln) = workload()
    resident_text(gpu_ctx, g, text)
    pm = PerfectMatch(None, gpu_ctx)
    ret, perfect, skip = pm.find_perfect_match_entry(flat, off, ln)
    assert np.all(ret == 0) and np.all(perfect["flags"] == 0) and np.all(skip == 0)
    pm = PerfectMatch(table(*CASES[0]), gpu_ctx)
    ret, _, _ = pm.find_perfect_match_entry(flat[:0], off[:0], ln[:0])
    assert len(ret) == 0
    pm.unload()


def test_gpu_probe_config1_shape_and_seeding_skip(gpu_ctx):
    """1 Mbp random reference, 10 k x 151 bp reads, a third of them error-free: probe == oracle, and
    the reads it answers drop out of seeding exactly as the reference's skip does (bwamem.cpp:672-689)."""
    from bwa_mem_scale_b200 import FMI_search, fmindex, seed_opt
    import fmi_cases as fc
    ref = synth.random_reference(1_000_000, 42)
    text = np.concatenate([ref, 3 - ref[::-1]]).astype(np.uint8)
    reads, _, _ = synth.simulate_reads(ref, 10000, seed=1, err=0.007)
    flat, off, ln = ol.pack_reads(reads)
    t = PerfectTable.build(ref, 151, 1.1)
    Extender(text, len(ref), [0], [len(ref)], ctx=gpu_ctx)
    pm = PerfectMatch(t, gpu_ctx)
    ret, perfect, skip = pm.find_perfect_match_entry(flat, off, ln)
    want = ol.oracle_emf_probe(t, text, flat, off, ln)
    assert np.array_equal(ret, want[0])
    assert np.array_equal(np.stack([perfect["flags"], perfect["location"]], axis=1), want[1])
    assert 2000 < skip.sum() < 5000
    # every answered read really is a verbatim copy of the text at the reported place
    for i in np.nonzero(skip)[0][:200]:
        loc = int(perfect["location"][i])
        w = ref[loc:loc + 151]
        assert np.array_equal(reads[i], w if ret[i] == 3 else 3 - w[::-1])
    idx = fmindex.build_fm_index(ref)
    fmi = FMI_search(idx, ctx=gpu_ctx)
    got = fmi.collect_smem(seed_opt(), flat, off, ln, skip)
    exp, _ = ol.oracle_collect_smem(idx, ol.default_seed_opt(), flat, off, ln, skip)
    assert np.array_equal(fc.smem_matrix(got), fc.smem_matrix(exp))
    assert not np.any(skip[got["rid"]])
    pm.unload()


# ---- the table built in HBM (bwagpu_emf_build_device) ---------------------------------------------------------------
def _lists(table, flags):
    """(fw list, rc list) a perfect.flags value dereferences to in `table` (GET_MULTI_FW_AND_RC, perfect.h:170-185)."""
    ml = int(flags) >> 2
    if ml == 0:
        return (), ()
    lt = table.loc_table
    head = int(lt[ml])
    if head & 0x80000000:
        st = head & 0x7FFFFFFF
        nfw, nrc, st = int(lt[st]), int(lt[st + 1]), st + 2
    else:
        nfw, nrc, st = (head >> 16) & 0xFFFF, head & 0xFFFF, ml + 1
    return tuple(lt[st:st + nfw]), tuple(lt[st + nfw:st + nfw + nrc])


@pytest.mark.parametrize("seed_len", [ec.SEED_LEN, 151, 64, 33])
def test_gpu_device_built_table_is_equivalent(gpu_ctx, tmp_path, seed_len):
    """Same content as the table of the host builder (which is byte-identical to the reference's perfect-index): every
    read gets the same code, location and strand bits, and its list id dereferences to the same fw / rc lists; the
    reference's own find_perfect_match_entry agrees when it runs on the device-built table."""
    ref, offs, lens, hole_off, hole_len = ec.genome()         # built WITHOUT the holes; the reads still probe around them
    text = fmindex.text_with_revcomp(ref)
    Extender(text, len(ref), offs, lens, ctx=gpu_ctx)
    want_t = PerfectTable.build(ref, seed_len, 1.1, contig_off=offs, contig_len=lens)
    flat, off, ln = ol.pack_reads(ec.reads(ref, offs, lens, hole_off, hole_len))
    pm_w = PerfectMatch(want_t, gpu_ctx)
    want = pm_w.find_perfect_match_entry(flat, off, ln)
    pm = PerfectMatch.build_on_device(gpu_ctx, seed_len, 1.1)
    got = pm.find_perfect_match_entry(flat, off, ln)
    got_t = pm.get_table()
    assert np.array_equal(got[0], want[0]) and np.array_equal(got[2], want[2])
    assert np.array_equal(got[1]["location"], want[1]["location"])
    assert np.array_equal(got[1]["flags"] & 3, want[1]["flags"] & 3)
    hit = np.flatnonzero((got[0] == 3) | (got[0] == 4))
    assert len(hit) > 500
    n_multi = 0
    for i in hit:
        a, b = _lists(got_t, got[1]["flags"][i]), _lists(want_t, want[1]["flags"][i])
        assert a == b, i
        n_multi += bool(a[0] or a[1])
    assert n_multi > 20 and any(len(_lists(got_t, got[1]["flags"][i])[0]) >= 256 for i in hit)     # the poly-A list
    # table invariants: same set of entries, one root per occupied slot, header statistics
    assert got_t.num_seed_entry == want_t.num_seed_entry and got_t.num_seed_used == want_t.num_seed_used
    assert got_t.num_seed_key == want_t.num_seed_key
    gv, wv = got_t.seed_table, want_t.seed_table
    key = lambda t: np.sort((t["location"][t["location"] != 0xFFFFFFFF].astype(np.int64) << 1) | (t["flags"][t["location"] != 0xFFFFFFFF] & 1))
    assert np.array_equal(key(gv), key(wv))
    assert np.array_equal((gv["location"] != 0xFFFFFFFF) & ((gv["flags"] & 2) == 0), (wv["location"] != 0xFFFFFFFF) & ((wv["flags"] & 2) == 0))
    # the reference's own probe on the device-built table
    if ol.ref_api("emf", "sse41") is not None:
        path, tpath = str(tmp_path / "dev.perfect"), str(tmp_path / "ref.0123")
        got_t.save(path)
        text.tofile(tpath)
        r = ol.RefEmf(path, tpath)
        rret, rperf = r.probe(flat, off, ln)
        r.close()
        assert np.array_equal(rret, got[0])
        assert np.array_equal(rperf[:, 0], got[1]["flags"]) and np.array_equal(rperf[:, 1], got[1]["location"])
    pm.unload()
