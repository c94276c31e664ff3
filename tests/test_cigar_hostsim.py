"""CPU tests: the per-thread CIGAR / NM / MD code of the CUDA kernel (csrc/cigar_core.cuh: mem_reg2aln's band loop,
bwa_gen_cigar2, ksw_global2 with backtrack), run on the host, against records of the reference's own mem_reg2aln
(golden; live against oracle/_ref when present)."""
import os
import tempfile

import numpy as np
import pytest

import cigar_cases as gc
import fmi_cases as fc
import hostsim_lib
import oracle_lib as ol
from bwa_mem_scale_b200 import cigar_opt, fmindex

GOLD = np.load(gc.GOLDEN)


def copt_of(o):
    return cigar_opt(o["a"], o["b"], o["o_del"], o["e_del"], o["o_ins"], o["e_ins"], o["w"])


@pytest.mark.parametrize("case", gc.GOLDEN_CASES, ids=[c[0] for c in gc.GOLDEN_CASES])
def test_cigar_code_matches_reference_records(case):
    name, kw = case
    ref, _ = fc.small_index()
    flat, off, ln = ol.pack_reads(gc.golden_reads(ref))
    assert len(ln) == int(GOLD["n_reads"])
    jobs = GOLD[f"{name}_jobs"]
    outs, cigar, md = hostsim_lib.reg2aln(hostsim_lib.load(), copt_of(ol.mem_opt(**kw)), len(ref), fmindex.text_with_revcomp(ref),
                                          flat, off, ln, jobs)
    assert np.array_equal(gc.out_matrix(outs), GOLD[f"{name}_outs"])
    assert np.array_equal(cigar, GOLD[f"{name}_cigar"]) and np.array_equal(md, GOLD[f"{name}_md"])
    if name == "retry":
        assert hostsim_lib.reg2aln.last_redo > 100          # the second pass (full rows) is exercised
    # properties: every CIGAR consumes the whole read and exactly the window minus a squeezed outer deletion
    for j in np.random.default_rng(0).choice(len(jobs), 500, replace=False):
        c = cigar[outs["cigar_off"][j]: outs["cigar_off"][j] + outs["n_cigar"][j]]
        op, ln_ = c & 0xf, c >> 4
        assert ln_[(op == 0) | (op == 1) | (op == 3)].sum() == ln[jobs["read"][j]]
        assert ln_[(op == 0) | (op == 2)].sum() <= jobs["re"][j] - jobs["rb"][j]
        assert np.all(op[1:] != op[:-1])


def test_rejected_windows():
    """Where the reference returns a NULL cigar (window bridging the strands, empty query) ok = 0."""
    ref, _ = fc.small_index()
    flat, off, ln = ol.pack_reads(gc.golden_reads(ref)[:4])
    jobs = np.zeros(3, ol.CIGAR_JOB_DTYPE)
    jobs["rb"], jobs["re"], jobs["qb"], jobs["qe"] = [len(ref) - 50, 100, 2 * len(ref) - 20], [len(ref) + 50, 100, 2 * len(ref) + 5], 0, [100, 50, 25]
    jobs["truesc"], jobs["w"] = 50, 10
    outs, cigar, md = hostsim_lib.reg2aln(hostsim_lib.load(), copt_of(ol.mem_opt()), len(ref), fmindex.text_with_revcomp(ref),
                                          flat, off, ln, jobs)
    assert list(outs["ok"]) == [0, 0, 0] and len(cigar) == 0


@pytest.mark.skipif(ol.ref_binary() is None, reason="oracle/_ref not built")
def test_cigar_code_live_against_reference():
    ref, idx = fc.small_index()
    rng = np.random.default_rng(77)
    from bwa_mem_scale_b200 import synth
    base, _, _ = synth.simulate_reads(ref, 600, seed=78, err=0.03)
    reads = [gc._indel(rng, r, 1 + i % 4) for i, r in enumerate(base)]
    flat, off, ln = ol.pack_reads(reads)
    o = ol.mem_opt(w=30)
    with tempfile.TemporaryDirectory() as td:
        prefix = os.path.join(td, "ref")
        fmindex.write_index_files(prefix, ref, idx)
        rm = ol.RefMem(prefix)
        chains, seeds, regs, _ = rm.core(o, flat, off, ln)
        jobs = gc.jobs_from_regs(regs, chains, ol.CIGAR_JOB_DTYPE)
        want = ol.reference_reg2aln(rm, o, flat, off, ln, jobs)
    got = hostsim_lib.reg2aln(hostsim_lib.load(), copt_of(o), len(ref), fmindex.text_with_revcomp(ref), flat, off, ln, jobs)
    assert np.array_equal(gc.out_matrix(got[0]), gc.out_matrix(want[0]))
    assert np.array_equal(got[1], want[1]) and np.array_equal(got[2], want[2])
