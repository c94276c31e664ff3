"""GPU parity tests (through the C ABI): CIGAR / NM / MD of final alignment regions (bwagpu_reg2aln_batch) against
records of the reference's own mem_reg2aln."""
import numpy as np
import pytest

import c2a_cases as cc
import cigar_cases as gc
import oracle_lib as ol
from bwa_mem_scale_b200 import Extender, RegToAln, cigar_opt, fmindex, pack_reads

pytestmark = pytest.mark.gpu
GOLD = np.load(gc.GOLDEN)


@pytest.fixture(scope="module")
def setup(gpu_ctx):
    ref, ref_string = cc.genome()
    Extender(ref_string, len(ref), [0], [len(ref)], ctx=gpu_ctx)
    flat, off, ln = pack_reads(gc.golden_reads(ref))
    return ref, RegToAln(gpu_ctx), flat, off, ln


@pytest.mark.parametrize("case", gc.GOLDEN_CASES, ids=[c[0] for c in gc.GOLDEN_CASES])
def test_gpu_cigar_matches_reference_records(setup, case):
    name, kw = case
    ref, r2a, flat, off, ln = setup
    o = ol.mem_opt(**kw)
    outs, cigar, md = r2a.mem_reg2aln(cigar_opt(o["a"], o["b"], o["o_del"], o["e_del"], o["o_ins"], o["e_ins"], o["w"]),
                                      flat, off, ln, GOLD[f"{name}_jobs"])
    assert np.array_equal(gc.out_matrix(outs), GOLD[f"{name}_outs"])
    assert np.array_equal(cigar, GOLD[f"{name}_cigar"]) and np.array_equal(md, GOLD[f"{name}_md"])


def test_gpu_cigar_empty_rejected_and_errors(setup, gpu_ctx):
    from bwa_mem_scale_b200 import BwaGpuError, CIGAR_JOB_DTYPE
    ref, r2a, flat, off, ln = setup
    outs, cigar, md = r2a.mem_reg2aln(cigar_opt(), flat, off, ln, np.zeros(0, CIGAR_JOB_DTYPE))
    assert len(outs) == 0 and len(cigar) == 0 and len(md) == 0
    jobs = np.zeros(2, CIGAR_JOB_DTYPE)
    jobs["rb"], jobs["re"], jobs["qe"], jobs["truesc"], jobs["w"] = [len(ref) - 50, 100], [len(ref) + 50, 100], [100, 50], 50, 10
    outs, cigar, md = r2a.mem_reg2aln(cigar_opt(), flat, off, ln, jobs)
    assert list(outs["ok"]) == [0, 0] and len(cigar) == 0
    jobs["read"] = len(ln)                                   # read index out of range
    with pytest.raises(BwaGpuError):
        r2a.mem_reg2aln(cigar_opt(), flat, off, ln, jobs)


@pytest.mark.parametrize("case", [gc.GOLDEN_CASES[0], gc.GOLDEN_CASES[3]], ids=["default", "retry"])
def test_gpu_cigar_device_resident_form(setup, gpu_ctx, case):
    name, kw = case
    ref, r2a, flat, off, ln = setup
    o = ol.mem_opt(**kw)
    jobs = np.ascontiguousarray(GOLD[f"{name}_jobs"])
    d = [gpu_ctx.upload(a) for a in (flat, off, ln, jobs)]
    res = r2a.mem_reg2aln_device(cigar_opt(o["a"], o["b"], o["o_del"], o["e_del"], o["o_ins"], o["e_ins"], o["w"]),
                                 d[0], d[1], d[2], len(ln), d[3], len(jobs))
    outs, cigar, md = r2a.fetch(res)
    assert np.array_equal(gc.out_matrix(outs), GOLD[f"{name}_outs"])
    assert np.array_equal(cigar, GOLD[f"{name}_cigar"]) and np.array_equal(md, GOLD[f"{name}_md"])
    assert 0 < res.n_gapfree < len(jobs)
    for p in d:
        gpu_ctx.free(p)
