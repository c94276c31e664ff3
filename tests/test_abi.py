"""CPU tests of the drop-in boundary: the shared library loads and exports every symbol that
include/bwagpu.h declares; with no CUDA device the product fails loudly (no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import pytest

from bwa_mem_scale_b200 import lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    src = open(os.path.join(ROOT, "include", "bwagpu.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(bwagpu_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported(built_lib):
    names = declared_functions()
    assert len(names) >= 10
    out = subprocess.run(["nm", "-D", "--defined-only", L.LIB_PATH], capture_output=True,
                         text=True, check=True).stdout
    exported = set(line.split()[-1] for line in out.splitlines() if line.strip())
    missing = [n for n in names if n not in exported]
    assert not missing, f"declared in bwagpu.h but not exported: {missing}"
    assert sorted(L.exported_symbols()) == names, "lib.py bindings out of sync with bwagpu.h"


def test_struct_layout_matches_reference():
    from bwa_mem_scale_b200.synth import SEQPAIR_DTYPE
    assert SEQPAIR_DTYPE.itemsize == 56            # sizeof(SeqPair), src/bandedSWA.h:90-99
    assert C.sizeof(L.BswParams) == 32


def test_no_device_fails_loudly(built_lib):
    if built_lib.bwagpu_device_count() > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(L.BwaGpuError, match="no CUDA device"):
        L.Context(0)


def test_product_does_not_touch_oracle():
    # the product tree must never import, link or execute anything under oracle/
    pkg = os.path.join(ROOT, "bwa_mem_scale_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle" not in txt.lower(), f"{os.path.join(dp, f)} mentions the oracle"
