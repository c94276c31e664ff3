"""Shared case table for the banded-SW tests (same list as tests/golden/make_golden.py)."""
import importlib.util
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
_mg = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_mg)
CASES = _mg.CASES
GOLDEN = np.load(os.path.join(HERE, "golden", "bsw_golden.npz"))


def results_matrix(pairs, fields):
    return np.stack([pairs[f] for f in fields], axis=1).astype(np.int64)
