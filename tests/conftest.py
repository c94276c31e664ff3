import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built_lib():
    """libbwagpu.so, built on demand (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g
    from bwa_mem_scale_b200 import lib
    if not os.path.exists(lib.LIB_PATH):
        g.build()
    return lib.load()


@pytest.fixture(scope="session")
def gpu_ctx(built_lib):
    from bwa_mem_scale_b200 import Context
    ctx = Context(0)          # raises loudly when there is no device: no CPU fallback
    yield ctx
    ctx.close()
