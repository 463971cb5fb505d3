import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build libsr_cuda.so and the oracle if they are missing (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g
    lib = os.path.join(ROOT, "software_renderer_b200", "libsr_cuda.so")
    orc = os.path.join(ROOT, "oracle", "libsr_oracle.so")
    if not (os.path.exists(lib) and os.path.exists(orc)):
        g.build()


@pytest.fixture(scope="session")
def renderer():
    import software_renderer_b200 as srb
    r = srb.CudaRenderer(0)
    yield r
    r.close()
