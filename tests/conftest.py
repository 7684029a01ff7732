import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built_artefacts():
    """The built artefacts are not in the repository: build what is missing (nvcc cross-compiles without a GPU).
    The oracle library builds itself on first use (tests/oracle_lib.py)."""
    need = [os.path.join(ROOT, "fdtd3d_b200", "libfdtd3d_b200.so"), os.path.join(ROOT, "fdtd3d_b200", "bin", "fdtd3d")]
    if not all(os.path.exists(p) for p in need):
        from fdtd3d_b200 import build as b
        b.build()
