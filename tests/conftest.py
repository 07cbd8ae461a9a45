import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "3d-gan_b200"))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu)")


@pytest.fixture(scope="session")
def oracle_path():
    """Builds (if needed) and returns the CPU oracle shared library."""
    import subprocess
    out = os.path.join(ROOT, "oracle", "_build", "libb2c_oracle.so")
    src = os.path.join(ROOT, "oracle", "b2c_oracle.c")
    if not os.path.exists(out) or os.path.getmtime(out) < os.path.getmtime(src):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["gcc", "-O2", "-fopenmp", "-shared", "-fPIC", "-o", out, src, "-lm"])
    return out


@pytest.fixture(scope="session")
def product_path():
    from b200conv3d import _lib
    if not os.path.exists(_lib.PRODUCT_LIB):
        import __graft_entry__
        __graft_entry__.build()
    return _lib.PRODUCT_LIB
