import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (oracle/libfhe_oracle.so), built on demand."""
    from tests import oracle_py
    return oracle_py.load()


@pytest.fixture(scope="session")
def pkg():
    lib = os.path.join(ROOT, "upmem--openfhe_b200", "lib", "libfhe_b200.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-C", ROOT, "lib"])
    return importlib.import_module("upmem--openfhe_b200")


@pytest.fixture(scope="session")
def backend(pkg):
    be = pkg.Backend(0)
    yield be
    be.close()
