import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    o = entry.load_oracle()
    o.build()
    return o


@pytest.fixture(scope="session")
def pkg():
    return entry.load_package()


@pytest.fixture(scope="session")
def capi(pkg):
    """The ctypes binding with the shared library loaded.  A missing .so is built (nvcc) — never
    skipped or substituted: the product has no fallback, so if it cannot be built the test fails."""
    if not os.path.exists(pkg.capi.LIB_PATH):
        import subprocess
        subprocess.check_call(["make", "-s", "-C", entry.PKG_DIR, "lib", "host"])
    pkg.capi.lib()
    return pkg.capi


@pytest.fixture(scope="session")
def gpu(capi):
    n = capi.device_count()
    if n < 1:
        pytest.fail("test is marked gpu but no CUDA device is visible")
    return n


GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(GOLDEN_DIR, "golden.json")) as f:
        return json.load(f)
