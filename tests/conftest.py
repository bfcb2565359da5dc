"""pytest configuration: the `gpu` marker, import paths, and the oracle build.

`-m "not gpu"` tests: oracle vs numpy / golden vectors, host-side logic, C-ABI symbol table.
`-m gpu` tests: parity of the CUDA path (through the C ABI) against the oracle.
"""
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
def oracle():
    import oracle_lib as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def cuda_lib():
    """libcrest_cuda.so, built on demand (nvcc cross-compiles without a GPU)."""
    from crest_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import subprocess
        subprocess.run(["make", "-s", "-j", str(max(1, os.cpu_count() or 1)), "-C",
                        os.path.join(ROOT, "crest_b200", "csrc")], check=True, stdout=subprocess.DEVNULL)
    return _lib.lib()


def has_gpu():
    try:
        from crest_b200 import _lib
        return _lib.lib().crest_cuda_device_count() > 0
    except Exception:
        return False


GOLDEN = os.path.join(ROOT, "tests", "golden")
