import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The oracle is cheap to build anywhere (gcc); librcull.so needs nvcc and is
    built by __graft_entry__.build() -- here only if it is missing and nvcc exists."""
    oracle_so = os.path.join(ROOT, "oracle", "librcull_oracle.so")
    rc = subprocess.call(["make", "-s", "-C", os.path.join(ROOT, "oracle")])  # no-op when up to date
    if rc != 0 and not os.path.exists(oracle_so):
        raise RuntimeError("the CPU oracle could not be built (make -C oracle)")
    lib = os.path.join(ROOT, "renderer_b200", "librcull.so")
    if not os.path.exists(lib) and os.path.exists("/usr/local/cuda/bin/nvcc"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "renderer_b200", "csrc")])
    yield
