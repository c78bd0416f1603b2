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
    if not os.path.exists(os.path.join(ROOT, "oracle", "librcull_oracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    lib = os.path.join(ROOT, "renderer_b200", "librcull.so")
    if not os.path.exists(lib) and os.path.exists("/usr/local/cuda/bin/nvcc"):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "renderer_b200", "csrc")])
    yield
