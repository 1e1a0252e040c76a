import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session", autouse=True)
def built_library():
    """The C-ABI library normally arrives prebuilt (build.sh / __graft_entry__.build()); if this checkout
    has none and nvcc is on the box, build it once -- the tests never run against anything else."""
    lib = os.path.join(ROOT, "terraform-provider-iterative_b200", "libtpimd5.so")
    if not os.path.exists(lib):
        import shutil
        import subprocess
        if shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc"):
            subprocess.check_call(["bash", os.path.join(ROOT, "build.sh")], cwd=ROOT)
    yield
