import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# the parity tests read the dpt tendency back (gmc_get_field); production runs do not store it
os.environ.setdefault("GMCORE_DEBUG_FIELDS", "1")
sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # host + oracle libraries are CPU code; build them if the in-tree copies are missing
    need = [os.path.join(ROOT, "gmcore_b200/lib/libgmcore_host.so"), os.path.join(ROOT, "oracle/_build/liboracle.so"),
            os.path.join(ROOT, "oracle/_build/libfastmath_host.so")]
    if not all(os.path.exists(p) for p in need):
        subprocess.check_call(["make", "-C", ROOT, "host", "oracle"], stdout=subprocess.DEVNULL)


def has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu_required():
    if not has_gpu():
        pytest.skip("no CUDA device")
