import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def lib():
    """The C-ABI library, built in-tree if needed.  GPU tests must run through it."""
    from miraculix_b200 import api, build
    build.build()
    return api.load_library()


@pytest.fixture(scope="session")
def gpu_lib(lib):
    import ctypes
    try:
        cudart_ok = True
        import torch  # only used to ask whether a device exists
        cudart_ok = torch.cuda.is_available()
    except Exception:
        cudart_ok = False
    if not cudart_ok:
        pytest.fail("GPU test selected but no CUDA device is visible: the product has no CPU path")
    # The GPU suite's small shapes are there to exercise the tensor-core kernels and the host
    # pipeline on ragged edges; by default the host-pointer call would route them to the one-kernel
    # latency path (csrc/dgemm_small.cu).  That path has its own file, tests/test_gpu_dgemm_small.py,
    # which switches it on and back off; smoke() and bench.py run the default configuration.
    lib.mb200_set_dgemm_small_limit(0.0)
    return lib
