import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def lib():
    """The C-ABI library, built in-tree if needed (nvcc cross-compiles without a GPU)."""
    from optical_flow_b200 import _lib, build
    build.build_library()
    return _lib.load()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from optical_flow_b200 import build
    build.build_library()
    return torch.device("cuda:0")


@pytest.fixture
def variant():
    """``variant(name, value)`` selects a kernel variant through of_set_option (the library reads no environment
    variables); everything set during the test is unset afterwards."""
    from optical_flow_b200 import _lib
    lib = _lib.load()
    used = []

    def set_(name, value="1"):
        used.append(name)
        _lib.check(lib.of_set_option(name.encode(), None if value is None else str(value).encode()), "of_set_option")

    yield set_
    for name in used:
        lib.of_set_option(name.encode(), None)
