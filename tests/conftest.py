import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_cls():
    """The plain-C restatement (oracle/cog_oracle.c); builds on demand with gcc."""
    from oracle import port
    port.build()
    return port.OracleVec


@pytest.fixture(scope="session", params=["port", "ref"])
def gpu_oracle_cls(request):
    """Both checkers for the GPU parity tests: the C restatement (oracle/cog_oracle.c) and the compiled
    UNMODIFIED reference (oracle/_ref/libcog_ref.so); the test id says which one was used."""
    if request.param == "port":
        from oracle import port
        port.build()
        return port.OracleVec
    import parity
    return parity.ref_oracle_cls()


@pytest.fixture(scope="session")
def cuda_lib():
    """The product library. No fallback: a missing .so or device is an error, not a skip."""
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from gym_eldorado_b200 import capi
    return capi.lib()


@pytest.fixture(scope="session")
def emu():
    """TEST-ONLY build of the kernels under the lock-step warp emulator (tests/emu)."""
    import parity
    return parity.emu_lib()
