import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    import helpers
    return helpers.load_oracle()


@pytest.fixture(scope="session")
def nsof():
    import nsof_b200
    return nsof_b200


@pytest.fixture(scope="session")
def emu_engine(nsof):
    """Engine over the TEST-ONLY block-emulator build of the kernel sources (tests/emu)."""
    import helpers
    lib = nsof.Library(helpers.build_emu())
    eng = nsof.Engine(0, lib=lib, _emulated=True)
    yield eng
    eng.close()


@pytest.fixture(scope="session")
def gpu_engine(nsof):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    eng = nsof.Engine(0)
    yield eng
    eng.close()


@pytest.fixture(scope="session")
def golden():
    import helpers
    return helpers.Golden()
