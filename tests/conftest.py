import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def oracle():
    """the CPU oracle (tests are the only place allowed to use it)."""
    import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def ref_cpu(oracle):
    r = oracle.ref_cpu()
    if r is None:
        pytest.skip("oracle/_ref/libcoloc_ref_cpu.so not built (needs /root/reference at build time)")
    return oracle


@pytest.fixture(scope="session")
def clb():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import coloc_b200
    coloc_b200.lib()          # raises if the extension is missing: no fallback
    return coloc_b200
