"""pytest configuration: registers the `gpu` marker and puts the host package + oracle on sys.path.

`-m "not gpu"`: oracle vs golden vectors / OpenBLAS, host logic, C-ABI symbol surface.
`-m gpu`      : parity tests proper -- the CUDA path through the C ABI against the oracle.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "accelerate-blas_b200", "python"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test selected but no CUDA device is visible")
    return torch.device("cuda:0")
