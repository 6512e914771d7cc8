import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: long-running")


@pytest.fixture(scope="session")
def built_lib():
    """libb200tsdf.so, (re)built when stale; nvcc cross-compiles without a GPU."""
    from mast3r_slam_quality_dualtsdf_b200 import build

    return build.build()


@pytest.fixture(scope="session")
def cuda_device(built_lib):
    import torch

    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test selected but no CUDA device is visible")
    return torch.device("cuda:0")
