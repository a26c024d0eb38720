import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def built_lib():
    """The product library; built on demand (nvcc cross-compiles without a GPU)."""
    from byox_vllm_b200 import build

    return build.build()


@pytest.fixture(scope="session")
def cuda_dev(built_lib):
    """GPU tests must run the CUDA path or fail: no skip, no fallback."""
    import torch

    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    from byox_vllm_b200 import _lib

    _lib.load()
    return torch.device("cuda:0")
