import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLD = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


def load_golden(name: str):
    """-> (x, state_dict, outputs) as torch tensors, from tests/golden/<name>.npz
    (written by oracle/pin_against_reference.py from the REFERENCE's own code)."""
    z = np.load(GOLD / f"{name}.npz")
    x = torch.from_numpy(z["x"])
    sd = {k[2:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("w:")}
    outs = {k[4:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("out:")}
    return x, sd, outs


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(scope="session")
def lib_built():
    """Builds the C-ABI library if it is missing (nvcc cross-compiles without a GPU)."""
    from etss_07_imageenhance_b200.build import build_library
    return build_library()


@pytest.fixture(scope="session")
def cuda_dev(lib_built):
    if not torch.cuda.is_available():
        pytest.fail("a -m gpu test was selected but CUDA is not available")
    import etss_07_imageenhance_b200 as P
    P.ops.require_device()
    return torch.device("cuda:0")
