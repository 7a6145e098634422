import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    import torch
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden_warp():
    return np.load(os.path.join(GOLDEN, "warp_reference.npz"))


@pytest.fixture(scope="session")
def golden_corr():
    return np.load(os.path.join(GOLDEN, "corr_torchvision.npz"))


@pytest.fixture(scope="session")
def golden_upsample():
    return np.load(os.path.join(GOLDEN, "upsample_reference.npz"))


@pytest.fixture(scope="session")
def golden_utils_warp():
    return np.load(os.path.join(GOLDEN, "utils_warp_reference.npz"))


@pytest.fixture(scope="session")
def golden_ncc():
    return np.load(os.path.join(GOLDEN, "ncc_reference.npz"))


NCC_CASES = ["ncc_noise_40", "ncc_smooth_48x37", "ncc_shift_64", "ncc_tiny_5x7", "ncc_same_33x70"]
UTILS_WARP_CASES = ["uw_24", "uw_17x40", "uw_oob_16", "uw_zero_12"]
UPSAMPLE_CASES = ["u_6x5", "u_12x12", "u_9x33"]
WARP_CASES = ["zero_24", "integer_32x40", "half_16", "noise_48", "smooth_64x48", "oob_20"]
CORR_CASES = ["c16_16x16", "c8_24x24", "c32_16x24"]


def rel_err(a, b):
    """max |a-b| / max(|b|, tiny): the norm every tolerance in tests/ is stated in."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))
