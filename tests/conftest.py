import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = Path(__file__).resolve().parent / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def synth_image(seed: int, h: int, w: int) -> np.ndarray:
    """Same generator as tests/golden/make_golden.py (inputs are seeded, only outputs are stored)."""
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w]
    grad = ((yy * 255 // max(h - 1, 1)) ^ (xx * 255 // max(w - 1, 1))).astype(np.uint8)
    img[: h // 3, :, 1] = grad[: h // 3]
    img[h // 2: h // 2 + max(h // 8, 1), : w // 2] = 255
    img[h // 2: h // 2 + max(h // 8, 1), w // 2:] = 0
    return img


@pytest.fixture(scope="session")
def golden_pre():
    return np.load(GOLDEN / "preprocess.npz")


@pytest.fixture(scope="session")
def golden_cls():
    return np.load(GOLDEN / "classify.npz")


@pytest.fixture(scope="session")
def golden_enc():
    return np.load(GOLDEN / "encoder.npz")
