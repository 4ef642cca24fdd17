import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:  # pragma: no cover
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def stft_golden():
    return np.load(os.path.join(GOLDEN_DIR, "stft_golden.npz"))


@pytest.fixture(scope="session")
def melspec_golden():
    return np.load(os.path.join(GOLDEN_DIR, "melspec_golden.npz"))


@pytest.fixture(scope="session")
def fronthalf_golden():
    return np.load(os.path.join(GOLDEN_DIR, "fronthalf_golden.npz"))


@pytest.fixture(scope="session")
def melspec_xcheck():
    return np.load(os.path.join(GOLDEN_DIR, "melspec_xcheck.npz"))


def eeg_like(rng, shape, walk=0.3, noise=1.0):
    """Synthetic EEG of SURVEY.md section 8d: random walk + 10 Hz sine + white noise, clipped."""
    x = np.cumsum(rng.standard_normal(shape), axis=-1) * walk
    t = np.arange(shape[-1]) / 200.0
    x = x + 10.0 * np.sin(2 * np.pi * 10.0 * t) + noise * rng.standard_normal(shape)
    return np.clip(x, -1024, 1024).astype(np.float32)


class Cfg:
    """Stand-in for the reference's CFG: only ``.diffEEG_trainer[...]`` is read."""

    def __init__(self, n_fft=64, hop_length=32, window="hann"):
        self.diffEEG_trainer = {"stft_n_fft": n_fft, "stft_hop_length": hop_length,
                                "stft_window": window}


@pytest.fixture(scope="session")
def eegtransform_golden():
    return np.load(os.path.join(GOLDEN_DIR, "eegtransform_golden.npz"))
