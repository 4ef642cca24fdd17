"""GPU parity, rows (f)3 / (f)4: ``process_eeg`` / ``process_spectrogram`` against the reference's own outputs.

EEG: the reference filters in float64; the kernel runs the same direct-form recurrences in float64 and keeps the
decimated rows in float32 before the z-score -- tolerance 1e-4 on z-scored values (|x| up to ~6; measured ~1e-6).
Spectrogram: fp32 intermediates, tolerance 1e-4 on the [0, 1] image.  Integer facts (row order, decimation phase,
padding, plane layout) are exact."""
import os
import sys

import numpy as np
import pytest
import torch

import b200spec
from oracle import classifier_oracle as co

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden import CLASSIFIER_EEG_CASES, CLASSIFIER_SPEC_CASES, classifier_eeg_input  # noqa: E402

pytestmark = pytest.mark.gpu
ATOL = 1e-4


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "classifier_golden.npz"))


@pytest.mark.parametrize("name", sorted(CLASSIFIER_EEG_CASES))
def test_process_eeg_against_reference_golden(gold, name):
    x = classifier_eeg_input(gold, name)
    got = b200spec.process_eeg(torch.from_numpy(x).cuda())
    torch.cuda.synchronize()
    want = gold[f"eeg/{name}/y"]
    assert got.shape == (1, 37, 3000) and got.dtype == torch.float32
    got = got.cpu().numpy()
    n_dec = len(range(0, x.shape[0] - 1, 4))
    assert np.all(got[:, :, min(n_dec, 3000):] == 0)             # pad_or_truncate
    err = np.abs(got - want).max()
    assert err <= ATOL, f"{name}: max abs err {err:.3e}"


@pytest.mark.parametrize("n", [10000, 10002, 10003, 9999, 2400, 40, 14001])
def test_process_eeg_lengths_and_batch(n):
    """n % 4 == 2 / 3: np.roll without an axis lets the last decimated sample see the next row's first samples."""
    rng = np.random.default_rng(n)
    x = (np.cumsum(rng.standard_normal((3, n, 20)), 1) * 0.5 + 5 * rng.standard_normal((3, n, 20))).astype(np.float32)
    x[1, n // 3:n // 3 + 5, 9] = np.nan
    want = np.stack([co.process_eeg_oracle(r) for r in x])
    got = b200spec.process_eeg(torch.from_numpy(x).cuda()).cpu().numpy()
    assert got.shape == want.shape == (3, 1, 37, 3000)
    assert np.abs(got - want).max() <= ATOL
    one = b200spec.process_eeg(torch.from_numpy(x[2]))             # host tensor, single record
    assert one.shape == (1, 37, 3000) and np.array_equal(one.cpu().numpy(), got[2])


def test_process_eeg_refuses_what_the_reference_mangles():
    x = torch.randn(2, 4000, 20)
    x[1, 0, 4] = float("nan")
    with pytest.raises(ValueError):
        b200spec.process_eeg(x.cuda())
    with pytest.raises(ValueError):
        co.process_eeg_oracle(x[1].numpy())


@pytest.mark.parametrize("name", sorted(CLASSIFIER_SPEC_CASES))
def test_process_spectrogram_against_reference_golden(gold, name):
    raw = torch.from_numpy(gold[f"spec/{name}/raw_f16"].astype(np.float32)).cuda()
    off = int(gold[f"spec/{name}/offset"])
    win = b200spec.spectrogram_window(raw, None if off < 0 else off)
    np.testing.assert_array_equal(win.cpu().numpy(), co.spectrogram_window(raw.cpu().numpy(), None if off < 0 else off))
    got = b200spec.process_spectrogram(win)
    torch.cuda.synchronize()
    assert got.shape == (3, 400, 300) and got.dtype == torch.float32
    assert torch.equal(got[0], got[1]) and torch.equal(got[0], got[2])
    err = np.abs(got[0].cpu().numpy() - gold[f"spec/{name}/y_plane"]).max()
    assert err <= ATOL, f"{name}: max abs err {err:.3e}"
    assert float(got.min()) >= 0.0 and float(got.max()) <= 1.0


def test_process_spectrogram_batch_and_refusal():
    rng = np.random.default_rng(12)
    win = rng.gamma(2.0, 8.0, size=(5, 400, 300)).astype(np.float32)
    win[0, 3, 7] = np.nan
    win[2, 100:110, :50] = np.nan
    win[:, 300:] = 0.0
    want = np.stack([co.process_spectrogram_oracle(w) for w in win])
    got = b200spec.process_spectrogram(torch.from_numpy(win).cuda()).cpu().numpy()
    assert got.shape == (5, 3, 400, 300) and np.abs(got - want).max() <= ATOL
    win[4, 17, :] = np.nan
    with pytest.raises(ValueError):
        b200spec.process_spectrogram(torch.from_numpy(win).cuda())
