"""GPU parity, BASELINE configs[3]: one long record cut into overlapping 50-s windows (mode N).

Window ``w`` = samples ``[w*stride, w*stride + 10000)`` (the window ``cfg_utils.py:369-371`` cuts,
taken every ``stride`` samples).  The sliding front door reads the windows in place; every image
must equal ``eeg_to_spectrogram`` of the materialised slice BIT FOR BIT and the oracle to 1e-3."""
import numpy as np
import pytest
import torch

from conftest import eeg_like
from oracle import melspec_oracle as mo

import b200spec

pytestmark = pytest.mark.gpu
ATOL = 1e-3
L_HOUR = 720_000


@pytest.fixture(scope="module")
def hour():
    x = eeg_like(np.random.default_rng(404), (20, L_HOUR))
    x[3, 100_000:100_050] = np.nan          # holes inside some windows only
    x[17, 500_000:530_000] = np.nan         # a dead electrode for 150 s
    return x


def test_window_count():
    assert b200spec.sliding_window_count(L_HOUR, 10000, 400) == 1776      # SURVEY 8d cfg 4
    assert b200spec.sliding_window_count(10000) == 1 and b200spec.sliding_window_count(9999) == 0
    assert b200spec.sliding_window_count(10399) == 1 and b200spec.sliding_window_count(10400) == 2


def test_hour_record_equals_materialised_windows(hour):
    xd = torch.from_numpy(hour).cuda()
    got = b200spec.eeg_to_spectrogram_sliding(xd)
    torch.cuda.synchronize()
    assert got.shape == (1776, 128, 256, 4) and got.dtype == torch.float32
    picks = [0, 1, 7, 249, 250, 251, 887, 1249, 1300, 1774, 1775]          # first, last, NaN regions, odd offsets
    mat = torch.stack([xd[:, 400 * w:400 * w + 10000] for w in picks]).contiguous()
    ref = b200spec.eeg_to_spectrogram(mat)
    assert torch.equal(got[picks], ref)
    want = mo.eeg_to_spectrogram_batch_oracle(hour[None, :, 400 * 250:400 * 250 + 10000])
    assert np.abs(got[250].cpu().numpy() - want[0]).max() <= ATOL
    want = mo.eeg_to_spectrogram_batch_oracle(hour[None, :, 400 * 1775:400 * 1775 + 10000])
    assert np.abs(got[1775].cpu().numpy() - want[0]).max() <= ATOL
    assert torch.isfinite(got).all()


def test_time_major_and_host_tensor(hour):
    part = np.ascontiguousarray(hour[:, 90_000:131_000])                  # 41 000 samples -> 78 windows
    n = b200spec.sliding_window_count(part.shape[1])
    assert n == 78
    dev = b200spec.eeg_to_spectrogram_sliding(torch.from_numpy(part).cuda())
    tm = b200spec.eeg_to_spectrogram_sliding(torch.from_numpy(np.ascontiguousarray(part.T)).cuda(), time_major=True)
    host = b200spec.eeg_to_spectrogram_sliding(torch.from_numpy(part))
    assert not host.is_cuda and host.shape == (78, 128, 256, 4)
    assert torch.equal(dev, tm) and torch.equal(dev.cpu(), host)


@pytest.mark.parametrize("stride", [4, 39, 401, 10000, 12345])
def test_strides(hour, stride):
    """Strides that are not a multiple of 4 samples leave the TMA-aligned engine (still within tolerance)."""
    part = torch.from_numpy(np.ascontiguousarray(hour[:, :40_000])).cuda()
    k = min(4, b200spec.sliding_window_count(40_000, 10000, stride))
    got = b200spec.eeg_to_spectrogram_sliding(part[:, :10000 + stride * (k - 1)], stride=stride)
    assert got.shape[0] == k
    mat = torch.stack([part[:, stride * w:stride * w + 10000] for w in range(k)]).contiguous()
    ref = b200spec.eeg_to_spectrogram(mat)
    assert (got - ref).abs().max().item() <= (0.0 if stride % 4 == 0 else ATOL)


def test_short_record_and_bad_arguments():
    x = torch.zeros(20, 9999, device="cuda")
    assert b200spec.eeg_to_spectrogram_sliding(x).shape == (0, 128, 256, 4)
    with pytest.raises(ValueError):
        b200spec.eeg_to_spectrogram_sliding(torch.zeros(2, 20, 10000, device="cuda"))
    with pytest.raises(ValueError):
        b200spec.eeg_to_spectrogram_sliding(torch.zeros(20, 10000, device="cuda"), stride=0)
