"""GPU parity, row (f)1: ``EEGTransformer.transform`` against the reference's own outputs.

Integer facts (shape, channel selection, decimation phase) are exact.  Values: the reference
filters in float64, the kernel runs the recurrence in float32 with double-precision state
propagation; tolerance |gpu - reference| <= 2e-4 on values of magnitude <= 32 (measured ~1e-5)."""
import numpy as np
import pytest
import torch

import b200spec
from oracle import eegtransform_oracle as eo
from test_oracle_eegtransform import CASES

pytestmark = pytest.mark.gpu
ATOL = 2e-4


@pytest.mark.parametrize("name", sorted(CASES))
def test_against_reference_golden(eegtransform_golden, name):
    x, want = eegtransform_golden[f"{name}/x"], eegtransform_golden[f"{name}/y"]
    got = b200spec.EEGTransformer(**CASES[name]).transform(x)
    assert isinstance(got, np.ndarray) and got.dtype == np.float32 and got.shape == want.shape
    if name == "raw_scale":                      # un-normalised: magnitudes of a few hundred
        np.testing.assert_allclose(got, want, rtol=2e-5, atol=ATOL)
    else:
        err = np.abs(got - want).max()
        assert err <= ATOL, f"{name}: max abs err {err:.3e}"


def test_batch_tensor_layouts_and_determinism():
    rng = np.random.default_rng(31)
    x = (np.cumsum(rng.standard_normal((37, 10000, 20)), 1) * 2 + rng.standard_normal((37, 10000, 20)) * 30).astype(np.float32)
    want = eo.eeg_transform_batch_oracle(x[:4], n_feats=19, downsample=5)
    xt = torch.from_numpy(x).cuda()
    t = b200spec.EEGTransformer(n_feats=19, downsample=5)
    got = t.transform(xt)
    assert got.is_cuda and got.shape == (37, 2000, 19) and got.dtype == torch.float32
    assert np.abs(got[:4].cpu().numpy() - want).max() <= ATOL
    assert torch.equal(got, t.transform(xt))
    cm = b200spec.EEGTransformer(n_feats=19, downsample=5, channel_major=True).transform(xt)
    assert cm.shape == (37, 19, 2000) and torch.equal(cm.transpose(1, 2), got)
    one = t.transform(xt[5])
    assert one.shape == (2000, 19) and torch.equal(one, got[5])          # records independent
    cpu = t.transform(xt[:2].cpu())
    assert cpu.device.type == "cpu" and torch.equal(cpu, got[:2].cpu())


def test_feeds_compute_stft():
    """The DiffEEG path end to end on the device: raw EEG -> transformer -> compute_stft image."""
    rng = np.random.default_rng(32)
    x = (rng.standard_normal((3, 10000, 20)) * 50).astype(np.float32)

    class Cfg:
        diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}

    sig = b200spec.EEGTransformer(n_feats=19, downsample=5, channel_major=True).transform(torch.from_numpy(x).cuda())
    img = b200spec.compute_stft(sig, Cfg)
    assert img.shape == (3, 19, 33, 2000)
    from oracle import stft_oracle
    ref_sig = np.transpose(eo.eeg_transform_batch_oracle(x, n_feats=19, downsample=5), (0, 2, 1)).astype(np.float32)
    want = stft_oracle.compute_stft_port(ref_sig, Cfg)
    assert np.abs(img.cpu().numpy() - want).max() <= 1e-3


def test_errors():
    with pytest.raises(NotImplementedError):
        b200spec.EEGTransformer(n_feats=19, apply_mu_law_encoding=True)
    with pytest.raises(ValueError):
        b200spec.EEGTransformer(n_feats=19).transform(torch.zeros(100, 10, device="cuda"))      # too few channels
    with pytest.raises(ValueError):
        b200spec.EEGTransformer(n_feats=8, apply_chris_magic_ch8=True).transform(torch.zeros(500, 8, device="cuda"))
    with pytest.raises(NotImplementedError):
        b200spec.EEGTransformer(n_feats=19, downsample=16).transform(torch.zeros(1000, 20, device="cuda"))
