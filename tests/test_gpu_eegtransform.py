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


def test_fused_transform_stft_equals_the_two_calls():
    """eeg_transform_stft (one C-ABI call) == transform + compute_stft, bit for
    bit, and both match the two lifted reference pieces chained (oracle ports) within 1e-3."""
    rng = np.random.default_rng(35)
    x = (np.cumsum(rng.standard_normal((300, 10000, 19)), 1) + rng.standard_normal((300, 10000, 19)) * 30).astype(np.float32)

    class Cfg:
        diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}

    xt = torch.from_numpy(x).cuda()
    tr = b200spec.EEGTransformer(n_feats=19, downsample=5, channel_major=True)
    fused = b200spec.eeg_transform_stft(xt, Cfg)
    two = b200spec.compute_stft(tr.transform(xt), Cfg)
    assert fused.shape == (300, 19, 33, 2000) and torch.equal(fused, two)
    from oracle import stft_oracle
    ref_sig = np.transpose(eo.eeg_transform_batch_oracle(x[:2], n_feats=19, downsample=5), (0, 2, 1)).astype(np.float32)
    want = stft_oracle.compute_stft_port(ref_sig, Cfg)
    assert np.abs(fused[:2].cpu().numpy() - want).max() <= 1e-3


def test_errors():
    with pytest.raises(ValueError):
        b200spec.EEGTransformer(n_feats=19).transform(torch.zeros(100, 10, device="cuda"))      # too few channels
    with pytest.raises(ValueError):
        b200spec.EEGTransformer(n_feats=8, apply_chris_magic_ch8=True).transform(torch.zeros(500, 8, device="cuda"))
    with pytest.raises(NotImplementedError):
        b200spec.EEGTransformer(n_feats=19, downsample=16).transform(torch.zeros(1000, 20, device="cuda"))


def test_span_cuts_do_not_change_a_bit():
    """The default configuration runs the lane kernel, which cuts a record into spans to fill whole
    waves: the cut depends on the batch size.  The state at a chunk start is rebuilt from the
    zero-state ends of the three chunks before it, warm-up chunks included, so a record transformed
    alone (many spans), in a batch of 5 or in a batch of 300 (one span) is the same bit for bit --
    for the (L, 20) parquet rows and for the reference's (L, 19) windows."""
    rng = np.random.default_rng(33)
    x = (np.cumsum(rng.standard_normal((5, 10000, 20)), 1) + rng.standard_normal((5, 10000, 20)) * 30).astype(np.float32)
    x[1, 4000:4100, 3] = np.nan
    x[2, :, 7] = 5000.0                                   # clipped to 1024
    for cols in (20, 19):
        xt = torch.from_numpy(np.ascontiguousarray(x[:, :, :cols])).cuda()
        big = xt.repeat(60, 1, 1)
        for kw in (dict(n_feats=19, downsample=5), dict(n_feats=19, downsample=5, channel_major=True)):
            t = b200spec.EEGTransformer(**kw)
            cut = t.transform(xt)
            for r in range(5):
                assert torch.equal(cut[r], t.transform(xt[r]))
            assert torch.equal(t.transform(big)[295:300], cut)
            want = eo.eeg_transform_batch_oracle(x[:, :, :cols], n_feats=19, downsample=5)
            got = cut.transpose(1, 2) if kw.get("channel_major") else cut
            assert np.abs(got.cpu().numpy() - want).max() <= ATOL
    x8 = np.ascontiguousarray(x[:, :, :8])
    t8 = b200spec.EEGTransformer(n_feats=8, apply_chris_magic_ch8=True, downsample=5)
    got = t8.transform(x8)
    want = eo.eeg_transform_batch_oracle(x8, n_feats=8, apply_chris_magic_ch8=True, downsample=5)
    assert np.abs(got - want).max() <= ATOL


def test_window19_runs_without_a_padding_copy():
    """(B, 10000, 19) goes straight into the kernel: one launch, no torch kernel in front of it."""
    x = torch.randn(8, 10000, 19, device="cuda") * 40
    t = b200spec.EEGTransformer(n_feats=19, downsample=5)
    t.transform(x)
    torch.cuda.synchronize()
    before = b200spec.launch_count()
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
        y = t.transform(x)
        torch.cuda.synchronize()
    assert b200spec.launch_count() - before == 1 and y.shape == (8, 2000, 19)
    kernels = [e.key for e in prof.key_averages() if e.device_type == torch.autograd.DeviceType.CUDA]
    assert all("eeg_transform" in k for k in kernels), kernels


@pytest.mark.parametrize("length", [40, 1234, 9999, 10003])
def test_ragged_lengths(length):
    rng = np.random.default_rng(length)
    x = (rng.standard_normal((3, length, 20)) * 40).astype(np.float32)
    got = b200spec.EEGTransformer(n_feats=19, downsample=5).transform(x)
    want = eo.eeg_transform_batch_oracle(x, n_feats=19, downsample=5)
    assert got.shape == want.shape and np.abs(got - want).max() <= ATOL


def test_slowly_decaying_filter_takes_the_exact_scan():
    """Poles at radius 0.995: what a chunk leaves behind does not die within three chunks, so the
    launch must pick the block-scan kernel (exact carry) instead of the truncating span kernel."""
    from scipy.signal import lfilter
    r, th = 0.995, 0.3
    sec = np.array([1.0, -2 * r * np.cos(th), r * r])
    a = np.convolve(sec, sec)
    b = np.array([1e-4, 0, 0, 0, 0])
    rng = np.random.default_rng(34)
    x = (rng.standard_normal((2, 10000, 20)) * 40).astype(np.float32)
    t = b200spec.EEGTransformer(n_feats=19, downsample=5)
    for i in range(5):
        t.params.b[i], t.params.a[i] = b[i], a[i]
    got = t.transform(x)
    want = lfilter(b, a, np.clip(x[:, :, :19], -1024, 1024) / 32.0, axis=1)[:, ::5]
    assert np.abs(got - want).max() <= 2e-3 * max(1.0, np.abs(want).max())
