"""CPU: the mode-R oracle against the reference's own outputs (golden) and against SciPy."""
import numpy as np
import pytest

from conftest import Cfg, eeg_like
from oracle import ref_extract as rx
from oracle import stft_oracle as so

CASES = ["diffeeg_64_32", "hms50s_64_32", "nfft128", "hop16", "one_window", "nfft16",
         "nfft256_odd_len", "flat_rows", "nan_sample", "decimating", "gather"]


@pytest.mark.parametrize("name", CASES)
def test_port_reproduces_reference_bit_for_bit(stft_golden, name):
    x, y = stft_golden[f"{name}/x"], stft_golden[f"{name}/y"]
    nperseg, noverlap = (int(v) for v in stft_golden[f"{name}/cfg"])
    got = so.compute_stft_port(x, Cfg(nperseg, noverlap))
    assert got.shape == y.shape and got.dtype == np.float32
    np.testing.assert_array_equal(np.isnan(got), np.isnan(y))
    np.testing.assert_array_equal(np.nan_to_num(got), np.nan_to_num(y))


@pytest.mark.parametrize("name", CASES)
def test_first_principles_restatement_matches_reference(stft_golden, name):
    x, y = stft_golden[f"{name}/x"], stft_golden[f"{name}/y"]
    nperseg, noverlap = (int(v) for v in stft_golden[f"{name}/cfg"])
    got = so.compute_stft_restated(x, nperseg, noverlap)
    np.testing.assert_array_equal(np.isnan(got), np.isnan(y))
    # flat rows divide noise by eps; everything else agrees to float rounding
    assert np.nanmax(np.abs(got - y)) <= 2e-6


def test_nan_poisons_whole_channel_only(stft_golden):
    y = stft_golden["nan_sample/y"]
    assert np.isnan(y[0, 0]).all() and not np.isnan(y[0, 1]).any()


def test_single_port(stft_golden):
    mag, _ = so.logmag_rows_port(stft_golden["single/x"][None], 64, 32)
    np.testing.assert_array_equal(mag[0], stft_golden["single/y"])
    mag2 = so.logmag_rows_restated(stft_golden["single/x"][None], 64, 32)
    assert np.abs(mag2[0] - stft_golden["single/y"]).max() <= 1e-6


@pytest.mark.parametrize("length", [64, 65, 127, 2000, 2001, 10000, 70000])
@pytest.mark.parametrize("nperseg,noverlap", [(64, 32), (128, 64), (64, 48), (256, 128), (16, 3), (64, 0)])
def test_frame_plan_equals_scipy(length, nperseg, noverlap):
    from scipy.signal import stft

    if length < nperseg:
        pytest.skip("SciPy shrinks nperseg here")
    _, t, z = stft(np.zeros(length, np.float32), fs=200, nperseg=nperseg, noverlap=noverlap)
    plan = so.frame_plan(length, nperseg, noverlap)
    assert plan["n_frames"] == z.shape[1] == len(t)
    assert plan["n_freqs"] == z.shape[0]
    np.testing.assert_allclose(t, np.arange(plan["n_frames"]) * plan["hop"] / 200.0, rtol=0, atol=1e-9)


@pytest.mark.parametrize("n_frames", [2, 3, 64, 126, 314, 1999, 2000, 2001, 2189, 22501])
def test_interp_brackets_equal_numpy_search(n_frames):
    hop = 32
    t = np.arange(n_frames) * hop / 200.0
    grid = np.linspace(0, t[-1], 2000)
    expect = np.clip(np.searchsorted(t, grid, side="right") - 1, 0, n_frames - 2)
    got = so.interp_brackets(n_frames)
    if (n_frames - 1) % 1999 == 0:
        # every column sits exactly on a knot: either neighbour yields the same value
        assert np.abs(got - expect).max() <= 1
    else:
        np.testing.assert_array_equal(got, expect)


@pytest.mark.skipif(not rx.reference_available(), reason="/root/reference not mounted")
def test_port_equals_live_reference():
    import torch

    ref = rx.load_reference_compute_stft()
    rng = np.random.default_rng(7)
    x = eeg_like(rng, (2, 19, 2000))
    want = ref(torch.from_numpy(x), Cfg()).numpy()
    np.testing.assert_array_equal(so.compute_stft_port(x, Cfg()), want)
