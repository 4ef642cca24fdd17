"""GPU parity, row (f)2: the spectrogram-condition front end of DiffEEG (diffEEG.py:47-53, 108-110, 136-145).

Golden vectors come from the reference's own module (tests/golden/make_golden.py::make_diffeeg): ``fp32`` = its own
ConvTranspose2d layer in fp32 + the three forward lines (tolerance 2e-5: fp32 sums of 57 products in another
order), ``hooked`` = the tensor its ``forward`` hands to ``channel_expand`` under its bf16 autocast (2e-2).
Interpolation indices are integer work: checked bit for bit against ATen through a ramp."""
import os
import sys

import numpy as np
import pytest
import torch

import b200spec
from oracle import diffeeg_oracle as do

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden import diffeeg_inputs  # noqa: E402

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "diffeeg_golden.npz"))


@pytest.mark.parametrize("case", ["ship19", "ckpt8", "train19"])
def test_against_reference_module(gold, case):
    spec, c, hidden, _ = diffeeg_inputs(case)
    partner = torch.from_numpy(gold[f"{case}/partner"]).cuda() if case == "train19" else None
    got = b200spec.spectrogram_condition(torch.from_numpy(spec).cuda(), torch.from_numpy(gold[f"{case}/weight"]),
                                         torch.from_numpy(gold[f"{case}/bias"]), 2000, partner=partner)
    torch.cuda.synchronize()
    got = got.cpu().numpy()
    assert got.shape == (spec.shape[0], hidden // 2, 2000) and got.dtype == np.float32
    assert np.abs(got - gold[f"{case}/fp32"]).max() <= 2e-5
    assert np.abs(got - gold[f"{case}/hooked"]).max() <= 2e-2        # the reference's own forward, bf16 autocast


def test_interpolation_positions_are_atens():
    """A 1-channel image whose only non-zero weight picks in[c, f, t'] makes the output a known function of the
    gathered positions: compare with the oracle (F.conv_transpose2d + F.interpolate) exactly."""
    rng = np.random.default_rng(3)
    spec = rng.random((2, 1, 33, 2000), dtype=np.float32)
    w = np.zeros((1, 3, 3, 3), np.float32)
    w[0, 0, 1, 0], w[0, 1, 1, 1], w[0, 2, 1, 2] = 1.0, 1.0, 1.0     # one kernel column each, centre row
    b = np.zeros(3, np.float32)
    want = do.spectrogram_condition_oracle(spec, w, b, 2000).numpy()
    got = b200spec.spectrogram_condition(torch.from_numpy(spec).cuda(), torch.from_numpy(w), torch.from_numpy(b), 2000)
    assert np.array_equal(got.cpu().numpy(), want)                    # one product per position: bit exact


@pytest.mark.parametrize("shape,hidden,t_out", [((5, 19, 33, 2000), 32, 2000), ((3, 8, 33, 2000), 64, 1000),
                                                 ((2, 4, 17, 333), 10, 777), ((1, 19, 33, 2000), 256, 2000)])
def test_against_oracle_shapes(shape, hidden, t_out):
    rng = np.random.default_rng(shape[0] * 100 + hidden)
    spec = rng.random(shape, dtype=np.float32)
    w = (rng.standard_normal((shape[1], hidden // 2, 3, 3)) * 0.2).astype(np.float32)
    b = (rng.standard_normal(hidden // 2) * 0.1).astype(np.float32)
    want = do.spectrogram_condition_oracle(spec, w, b, t_out).numpy()
    got = b200spec.spectrogram_condition(torch.from_numpy(spec).cuda(), torch.from_numpy(w), torch.from_numpy(b), t_out)
    assert got.shape == want.shape and np.abs(got.cpu().numpy() - want).max() <= 5e-5


def test_feeds_from_compute_stft_and_errors():
    class Cfg:
        diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}

    x = torch.randn(4, 19, 2000, device="cuda")
    img = b200spec.compute_stft(x, Cfg)
    w = torch.randn(19, 16, 3, 3) * 0.1
    b = torch.zeros(16)
    cond = b200spec.spectrogram_condition(img, w, b, x.shape[-1])
    want = do.spectrogram_condition_oracle(img.cpu(), w, b, 2000)
    assert (cond.cpu() - want).abs().max().item() <= 5e-5
    with pytest.raises(ValueError):
        b200spec.spectrogram_condition(img, torch.randn(18, 16, 3, 3), b, 2000)
    with pytest.raises(ValueError):
        b200spec.spectrogram_condition(img, w, b, 2000, partner=torch.tensor([0, 1, 2, 9]))


@pytest.mark.parametrize("shape", [(64, 19, 2000), (5, 8, 2000), (3, 19, 10000)])
def test_stft_condition_without_the_image_is_bit_identical(shape):
    """compute_stft's kernel hands over its normalised knots and the condition kernel evaluates the pixels it needs
    from them: the same bits as spectrogram_condition(compute_stft(x)), training-mode mixing and a NaN channel included."""
    class Cfg:
        diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}

    g = torch.Generator(device="cuda").manual_seed(shape[0])
    x = torch.randn(shape, device="cuda", generator=g).cumsum(-1) * 0.3 + torch.randn(shape, device="cuda", generator=g)
    x[0, 1] = 0.0                      # a flat channel
    x[1, 2, 777] = float("nan")        # a poisoned channel: its image rows are NaN
    w = torch.randn((shape[1], 16, 3, 3), device="cuda", generator=g) * 0.1
    b = torch.randn((16,), device="cuda", generator=g) * 0.1
    img = b200spec.compute_stft(x, Cfg)
    partner = torch.randperm(shape[0], device="cuda", generator=g)
    for part in (None, partner):
        want = b200spec.spectrogram_condition(img, w, b, 2000, partner=part)
        got = b200spec.stft_condition(x, Cfg, w, b, 2000, partner=part)
        assert got.shape == want.shape == (shape[0], 16, 2000)
        assert torch.equal(torch.nan_to_num(got, nan=-7.0), torch.nan_to_num(want, nan=-7.0))
    plain = b200spec.stft_condition(x, Cfg, w, b, 2000)
    assert torch.isnan(plain[1]).any() and not torch.isnan(plain[0]).any() and not torch.isnan(plain[2]).any()
    ref = do.spectrogram_condition_oracle(img[2:3].cpu(), w.cpu(), b.cpu(), 2000)
    assert (plain[2:3].cpu() - ref).abs().max() <= 5e-5
