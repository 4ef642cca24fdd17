"""CPU: the row-(f)2 oracle against golden vectors of the reference's own DiffEEG module."""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import diffeeg_oracle as do
from oracle import ref_extract as rx

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden import diffeeg_inputs  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "diffeeg_golden.npz")


@pytest.mark.parametrize("case", ["ship19", "ckpt8", "train19"])
def test_oracle_reproduces_reference_module(case):
    gold = np.load(GOLD)
    spec, c, hidden, _ = diffeeg_inputs(case)
    partner = gold[f"{case}/partner"] if case == "train19" else None
    got = do.spectrogram_condition_oracle(spec, gold[f"{case}/weight"], gold[f"{case}/bias"], 2000, partner).numpy()
    assert np.abs(got - gold[f"{case}/fp32"]).max() <= 1e-6            # the reference's own layer object, fp32
    assert np.abs(got - gold[f"{case}/hooked"]).max() <= 2e-2          # its forward under bf16 autocast


def test_source_positions_match_aten():
    for in_len, out_len in [(33 * 15991, 2000), (33 * 15991, 1000), (17 * 2655, 777), (5, 9)]:
        ramp = torch.arange(in_len, dtype=torch.float32).view(1, 1, -1)
        want = torch.nn.functional.interpolate(ramp, size=out_len, mode="linear", align_corners=False)[0, 0].numpy()
        i0, i1, lam = do.source_positions(in_len, out_len)
        got = (np.float32(1) - lam) * i0.astype(np.float32) + lam * i1.astype(np.float32)
        np.testing.assert_allclose(got, want, rtol=0, atol=np.float32(in_len) * 2e-7)


@pytest.mark.skipif(not rx.reference_available(), reason="/root/reference not mounted")
def test_live_reference_layer_hyperparameters():
    m = rx.load_reference_diffeeg()
    up = m.spectrogram_upsample1
    assert up.stride == do.STRIDE and up.padding == do.PADDING and up.kernel_size == (3, 3)
    assert up.in_channels == 19 and up.out_channels == 16
