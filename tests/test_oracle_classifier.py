"""CPU: the rows (f)3 / (f)4 oracle against golden vectors of the reference's own ``CombinedDataset`` methods
(tests/golden/make_golden.py::make_classifier, AST-lifted from the notebook export)."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

from oracle import classifier_oracle as co
from oracle import ref_extract as rx

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
from make_golden import CLASSIFIER_EEG_CASES, CLASSIFIER_SPEC_CASES, classifier_eeg_input  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "classifier_golden.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.mark.parametrize("name", sorted(CLASSIFIER_EEG_CASES))
def test_process_eeg_oracle_is_the_reference(gold, name):
    got = co.process_eeg_oracle(classifier_eeg_input(gold, name))
    want = gold[f"eeg/{name}/y"]
    assert got.shape == want.shape == (1, 37, 3000) and got.dtype == np.float32
    np.testing.assert_array_equal(got, want)                 # same SciPy / NumPy calls: bit for bit


@pytest.mark.parametrize("name", sorted(CLASSIFIER_SPEC_CASES))
def test_process_spectrogram_oracle_is_the_reference(gold, name):
    raw = gold[f"spec/{name}/raw_f16"].astype(np.float32)
    off = int(gold[f"spec/{name}/offset"])
    win = co.spectrogram_window(raw, None if off < 0 else off)
    assert win.shape == (400, 300)
    got = co.process_spectrogram_oracle(win)
    assert got.shape == (3, 400, 300) and got.dtype == np.float32
    for plane in range(3):
        np.testing.assert_array_equal(got[plane], gold[f"spec/{name}/y_plane"])


def test_default_coefficients_are_scipys():
    from scipy.signal import butter, iirnotch, lfilter_zi

    from b200spec import _native as nat

    p = nat.ClsEegParams()
    nat.lib.eegspec_cls_eeg_params_default(C.byref(p))
    for order, n, b_c, a_c in ((5, p.n_taps1, p.b1, p.a1), (6, p.n_taps2, p.b2, p.a2)):
        b, a = co.bandpass(order)
        assert n == 2 * order + 1
        np.testing.assert_allclose(list(b_c)[:n], b, rtol=1e-15, atol=0)
        np.testing.assert_allclose(list(a_c)[:n], a, rtol=1e-15, atol=0)
        assert np.all(np.asarray(b)[1::2] == 0)                  # the kernel drops the odd numerator taps
    assert [(p.pair_a[i], p.pair_b[i]) for i in range(p.n_pairs)] == co.PAIRS
    assert (p.n_features, p.fixed_length) == (19, 3000)
    s = nat.ClsSpecParams()
    nat.lib.eegspec_cls_spec_params_default(C.byref(s))
    b, a = iirnotch(60, 30, 200)
    np.testing.assert_allclose(list(s.notch_b), b, rtol=1e-15)
    np.testing.assert_allclose(list(s.notch_a), a, rtol=1e-15)
    np.testing.assert_allclose(list(s.notch_zi), lfilter_zi(b, a), rtol=1e-13)
    assert (s.rows, s.cols) == co.IMAGE_SIZE


@pytest.mark.skipif(not rx.reference_available(), reason="/root/reference not mounted")
def test_against_the_live_reference():
    import pandas as pd

    from oracle import melspec_oracle as mo

    rng = np.random.default_rng(8)
    x = (np.cumsum(rng.standard_normal((10200, 20)), 0) + 4 * rng.standard_normal((10200, 20))).astype(np.float32)
    x[7000:7004, 5] = np.nan
    ds = rx.load_reference_combined_dataset({3: pd.DataFrame(x, columns=mo.EEG_COLUMNS)}, {})
    np.testing.assert_array_equal(co.process_eeg_oracle(x), ds.process_eeg({"eeg_id": 3}).numpy())
