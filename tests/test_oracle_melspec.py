"""CPU: the mode-N oracle -- front half pinned to the reference, back half cross-checked."""
import os

import numpy as np
import pytest

from conftest import eeg_like
from oracle import melspec_oracle as mo
from oracle import ref_extract as rx


def test_montage_matches_reference_config(melspec_golden):
    assert list(melspec_golden["ref_eeg_columns"]) == mo.EEG_COLUMNS
    assert [tuple(p) for p in melspec_golden["ref_map_features"]] == mo.MAP_FEATURES


@pytest.mark.skipif(not rx.reference_available(), reason="/root/reference not mounted")
def test_montage_matches_live_reference_yaml():
    import yaml

    with open(os.path.join(rx.REFERENCE_ROOT, "root", "config", "config.yml")) as fh:
        cfg = yaml.safe_load(fh)
    assert cfg["EEG_COLUMNS"] == mo.EEG_COLUMNS
    assert [tuple(p) for p in cfg["map_features"]] == mo.MAP_FEATURES


def test_crop_and_nan_fill_follow_reference_idiom():
    assert mo.crop_offset(10000) == 0 and mo.crop_offset(18000) == 4000 and mo.crop_offset(10001) == 0
    x = np.array([1.0, np.nan, 3.0, np.nan], np.float32)
    np.testing.assert_array_equal(mo.fill_missing(x), np.array([1, 2, 3, 2], np.float32))
    np.testing.assert_array_equal(mo.fill_missing(np.full(5, np.nan, np.float32)), np.zeros(5, np.float32))


def test_front_half_against_the_references_own_functions(fronthalf_golden):
    """``crop_offset`` / ``fill_missing`` / ``montage_differences`` against golden vectors produced by the
    reference's own ``_get_eeg_window`` (cfg_utils.py:354-389, run on a parquet file) and
    ``calculate_differential_signals`` (data_utils.py:183-215), lifted by ``oracle/ref_extract.py``."""
    g = fronthalf_golden
    for case in ("odd", "longodd"):
        x = g[f"window/{case}/x"]                                  # (n_pts, 20) parquet rows, EEG_COLUMNS order
        off = mo.crop_offset(x.shape[0])
        for variant in ("full19", "magic8"):
            want = g[f"window/{case}/{variant}"]
            cols = [mo.EEG_COLUMNS.index(c) for c in g[f"channels/{variant}"]]
            with np.errstate(all="ignore"):
                import warnings
                with warnings.catch_warnings():
                    warnings.simplefilter("ignore")
                    got = np.stack([mo.fill_missing(x[off:off + mo.EEG_PTS, c]) for c in cols], axis=1)
            assert want.shape == (mo.EEG_PTS, len(cols)) and want.dtype == np.float32
            np.testing.assert_array_equal(got, want)                # bit for bit, NaN-free
    x = g["diff/x"]
    want = g["diff/eeg_columns"]
    np.testing.assert_array_equal(want[:20], x)
    got = mo.montage_differences(x, mo.pair_indices(18))
    np.testing.assert_array_equal(got, want[20:].astype(np.float32))   # integer microvolts: exact in fp32
    want19 = g["diff/yaml_features"]                                # the YAML's own feature_to_index (19 channels)
    assert list(g["diff/yaml_feature_names"]) == mo.EEG_COLUMNS[:19]
    np.testing.assert_array_equal(want19[19:], want[20:])


@pytest.mark.skipif(not rx.reference_available(), reason="/root/reference not mounted")
def test_front_half_against_the_live_reference(tmp_path):
    import pandas as pd

    rng = np.random.default_rng(77)
    x = eeg_like(rng, (11111, 20))
    x[rng.random(11111) < 0.02, 6] = np.nan
    x[:, 10] = np.nan
    path = tmp_path / "eeg.parquet"
    pd.DataFrame(x, columns=mo.EEG_COLUMNS).to_parquet(path)
    fn, cfg = rx.load_reference_get_eeg_window()
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        want = fn(path)
        off = mo.crop_offset(11111)
        got = np.stack([mo.fill_missing(x[off:off + 10000, mo.EEG_COLUMNS.index(c)]) for c in cfg.channel_feats], 1)
    np.testing.assert_array_equal(got, want)


def test_oracle_against_independent_torchaudio_image(melspec_golden, melspec_xcheck):
    """A second implementation of the whole mode-N image from LIBRARY code (torchaudio MelSpectrogram +
    amplitude_to_DB; ``tests/golden/make_golden.py::xcheck_image``).  Narrows, does not close, "unpinned"."""
    for case in ("clean", "nan", "range"):
        got = mo.eeg_to_spectrogram_batch_oracle(melspec_golden[f"{case}/x"])
        err = np.abs(got - melspec_xcheck[f"{case}/y"]).max()
        assert err <= 2e-5, f"{case}: {err:.2e}"


def test_filterbank_structure_and_torchaudio():
    fb = mo.mel_filterbank()
    nz = np.argwhere(fb > 0)
    assert fb.shape == (128, 513) and nz[:, 1].min() == 1 and nz[:, 1].max() == 102
    assert (fb > 0).sum(axis=1).max() <= 2
    ta = pytest.importorskip("torchaudio")
    ref = ta.functional.melscale_fbanks(513, 0.0, 20.0, 128, 200, norm="slaney", mel_scale="slaney").T.numpy()
    assert np.abs(ref - fb).max() <= 1e-4 * fb.max()


def test_stft_power_against_torch_stft():
    import torch

    rng = np.random.default_rng(3)
    y = eeg_like(rng, (10000,))
    win = torch.zeros(1024, dtype=torch.float64)
    win[448:576] = torch.hann_window(128, periodic=True, dtype=torch.float64)
    for mode in ("constant", "reflect"):
        s = torch.stft(torch.from_numpy(y).double(), 1024, 39, 1024, window=win, center=True,
                       pad_mode=mode, return_complex=True).abs() ** 2
        p = mo.stft_power(y, pad_mode=mode)
        assert p.shape == (513, 257)
        assert np.abs(s.numpy() - p).max() <= 1e-6 * p.max()


def test_oracle_reproduces_golden(melspec_golden):
    for case, key, mode in [("clean", "y", "constant"), ("clean", "y_reflect", "reflect"),
                            ("nan", "y", "constant"), ("range", "y", "constant")]:
        got = mo.eeg_to_spectrogram_batch_oracle(melspec_golden[f"{case}/x"], pad_mode=mode)
        np.testing.assert_allclose(got, melspec_golden[f"{case}/{key}"], rtol=0, atol=1e-6)


def test_image_properties(melspec_golden):
    y = melspec_golden["range/y"]
    assert y.shape == (1, 128, 256, 4) and y.dtype == np.float32
    assert y.max() <= 1.0 + 1e-6 and y.min() >= -1.0 - 1e-6   # dB in [-80, 0] -> [-1, 1]
    # the electrode pop on Fp1 drives quiet pixels of pair Fp1-F7 onto the top_db floor
    x = melspec_golden["range/x"][0]
    db = mo.pair_melspec_db(mo.fill_missing(x[0] - x[4]))
    assert (db == -80.0).mean() > 0.05 and db.max() == 0.0


def test_generated_mel_epilogue_is_current_and_matches_the_bank():
    """csrc/mel_default_epilogue.h (literal taps of engine 3) must be what the generator emits
    for the oracle's Slaney bank."""
    import importlib.util
    import os

    from conftest import ROOT

    spec = importlib.util.spec_from_file_location("gen_mel", os.path.join(ROOT, "tools", "gen_mel_epilogue.py"))
    gen = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(gen)
    with open(gen.OUT) as fh:
        assert fh.read() == gen.render()
    # the generator derives the bank itself (product-side derivation); it must be the oracle's bank
    np.testing.assert_array_equal(gen.bank(), mo.mel_filterbank())
