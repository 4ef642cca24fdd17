"""Row (f)1 oracle: the restatement of ``_EEGTransformer.transform`` is PINNED -- against the
committed outputs of the reference's own class (tests/golden/eegtransform_golden.npz, made by
tests/golden/make_golden.py) and, when /root/reference is mounted, against the class itself."""
import ctypes as C

import numpy as np
import pytest

from oracle import eegtransform_oracle as eo
from oracle import ref_extract as rx

CASES = {
    "default19": dict(n_feats=19, downsample=5),
    "magic8": dict(n_feats=8, apply_chris_magic_ch8=True, downsample=5),
    "nan_clip": dict(n_feats=19, downsample=5),
    "no_downsample": dict(n_feats=19),
    "no_filter_ds4": dict(n_feats=19, apply_butter_lowpass_filter=False, downsample=4),
    "raw_scale": dict(n_feats=19, normalize=False, downsample=3),
    "mu_law": dict(n_feats=19, apply_mu_law_encoding=True, downsample=5),
    "mu_law_nofilter_nan": dict(n_feats=19, apply_butter_lowpass_filter=False, apply_mu_law_encoding=True, downsample=2),
    # (L, 19): the window _get_eeg_window hands the transformer (cfg_utils.py:366-373) -- no EKG column
    "window19": dict(n_feats=19, downsample=5),
    "window19_ds4": dict(n_feats=19, downsample=4),
    "window19_odd_length": dict(n_feats=19, downsample=5),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_golden(eegtransform_golden, name):
    x, want = eegtransform_golden[f"{name}/x"], eegtransform_golden[f"{name}/y"]
    got = eo.eeg_transform_oracle(x, **CASES[name])
    f64 = CASES[name].get("apply_butter_lowpass_filter", True) or CASES[name].get("apply_mu_law_encoding", False)
    assert got.shape == want.shape and got.dtype == (np.float64 if f64 else np.float32)
    np.testing.assert_allclose(got, want, rtol=0, atol=1e-12, equal_nan=True)   # same SciPy calls: bit-level agreement


@pytest.mark.parametrize("name", ["default19", "magic8", "nan_clip"])
def test_written_out_recurrence_is_lfilter(eegtransform_golden, name):
    """The explicit direct-form-II-transposed loop (what the CUDA kernel parallelises) is SciPy's lfilter."""
    x, want = eegtransform_golden[f"{name}/x"][:1500], eegtransform_golden[f"{name}/y"]
    kw = dict(CASES[name])
    got = eo.eeg_transform_oracle(x, use_scipy=False, **kw)
    np.testing.assert_allclose(got, want[:got.shape[0]], rtol=0, atol=1e-10)


@pytest.mark.skipif(not rx.reference_available(), reason="/root/reference is not mounted")
def test_oracle_matches_live_reference():
    ref = rx.load_reference_eeg_transformer()
    rng = np.random.default_rng(5)
    x = (rng.standard_normal((10000, 20)) * 60).astype(np.float32)
    x[17:30, 2] = np.nan
    for kw in CASES.values():
        xx = x[:, :8] if kw.get("apply_chris_magic_ch8") else x
        np.testing.assert_array_equal(eo.eeg_transform_oracle(xx, **kw), ref(**kw).transform(xx))
    cfg = rx.reference_cfg_channels()
    assert len(cfg.channel_feats) == eo.N_CHANNEL_FEATS and cfg.EEG_PTS == 10000
    codes = {f: i for i, f in enumerate(cfg.feats)}                       # dataset.py:110-121
    names = [("Fp1", "T3"), ("T3", "O1"), ("Fp1", "C3"), ("C3", "O1"), ("Fp2", "C4"), ("C4", "O2"), ("Fp2", "T4"), ("T4", "O2")]
    assert [(codes[a], codes[b]) for a, b in names] == eo.MAGIC8_PAIRS


def test_abi_default_coefficients_are_scipy_butter():
    """eegspec_xform_params_default derives butter(4, 0.2) itself (host code, no GPU needed)."""
    from b200spec import _native as nat

    p = nat.XformParams()
    nat.lib.eegspec_xform_params_default(C.byref(p))
    b, a = eo.butter_lowpass()
    np.testing.assert_allclose(np.array(p.b[:]), b, rtol=1e-12, atol=1e-15)
    np.testing.assert_allclose(np.array(p.a[:]), a, rtol=1e-12, atol=1e-15)
    assert (p.n_out, p.downsample, p.normalize, p.filter) == (19, 5, 1, 1)
    assert p.clip_abs == 1024.0 and p.scale == 1.0 / 32.0
    assert nat.lib.eegspec_xform_out_length(10000, C.byref(p)) == 2000
    assert nat.lib.eegspec_xform_out_length(10001, C.byref(p)) == 2001
