"""CPU restatement of the reference's raw-EEG transformer (row (f)1 of the scope table).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``): imported by ``tests/``, by
``__graft_entry__.smoke()`` and by ``bench.py``'s CPU-baseline leg, never by the product path.

Follows ``_EEGTransformer.transform`` (/root/reference/root/src/data/dataset.py:73-104):

1. channel selection ``x_[:, selected_indices]`` (:88-90) -- the indices are ``range(n)`` over
   ``CFG.channel_feats`` (:69-71) -- or the "magic 8" differences (:106-123) over ``CFG.feats``;
2. ``np.clip(x_, -1024, 1024)`` then ``np.nan_to_num(x_, nan=0) / 32.0`` (:92-94);
3. ``butter(4, 20 / 100, btype="low")`` + ``lfilter(b, a, data, axis=0)`` (:125-131): a causal IIR
   along time, SciPy's direct form II transposed, float64;
4. optional mu-law companding ``sign(x) log(1 + mu |x|) / log(mu + 1)`` with mu = 1 (:98-99, :133-147);
5. ``x_[::downsample, :]`` (:101-102).

PINNED: ``tests/test_oracle_eegtransform.py`` checks this restatement against the reference's own
class executed through ``oracle/ref_extract.py`` (when /root/reference is present) and against
the committed golden vectors generated from it (``tests/golden/eegtransform_golden.npz``).
"""
from __future__ import annotations

import numpy as np

# CFG.feats order (cfg_utils.py:87-90): Fp1 T3 C3 O1 Fp2 C4 T4 O2 -> dataset.py:110-121
N_CHANNEL_FEATS = 19   # len(CFG.channel_feats), cfg_utils.py:92-95: the selection ignores n_feats
MAGIC8_PAIRS = [(0, 1), (1, 3), (0, 2), (2, 3), (4, 5), (5, 7), (4, 6), (6, 7)]


def butter_lowpass(cutoff_freq=20.0, sampling_rate=200.0, order=4):
    """dataset.py:125-128.  Returns (b, a) as float64 arrays of length order + 1."""
    from scipy.signal import butter

    return butter(order, cutoff_freq / (0.5 * sampling_rate), btype="low", analog=False)


def lfilter_df2t(b, a, x):
    """SciPy's ``lfilter`` recurrence (direct form II transposed, zero initial state) along axis 0,
    written out: y[n] = b0 x[n] + z0;  z_i = b_{i+1} x[n] - a_{i+1} y[n] + z_{i+1}."""
    b = np.asarray(b, np.float64) / a[0]
    a = np.asarray(a, np.float64) / a[0]
    order = len(a) - 1
    x = np.asarray(x, np.float64)
    z = np.zeros((order + 1,) + x.shape[1:], np.float64)
    y = np.empty_like(x)
    for n in range(x.shape[0]):
        y[n] = b[0] * x[n] + z[0]
        for i in range(order):
            z[i] = b[i + 1] * x[n] - a[i + 1] * y[n] + z[i + 1]
    return y


def eeg_transform_oracle(x, n_feats=19, apply_chris_magic_ch8=False, normalize=True,
                         apply_butter_lowpass_filter=True, downsample=None, use_scipy=True,
                         apply_mu_law_encoding=False):
    """x: (L, C) raw EEG of one record -> (ceil(L / downsample), n_out) float64."""
    x_ = np.array(x, copy=True)
    if apply_chris_magic_ch8:
        tmp = np.zeros((x_.shape[0], n_feats), dtype="float32")
        for i, (p, q) in enumerate(MAGIC8_PAIRS):
            tmp[:, i] = x_[:, p] - x_[:, q]
        x_ = tmp
    else:
        x_ = x_[:, :N_CHANNEL_FEATS]
    if normalize:
        x_ = np.clip(x_, -1024, 1024)
        x_ = np.nan_to_num(x_, nan=0) / 32.0
    if apply_butter_lowpass_filter:
        b, a = butter_lowpass()
        if use_scipy:
            from scipy.signal import lfilter

            x_ = lfilter(b, a, x_, axis=0)
        else:
            x_ = lfilter_df2t(b, a, x_)
    if apply_mu_law_encoding:
        mu = 1                                         # dataset.py:99: self._quantize_data(x_, 1)
        x_ = np.sign(x_) * np.log(1 + mu * np.abs(x_)) / np.log(mu + 1)
    if downsample is not None:
        x_ = x_[::downsample, :]
    return x_


def eeg_transform_batch_oracle(x, **kw):
    return np.stack([eeg_transform_oracle(r, **kw) for r in x])
