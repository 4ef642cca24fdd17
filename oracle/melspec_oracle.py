"""Mode N oracle: raw EEG -> 4-chain x 128-mel log spectrogram image.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

PARITY UNPINNED for the mel back half: ``/root/reference`` contains no mel / librosa
code at all; the recipe is the one ``/root/repo/BASELINE.json`` (``north_star``)
describes.  The third-party arithmetic is **librosa** (not pinned by the reference's
``requirements.txt``, not installed here); this file restates its published algorithm,
0.10.x semantics: ``librosa.feature.melspectrogram`` -> ``librosa.stft`` (centered,
``pad_mode='constant'``; ``'reflect'`` for librosa < 0.10 is available as an option),
``librosa.filters.mel`` (Slaney scale, Slaney norm, fp32), ``librosa.power_to_db``
(``ref=np.max``, ``amin=1e-10``, ``top_db=80``).

The front half IS pinned to the reference:
  * channel order  = ``EEG_COLUMNS``        /root/reference/root/config/config.yml:34-54
  * bipolar pairs  = ``map_features`` 0..15 /root/reference/root/config/config.yml:529-545
    (same differencing as ``calculate_differential_signals``,
    /root/reference/root/src/utils/data_utils.py:183-215)
  * middle crop    = ``(n - 10000)//2``     /root/reference/root/src/utils/cfg_utils.py:369-371
  * NaN fill       = nanmean / all-NaN->0   /root/reference/root/src/utils/cfg_utils.py:380-385
"""
from __future__ import annotations

import numpy as np

# /root/reference/root/config/config.yml:34-54
EEG_COLUMNS = ["Fp1", "F3", "C3", "P3", "F7", "T3", "T5", "O1", "Fz", "Cz", "Pz",
               "Fp2", "F4", "C4", "P4", "F8", "T4", "T6", "O2", "EKG"]
# /root/reference/root/config/config.yml:529-547 (18 pairs; the first 16 are the four
# chains LL, LP, RP, RR of the north_star, four consecutive differences each)
MAP_FEATURES = [("Fp1", "F7"), ("F7", "T3"), ("T3", "T5"), ("T5", "O1"),
                ("Fp1", "F3"), ("F3", "C3"), ("C3", "P3"), ("P3", "O1"),
                ("Fp2", "F8"), ("F8", "T4"), ("T4", "T6"), ("T6", "O2"),
                ("Fp2", "F4"), ("F4", "C4"), ("C4", "P4"), ("P4", "O2"),
                ("Fz", "Cz"), ("Cz", "Pz")]
CHAIN_NAMES = ["LL", "LP", "RP", "RR"]

SR = 200
EEG_PTS = 10_000     # /root/reference/root/src/utils/cfg_utils.py:73-75 (EEG_PTS)
N_FFT = 1024
WIN_LENGTH = 128
N_MELS = 128
FMIN, FMAX = 0.0, 20.0
N_TIME = 256
AMIN = 1e-10
TOP_DB = 80.0


def pair_indices(n_pairs: int = 16) -> np.ndarray:
    col = {name: i for i, name in enumerate(EEG_COLUMNS)}
    return np.array([[col[a], col[b]] for a, b in MAP_FEATURES[:n_pairs]], dtype=np.int32)


def crop_offset(n_pts: int, window: int = EEG_PTS) -> int:
    """cfg_utils.py:369-371 -- ``offset = (n_pts - EEG_PTS)//2``."""
    return (n_pts - window) // 2


def montage_differences(eeg: np.ndarray, pairs: np.ndarray) -> np.ndarray:
    """``diff[i] = x[a_i] - x[b_i]`` (data_utils.py:183-215; fp32 as dataset.py:106-123)."""
    eeg = np.asarray(eeg, dtype=np.float32)
    return eeg[pairs[:, 0]] - eeg[pairs[:, 1]]


def fill_missing(x: np.ndarray) -> np.ndarray:
    """cfg_utils.py:380-385 applied to one fp32 series."""
    x = np.array(x, dtype=np.float32, copy=True)
    nan = np.isnan(x)
    if nan.mean() < 1:
        if nan.any():
            mean = np.nanmean(x)
            x = np.nan_to_num(x, nan=mean)
    else:
        x[:] = 0
    return x


# --------------------------------------------------------------------------- librosa restated
def _hz_to_mel_slaney(f):
    f = np.asanyarray(f, dtype=np.float64)
    f_sp = 200.0 / 3
    mels = f / f_sp
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    with np.errstate(divide="ignore", invalid="ignore"):
        log_t = f >= min_log_hz
        mels = np.where(log_t, min_log_mel + np.log(np.maximum(f, 1e-300) / min_log_hz) / logstep, mels)
    return mels


def _mel_to_hz_slaney(m):
    m = np.asanyarray(m, dtype=np.float64)
    f_sp = 200.0 / 3
    freqs = f_sp * m
    min_log_hz = 1000.0
    min_log_mel = min_log_hz / f_sp
    logstep = np.log(6.4) / 27.0
    log_t = m >= min_log_mel
    return np.where(log_t, min_log_hz * np.exp(logstep * (m - min_log_mel)), freqs)


def mel_filterbank(sr: int = SR, n_fft: int = N_FFT, n_mels: int = N_MELS,
                   fmin: float = FMIN, fmax: float = FMAX) -> np.ndarray:
    """``librosa.filters.mel(htk=False, norm='slaney', dtype=float32)`` -> (n_mels, n_fft//2+1)."""
    fftfreqs = np.fft.rfftfreq(n=n_fft, d=1.0 / sr)
    mel_pts = np.linspace(_hz_to_mel_slaney(fmin), _hz_to_mel_slaney(fmax), n_mels + 2)
    mel_f = _mel_to_hz_slaney(mel_pts)
    fdiff = np.diff(mel_f)
    ramps = np.subtract.outer(mel_f, fftfreqs)
    weights = np.zeros((n_mels, 1 + n_fft // 2), dtype=np.float32)
    for i in range(n_mels):
        lower = -ramps[i] / fdiff[i]
        upper = ramps[i + 2] / fdiff[i + 1]
        weights[i] = np.maximum(0, np.minimum(lower, upper))
    enorm = 2.0 / (mel_f[2:n_mels + 2] - mel_f[:n_mels])
    weights *= enorm[:, np.newaxis]
    return weights


def stft_power(y: np.ndarray, n_fft: int = N_FFT, hop: int | None = None,
               win_length: int = WIN_LENGTH, pad_mode: str = "constant") -> np.ndarray:
    """``np.abs(librosa.stft(y, center=True))**2`` -> (n_fft//2+1, n_frames) fp32.

    librosa multiplies fp32 frames by an fp64 periodic-Hann window zero-padded (centred)
    to ``n_fft``, runs the rFFT in fp64 and stores complex64."""
    y = np.asarray(y, dtype=np.float32)
    hop = hop if hop is not None else len(y) // N_TIME
    win = 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(win_length) / win_length)
    lpad = (n_fft - win_length) // 2
    fft_window = np.zeros(n_fft)
    fft_window[lpad:lpad + win_length] = win
    ypad = np.pad(y, n_fft // 2, mode=pad_mode)
    n_frames = 1 + (len(ypad) - n_fft) // hop
    idx = (np.arange(n_frames) * hop)[:, None] + np.arange(n_fft)[None, :]
    spec = np.fft.rfft(ypad[idx].astype(np.float64) * fft_window, axis=-1).astype(np.complex64)
    return (np.abs(spec) ** 2).T.astype(np.float32)


def power_to_db_refmax(s: np.ndarray, amin: float = AMIN, top_db: float = TOP_DB) -> np.ndarray:
    """``librosa.power_to_db(S, ref=np.max)`` in fp32."""
    s = np.asarray(s, dtype=np.float32)
    ref = np.max(s)
    log_spec = 10.0 * np.log10(np.maximum(np.float32(amin), s))
    log_spec -= 10.0 * np.log10(np.maximum(np.float32(amin), ref))
    return np.maximum(log_spec, log_spec.max() - np.float32(top_db)).astype(np.float32)


def pair_melspec_db(y: np.ndarray, fb: np.ndarray | None = None, pad_mode: str = "constant") -> np.ndarray:
    """One montage difference (already NaN-filled) -> (128, 257) dB image."""
    fb = mel_filterbank() if fb is None else fb
    power = stft_power(y, pad_mode=pad_mode)
    mel = (fb @ power).astype(np.float32)
    return power_to_db_refmax(mel)


def eeg_to_spectrogram_oracle(eeg: np.ndarray, pad_mode: str = "constant",
                              fb: np.ndarray | None = None) -> np.ndarray:
    """One record ``(20, L>=10000)`` channel-major fp32 -> ``(128, 256, 4)`` fp32.

    Chain ``k`` averages four consecutive pairs ``4k..4k+3``; every pair image is
    ``(dB + 40)/40`` with the dB referenced to that pair's own maximum."""
    eeg = np.asarray(eeg, dtype=np.float32)
    fb = mel_filterbank() if fb is None else fb
    off = crop_offset(eeg.shape[-1])
    win = eeg[:, off:off + EEG_PTS]
    diffs = montage_differences(win, pair_indices(16))
    img = np.zeros((N_MELS, N_TIME, 4), dtype=np.float32)
    for k in range(4):
        for kk in range(4):
            y = fill_missing(diffs[4 * k + kk])
            db = pair_melspec_db(y, fb, pad_mode)
            width = (db.shape[1] // 32) * 32
            img[:, :, k] += (db[:, :width] + 40) / 40
        img[:, :, k] /= 4.0
    return img


def eeg_to_spectrogram_batch_oracle(eeg: np.ndarray, pad_mode: str = "constant") -> np.ndarray:
    fb = mel_filterbank()
    return np.stack([eeg_to_spectrogram_oracle(r, pad_mode, fb) for r in np.asarray(eeg)], axis=0)
