"""Rows (f)3 / (f)4 oracle: the classifier-side preprocessing of the multimodal model's two inputs.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Reference: ``CombinedDataset`` in ``/root/reference/root/jupyter_notebooks/XAI_Multimodality.py`` --
``process_eeg`` :1148-1164 with its helpers :1212-1286 (same code as ``root/src/utils/data_utils.py:183-249``)
and ``process_spectrogram`` :1166-1204 with its helpers :1286-1307 (``data_utils.py:129-181``).  The SHAP
experiments explain exactly these tensors (``shap.GradientExplainer`` on the EEG branch, :2283-2290).

PINNED: ``tests/golden/classifier_golden.npz`` holds inputs and outputs of the reference's own methods, executed
through ``oracle/ref_extract.py::load_reference_combined_dataset`` (AST lift of the notebook export).  One call
of the reference cannot run here: ``skimage.transform.resize`` (scikit-image is not installed); for an image that
already has the target shape (400, 300) it is the identity, the only case the reference's pipeline produces
unless a spectrogram row is entirely NaN -- that case is refused, not guessed.
"""
from __future__ import annotations

import numpy as np
from scipy.ndimage import gaussian_filter
from scipy.signal import butter, filtfilt, iirnotch, lfilter

FS = 200                       # CFG.sampling_rate   (XAI_Multimodality.py:231)
BAND = (0.5, 20.0)             # CFG.bandpass_filter (:205)
FIXED_LENGTH = 3000            # CFG.fixed_length    (:245)
IMAGE_SIZE = (400, 300)        # CFG.image_size      (:246)
# CFG.map_features (:211-217) over CFG.eeg_features = EEG_COLUMNS without EKG (:122), as row indices
PAIRS = [(0, 4), (4, 5), (5, 6), (6, 7), (0, 1), (1, 2), (2, 3), (3, 7),
         (11, 15), (15, 16), (16, 17), (17, 18), (11, 12), (12, 13), (13, 14), (14, 18), (8, 9), (9, 10)]


def bandpass(order: int):
    """``butter(order, [low / nyq, high / nyq], btype='band')`` (:1247-1251): 2 * order + 1 coefficients."""
    nyq = 0.5 * FS
    return butter(order, [BAND[0] / nyq, BAND[1] / nyq], btype="band")


def process_eeg_oracle(values: np.ndarray) -> np.ndarray:
    """``eeg.values`` of one parquet frame, (n, 20) -> the (1, 37, 3000) fp32 tensor of ``process_eeg``.

    Raises if a filtered row is entirely NaN (first sample NaN): the reference then DROPS the row (:1214) and
    every later channel index shifts by one -- behaviour nobody should reproduce."""
    waves = np.asarray(values).T                                   # :1151
    b, a = bandpass(5)                                             # :1156 (default order=5, :1253)
    waves = lfilter(b, a, waves, axis=1)
    if np.isnan(waves).all(axis=1).any():
        raise ValueError("a row is all NaN after the first filter: the reference drops it and misaligns the channels")
    where = np.isnan(waves)                                        # handle_nan :1212-1222
    if where.any():
        mean = np.nanmean(waves, axis=1, keepdims=True)
        waves[where] = np.take(mean, np.where(where)[0])
    diff = np.stack([waves[i] - waves[j] for i, j in PAIRS])       # :1257-1265
    waves = np.vstack((waves, diff))                               # (38, n)
    b, a = bandpass(6)                                             # denoise_filter :1267-1271
    y = lfilter(b, a, waves, axis=1)
    y = (y + np.roll(y, -1) + np.roll(y, -2) + np.roll(y, -3)) / 4   # np.roll WITHOUT axis: rolls the flattened array
    y = y[:, 0:-1:4]
    y = (y - y.mean(axis=1, keepdims=True)) / (y.std(axis=1, keepdims=True) + 1e-6)   # :1273-1276
    # select_and_map_channels :1278-1282: rows 0..18, then 18 rows from index len(feature_to_index) = 19 --
    # i.e. EKG and the first 17 differences (the differences start at row 20; Cz-Pz is lost)
    y = y[list(range(19)) + list(range(19, 19 + len(PAIRS)))]
    if y.shape[1] < FIXED_LENGTH:                                  # pad_or_truncate :1224-1245
        y = np.hstack((y, np.zeros((y.shape[0], FIXED_LENGTH - y.shape[1]))))
    else:
        y = y[:, :FIXED_LENGTH]
    return y[np.newaxis].astype(np.float32)                        # :1162-1163


def spectrogram_window(raw: np.ndarray, offset_seconds=None) -> np.ndarray:
    """:1173-1184: the (400, 300) array the numeric chain starts from.  NB the reference slices COLUMNS of the
    (time, 401) parquet array by the label offset and then transposes; kept as it is."""
    raw = np.asarray(raw)
    if offset_seconds is not None:
        off = int(offset_seconds // 2)
        basic = raw[:, off:off + 300]
        basic = np.pad(basic, ((0, 0), (0, max(0, 300 - basic.shape[1]))), mode="constant")
    else:
        basic = raw
    s = basic.T
    rows, cols = IMAGE_SIZE
    s = np.vstack((s, np.zeros((rows - s.shape[0], s.shape[1])))) if s.shape[0] < rows else s[:rows]
    s = np.hstack((s, np.zeros((s.shape[0], cols - s.shape[1])))) if s.shape[1] < cols else s[:, :cols]
    return s


def process_spectrogram_oracle(window: np.ndarray) -> np.ndarray:
    """The numeric chain of ``process_spectrogram`` on the (400, 300) window -> (3, 400, 300) fp32."""
    s = np.array(window, dtype=np.float64)
    if np.isnan(s).all(axis=1).any():
        raise ValueError("an all-NaN row: the reference drops it and resizes with scikit-image (not available)")
    where = np.isnan(s)                                            # handle_nan
    if where.any():
        mean = np.nanmean(s, axis=1, keepdims=True)
        s[where] = np.take(mean, np.where(where)[0])
    s = s - np.mean(s, axis=0)                                     # baseline_correction :1286-1288
    b, a = iirnotch(60, 30, FS)                                    # apply_notch_filter :1294-1297
    s = filtfilt(b, a, s, axis=0)
    s = gaussian_filter(s, sigma=1.0)                              # smooth_spectrogram :1299-1301
    s = (s - np.min(s)) / (np.max(s) - np.min(s) + 1e-6)           # normalize_signal :1290-1292
    s = np.tile(s[..., None], (1, 1, 3)).astype(np.float32)        # :1190, :1201 (resize to the same shape: identity)
    return np.transpose(s, (2, 0, 1))                              # :1202
