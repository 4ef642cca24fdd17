"""Mode R oracle: CPU restatement of the reference's ``compute_stft``.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Parity: PINNED by
``tests/golden/stft_*.npz`` (outputs of the reference function itself).

Two independent restatements of ``/root/reference/root/src/utils/DiffEEG_utils.py:27-76``:

``compute_stft_port``
    keeps the reference's call structure -- one ``scipy.signal.stft`` call per
    (record, channel) row (ref :46-51), ``log1p(abs(.))`` (ref :54), one ``np.interp`` per
    frequency row onto ``linspace(0, t[-1], 2000)`` (ref :57-61), fp64 stack cast to fp32
    (ref :66-69) and a per-row min-max (ref :72-74).  This is the function whose wall time
    is reported as the CPU baseline (``kind: "port"``): it does the same third-party
    calls, the same number of times, as the reference.

``compute_stft_restated``
    re-derives SciPy's legacy ``_spectral_helper`` semantics (zero boundary extension,
    zero padding to a whole number of hops, periodic Hann rounded to fp32, fp64 rFFT,
    ``scaling='spectrum'``, complex64 result) and ``np.interp`` from first principles in
    vectorised NumPy.  Used to attribute errors and to cover shapes where thousands of
    SciPy calls would be slow.

Integer facts (frame count, padding, interpolation brackets) are exposed separately so
that the GPU path can be checked bit-exactly.
"""
from __future__ import annotations

import numpy as np

FS = 200.0          # literal in the reference (DiffEEG_utils.py:48)
OUT_COLS = 2000     # literal in the reference (DiffEEG_utils.py:57-58)
EPS = 1e-8          # literal in the reference (DiffEEG_utils.py:74)


# --------------------------------------------------------------------------- integers
def frame_plan(length: int, nperseg: int, noverlap: int) -> dict:
    """SciPy ``stft`` framing with ``boundary='zeros', padded=True`` (the defaults the
    reference relies on, DiffEEG_utils.py:46-51).  Pure integer arithmetic."""
    if not 0 <= noverlap < nperseg:
        raise ValueError("noverlap must be less than nperseg.")
    if length < nperseg:
        # SciPy would silently shrink nperseg to the input length (with a warning);
        # neither the reference nor this project ever runs in that regime.
        raise NotImplementedError("signal shorter than nperseg")
    hop = nperseg - noverlap
    half = nperseg // 2
    ext = length + 2 * half
    nadd = (-(ext - nperseg) % hop) % nperseg
    n_frames = (ext + nadd - nperseg) // hop + 1
    return {
        "hop": hop,
        "half": half,
        "nadd": nadd,
        "n_frames": n_frames,
        "n_freqs": nperseg // 2 + 1,
        "padded_len": ext + nadd,
    }


def interp_brackets(n_frames: int, out_cols: int = OUT_COLS) -> np.ndarray:
    """Left knot index used by ``np.interp`` for every output column (ref :57-61).

    Column ``j`` sits at ``j*t[-1]/(out_cols-1)``; knots at ``k*hop/fs``; hence the
    bracket is ``floor(j*(n-1)/(out_cols-1))`` clamped to ``n-2``."""
    if n_frames < 2:
        return np.zeros(out_cols, dtype=np.int64)
    j = np.arange(out_cols, dtype=np.int64)
    i = (j * (n_frames - 1)) // (out_cols - 1)
    return np.minimum(i, n_frames - 2)


# --------------------------------------------------------------------------- the port
def _hop_as_noverlap(cfg) -> tuple[int, int, str]:
    tr = cfg.diffEEG_trainer
    # NB: the reference passes ``stft_hop_length`` as SciPy's ``noverlap`` (ref :50).
    return int(tr["stft_n_fft"]), int(tr["stft_hop_length"]), tr["stft_window"]


def logmag_rows_port(rows: np.ndarray, nperseg: int, noverlap: int, window: str = "hann"):
    """``log1p(|STFT|)`` per row, plus the frame-centre times (ref :46-54, :504-509)."""
    from scipy.signal import stft

    rows = np.asarray(rows, dtype=np.float32)
    out, t = [], None
    for r in rows:
        _, t, z = stft(r, fs=FS, nperseg=nperseg, noverlap=noverlap, window=window)
        out.append(np.log1p(np.abs(z)))
    return np.stack(out, axis=0), t


def compute_stft_port(signal, cfg, out_cols: int = OUT_COLS) -> np.ndarray:
    """(B, C, L) fp32 -> (B, C, nperseg//2+1, out_cols) fp32, the reference's recipe."""
    from scipy.signal import stft

    x = np.asarray(signal, dtype=np.float32)
    bsz, nch, _ = x.shape
    nperseg, noverlap, window = _hop_as_noverlap(cfg)
    images = np.empty((bsz * nch, nperseg // 2 + 1, out_cols), dtype=np.float64)
    for r, row in enumerate(x.reshape(bsz * nch, -1)):
        _, t, z = stft(row, fs=FS, nperseg=nperseg, noverlap=noverlap, window=window)
        mag = np.log1p(np.abs(z))
        grid = np.linspace(0, t[-1], out_cols)
        for f in range(mag.shape[0]):
            images[r, f] = np.interp(grid, t, mag[f])
    img = images.astype(np.float32)
    lo = img.min(axis=-1, keepdims=True)
    hi = img.max(axis=-1, keepdims=True)
    img = (img - lo) / (hi - lo + np.float32(EPS))
    return img.reshape(bsz, nch, nperseg // 2 + 1, out_cols)


# --------------------------------------------------------------------------- first principles
def hann_periodic(n: int) -> np.ndarray:
    """``scipy.signal.get_window('hann', n)`` (fftbins=True => periodic)."""
    return 0.5 - 0.5 * np.cos(2.0 * np.pi * np.arange(n) / n)


def logmag_rows_restated(rows: np.ndarray, nperseg: int, noverlap: int) -> np.ndarray:
    """(R, L) fp32 -> (R, F, n_frames) fp32 ``log1p(|Z|)`` without calling SciPy.

    SciPy's legacy helper promotes fp32 data to fp64 when it appends the padding zeros,
    rounds the window to fp32 (``win.astype(complex64)``), runs the rFFT in fp64, scales
    by ``1/sum(win)`` and only then rounds to complex64 -- restated here step by step."""
    rows = np.asarray(rows, dtype=np.float32)
    plan = frame_plan(rows.shape[-1], nperseg, noverlap)
    half, hop, n = plan["half"], plan["hop"], plan["n_frames"]
    ext = np.zeros((rows.shape[0], plan["padded_len"]), dtype=np.float64)
    ext[:, half:half + rows.shape[-1]] = rows
    win = hann_periodic(nperseg).astype(np.float32).astype(np.float64)
    idx = (np.arange(n) * hop)[:, None] + np.arange(nperseg)[None, :]
    spec = np.fft.rfft(ext[:, idx] * win, axis=-1)           # (R, n, F) complex128
    spec = (spec / win.sum()).astype(np.complex64)
    mag = np.log1p(np.abs(spec))                              # fp32
    return np.ascontiguousarray(np.swapaxes(mag, 1, 2))


def interp_image_restated(mag: np.ndarray, hop: int, out_cols: int = OUT_COLS) -> np.ndarray:
    """Piecewise-linear resample of (..., n) -> (..., out_cols) in fp64 (ref :57-61)."""
    n = mag.shape[-1]
    t = np.arange(n, dtype=np.float64) * hop / FS
    grid = np.linspace(0, t[-1], out_cols)
    if n == 1:
        return np.repeat(mag.astype(np.float64), out_cols, axis=-1)
    i = interp_brackets(n, out_cols)
    lo = mag[..., i].astype(np.float64)
    hi = mag[..., i + 1].astype(np.float64)
    slope = (hi - lo) / (t[i + 1] - t[i])
    out = slope * (grid - t[i]) + lo
    out[..., -1] = mag[..., -1]                               # x == xp[-1] returns fp[-1]
    return out


def compute_stft_restated(signal, nperseg: int = 64, noverlap: int = 32,
                          out_cols: int = OUT_COLS) -> np.ndarray:
    x = np.asarray(signal, dtype=np.float32)
    bsz, nch, length = x.shape
    plan = frame_plan(length, nperseg, noverlap)
    mag = logmag_rows_restated(x.reshape(bsz * nch, length), nperseg, noverlap)
    img = interp_image_restated(mag, plan["hop"], out_cols).astype(np.float32)
    lo = img.min(axis=-1, keepdims=True)
    hi = img.max(axis=-1, keepdims=True)
    img = (img - lo) / (hi - lo + np.float32(EPS))
    return img.reshape(bsz, nch, plan["n_freqs"], out_cols)
