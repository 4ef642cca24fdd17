"""GPU mirror of the reference's raw-EEG transformer (row (f)1 of the scope table).

``EEGTransformer`` has the constructor and the ``transform`` method of ``_EEGTransformer``
(/root/reference/root/src/data/dataset.py:48-104): same argument names and meaning, same
``(L, C) -> (L / downsample, n_out)`` layout per record.  Differences, all explicit:

* it also accepts a batch ``(B, L, C)`` (the reference loops over records in Python);
* the result is float32 on the input's device (the reference returns float64 from
  ``scipy.signal.lfilter``); NumPy in -> NumPy out, tensor in -> tensor out;
* ``apply_mu_law_encoding=True`` compands with mu = 1 like the reference (``_quantize_data(x_, 1)``);
* ``channel_major=True`` returns ``(B, n_out, L_out)``, the layout ``compute_stft`` consumes.

The arithmetic runs in ``csrc/eegtransform.cu`` behind ``eegspec_eeg_transform_f32``.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np
import torch

from . import _native as nat
from .transform import _ptr, _require_cuda, _stream_ptr

# dataset.py:110-121 over CFG.feats = Fp1 T3 C3 O1 Fp2 C4 T4 O2 (cfg_utils.py:87-90)
MAGIC8_PAIRS = ((0, 1), (1, 3), (0, 2), (2, 3), (4, 5), (5, 7), (4, 6), (6, 7))
N_CHANNEL_FEATS = 19   # len(CFG.channel_feats), cfg_utils.py:92-95
EEG_PTS = 10000        # CFG.EEG_PTS, cfg_utils.py:75


def xform_params(n_feats: int = 19, apply_chris_magic_ch8: bool = False, normalize: bool = True,
                 apply_butter_lowpass_filter: bool = True, downsample: Optional[int] = None,
                 channel_major: bool = False, apply_mu_law_encoding: bool = False) -> nat.XformParams:
    p = nat.XformParams()
    nat.lib.eegspec_xform_params_default(C.byref(p))     # Butterworth (4, 20 Hz @ 200 Hz) in double precision
    if apply_chris_magic_ch8:
        if n_feats != len(MAGIC8_PAIRS):
            raise ValueError("the magic-8 montage has 8 features (dataset.py:108-121)")
        p.n_out = len(MAGIC8_PAIRS)
        for i, (a, b) in enumerate(MAGIC8_PAIRS):
            p.sel_a[i], p.sel_b[i] = a, b
    else:
        p.n_out = N_CHANNEL_FEATS                         # the selection ignores n_feats (dataset.py:88-90)
        for i in range(N_CHANNEL_FEATS):
            p.sel_a[i], p.sel_b[i] = i, -1
    p.normalize = int(bool(normalize))
    p.filter = int(bool(apply_butter_lowpass_filter))
    p.downsample = 1 if downsample is None else int(downsample)
    if p.downsample < 1:
        raise ValueError("downsample must be a positive integer or None")
    p.out_channel_major = int(bool(channel_major))
    p.mu_law = int(bool(apply_mu_law_encoding))       # dataset.py:98-99: _quantize_data(x_, 1) -> mu = 1
    p.mu = 1.0
    return p


class EEGTransformer:
    """Drop-in for ``_EEGTransformer`` (dataset.py:48-104) running on the GPU."""

    def __init__(self, n_feats: int, apply_chris_magic_ch8: bool = False, normalize: bool = True,
                 apply_butter_lowpass_filter: bool = True, apply_mu_law_encoding: bool = False,
                 downsample: Optional[int] = None, channel_major: bool = False) -> None:
        self.n_feats = n_feats
        self.apply_chris_magic_ch8 = apply_chris_magic_ch8
        self.normalize = normalize
        self.apply_butter_lowpass_filter = apply_butter_lowpass_filter
        self.apply_mu_law_encoding = apply_mu_law_encoding
        self.downsample = downsample
        self.params = xform_params(n_feats, apply_chris_magic_ch8, normalize, apply_butter_lowpass_filter,
                                   downsample, channel_major, apply_mu_law_encoding)

    def transform(self, x):
        """x: (L, C) or (B, L, C) raw EEG, time-major like the reference's arrays."""
        _require_cuda()
        as_numpy = isinstance(x, np.ndarray)
        t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32)) if as_numpy else x
        if t.dim() not in (2, 3):
            raise ValueError(f"expected (L, C) or (B, L, C), got {tuple(t.shape)}")
        single = t.dim() == 2
        if single:
            t = t.unsqueeze(0)
        if self.apply_chris_magic_ch8 and t.shape[1] != EEG_PTS:
            raise ValueError(f"the magic-8 montage is defined on {EEG_PTS}-sample windows (dataset.py:108)")
        src_device = t.device
        dev = t.device if t.is_cuda else torch.device("cuda", torch.cuda.current_device())
        t = t.to(device=dev, dtype=torch.float32)
        batch, length, n_ch = t.shape
        need = max(max(self.params.sel_a[i], self.params.sel_b[i]) for i in range(self.params.n_out)) + 1
        if n_ch < need:
            raise ValueError(f"{n_ch} channels given, the selection needs {need}")
        if n_ch % 4 and length % 4:     # 16-byte row copies need whole words per record: the reference's
            t = torch.nn.functional.pad(t, (0, 4 - n_ch % 4))   # (10000, 19) windows go in as they are,
            n_ch = t.shape[2]                                   # only odd lengths of odd widths are padded
        t = t.contiguous()
        l_out = int(nat.lib.eegspec_xform_out_length(length, C.byref(self.params)))
        shape = (batch, self.params.n_out, l_out) if self.params.out_channel_major else (batch, l_out, self.params.n_out)
        out = torch.empty(shape, device=dev, dtype=torch.float32)
        with torch.cuda.device(dev):
            nat.check(nat.lib.eegspec_eeg_transform_f32(_ptr(t), batch, length, n_ch, length * n_ch,
                                                        C.byref(self.params), _ptr(out), _stream_ptr(dev)),
                      "eegspec_eeg_transform_f32")
        if single:
            out = out[0]
        if as_numpy:
            return out.cpu().numpy()
        return out.to(src_device)


def eeg_transform_stft(x: torch.Tensor, config, transformer: Optional[EEGTransformer] = None) -> torch.Tensor:
    """Raw EEG windows -> the DiffEEG spectrogram condition in one call.

    ``x``: (B, L, C) time-major raw EEG on the device (the (10000, 19) windows of ``_get_eeg_window``).
    Equivalent to ``compute_stft(transformer.transform(x, channel_major), config)`` -- the reference's
    ``EEGDataset`` transform (dataset.py:73-104, :213-228) followed by ``compute_stft``
    (DiffEEG_utils.py:27-76, called at DiffEEG_train.py:131): one C-ABI call, two launches per pass of
    ~2 000 records, the transformed (C, L/5) rows (3 % of the image bytes) in a workspace between them.
    Returns (B, n_out, F, 2000)."""
    from .transform import EPS, OUT_COLS, _stft_keys

    _require_cuda()
    tr = transformer if transformer is not None else EEGTransformer(n_feats=19, downsample=5)
    nperseg, noverlap = _stft_keys(config)
    if x.dim() != 3 or not x.is_cuda:
        raise ValueError("eeg_transform_stft expects a (B, L, C) tensor on the device")
    t = x.to(dtype=torch.float32).contiguous()
    batch, length, n_ch = t.shape
    if n_ch % 4 and length % 4:
        t = torch.nn.functional.pad(t, (0, 4 - n_ch % 4)).contiguous()
        n_ch = t.shape[2]
    p = tr.params
    n_freqs = nperseg // 2 + 1
    out = torch.empty((batch, p.n_out, n_freqs, OUT_COLS), device=t.device, dtype=torch.float32)
    if batch == 0:
        return out
    with torch.cuda.device(t.device):
        ws_bytes = nat.lib.eegspec_eeg_transform_stft_workspace(batch, length, C.byref(p), nperseg, noverlap, OUT_COLS)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=t.device)
        nat.check(nat.lib.eegspec_eeg_transform_stft_f32(_ptr(t), batch, length, n_ch, length * n_ch, C.byref(p),
                                                         nperseg, noverlap, OUT_COLS, EPS, _ptr(out), _ptr(ws),
                                                         ws_bytes, _stream_ptr(t.device)),
                  "eeg_transform_stft")
    return out
