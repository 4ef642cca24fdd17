"""Rows (f)3 / (f)4: the classifier-side preprocessing of the multimodal model's two inputs on the GPU.

Mirrors of ``CombinedDataset.process_eeg`` and ``CombinedDataset.process_spectrogram``
(/root/reference/root/jupyter_notebooks/XAI_Multimodality.py:1148-1204; helpers :1212-1307, the same code as
/root/reference/root/src/utils/data_utils.py:129-249), batched.  The arithmetic runs in
``csrc/classifier_eeg.cu`` / ``csrc/classifier_spec.cu`` behind ``eegspec_classifier_eeg_f32`` /
``eegspec_classifier_spec_f32``.

Two behaviours of the reference are refused rather than reproduced (``ValueError``): a channel whose first
sample is NaN (its filtered row is all NaN, ``handle_nan`` drops the row and every later channel index shifts),
and a spectrogram row that is entirely NaN (dropped, then the image is resized with scikit-image).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _native as nat
from .transform import _ptr, _require_cuda, _stream_ptr

IMAGE_SIZE = (400, 300)     # CFG.image_size (XAI_Multimodality.py:246)


def _device_f32(x: torch.Tensor) -> torch.Tensor:
    dev = x.device if x.is_cuda else torch.device("cuda", torch.cuda.current_device())
    return x.to(device=dev, dtype=torch.float32).contiguous()


def process_eeg(values: torch.Tensor, check: bool = True) -> torch.Tensor:
    """``values``: ``eeg.values`` of one parquet frame, (n, 20) in ``EEG_COLUMNS`` order, or a batch (B, n, 20).
    Returns the (B, 1, 37, 3000) fp32 tensor ``process_eeg`` builds (19 channels, EKG, 17 montage differences;
    band-passed, decimated by 4, z-scored, padded to 3000) on the device (one record -> (1, 37, 3000)).
    ``check=True`` synchronises once to raise on records the reference would mangle."""
    _require_cuda()
    single = values.dim() == 2
    x = _device_f32(values.unsqueeze(0) if single else values)
    if x.dim() != 3:
        raise ValueError("values must be (n, 20) or (B, n, 20)")
    batch, length, n_ch = x.shape
    p = nat.ClsEegParams()
    nat.lib.eegspec_cls_eeg_params_default(C.byref(p))
    out = torch.empty((batch, 1, p.n_features + p.n_pairs, p.fixed_length), dtype=torch.float32, device=x.device)
    if batch == 0:
        return out
    with torch.cuda.device(x.device):
        status = torch.empty(batch, dtype=torch.int32, device=x.device)
        ws_bytes = nat.lib.eegspec_classifier_eeg_workspace(batch, length, n_ch)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
        nat.check(nat.lib.eegspec_classifier_eeg_f32(_ptr(x), batch, length, n_ch, length * n_ch, C.byref(p), _ptr(out),
                                                     _ptr(status), _ptr(ws), ws_bytes, _stream_ptr(x.device)),
                  "process_eeg")
        if check and bool(status.any()):
            bad = status.nonzero().flatten().tolist()
            raise ValueError(f"records {bad}: a channel starts with NaN -- the reference drops that row and misaligns "
                             "every later channel (XAI_Multimodality.py:1214); not reproduced")
    return out[0] if single else out


def spectrogram_window(raw: torch.Tensor, offset_seconds: Optional[float] = None) -> torch.Tensor:
    """XAI_Multimodality.py:1173-1183: the (400, 300) array the numeric chain starts from, cut out of one raw
    (time, 401) parquet array.  Pure indexing (NB: the reference slices COLUMNS by the label offset, then
    transposes -- kept as it is)."""
    if offset_seconds is not None:
        off = int(offset_seconds // 2)
        basic = raw[:, off:off + 300]
        if basic.shape[1] < 300:
            basic = torch.nn.functional.pad(basic, (0, 300 - basic.shape[1]))
    else:
        basic = raw
    s = basic.transpose(0, 1)
    rows, cols = IMAGE_SIZE
    s = s[:rows, :cols]
    return torch.nn.functional.pad(s, (0, cols - s.shape[1], 0, rows - s.shape[0])).contiguous()


def process_spectrogram(window: torch.Tensor, check: bool = True) -> torch.Tensor:
    """``window``: (400, 300) from ``spectrogram_window`` or a batch (B, 400, 300), NaNs allowed.  Returns the
    (B, 3, 400, 300) fp32 tensor of ``process_spectrogram`` (NaN fill, baseline, 60 Hz notch forward-backward,
    Gaussian, min-max to [0, 1], three identical planes) on the device."""
    _require_cuda()
    single = window.dim() == 2
    w = _device_f32(window.unsqueeze(0) if single else window)
    if w.dim() != 3:
        raise ValueError("window must be (rows, cols) or (B, rows, cols)")
    batch, rows, cols = w.shape
    p = nat.ClsSpecParams()
    nat.lib.eegspec_cls_spec_params_default(C.byref(p))
    p.rows, p.cols = rows, cols
    out = torch.empty((batch, 3, rows, cols), dtype=torch.float32, device=w.device)
    if batch == 0:
        return out
    with torch.cuda.device(w.device):
        status = torch.empty(batch, dtype=torch.int32, device=w.device)
        ws_bytes = nat.lib.eegspec_classifier_spec_workspace(batch, C.byref(p))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=w.device)
        nat.check(nat.lib.eegspec_classifier_spec_f32(_ptr(w), batch, C.byref(p), _ptr(out), _ptr(status), _ptr(ws),
                                                      ws_bytes, _stream_ptr(w.device)),
                  "process_spectrogram")
        if check and bool(status.any()):
            bad = status.nonzero().flatten().tolist()
            raise ValueError(f"records {bad}: a spectrogram row is entirely NaN -- the reference drops it and resizes "
                             "the image with scikit-image (XAI_Multimodality.py:1214, :1303-1305); not reproduced")
    return out[0] if single else out
