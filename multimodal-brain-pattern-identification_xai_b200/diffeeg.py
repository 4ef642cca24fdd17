"""Row (f)2: the spectrogram-condition front end of the DiffEEG denoiser on the GPU.

Mirror of the condition branch of ``DiffEEG.forward`` (/root/reference/root/src/models/diffEEG.py):

    spec = self.spectrogram_upsample1(spectrogram)                           # :108   ConvTranspose2d, :47-53
    spec = F.relu(spec).flatten(start_dim=2)                                 # :109
    spec = F.interpolate(spec, size=T, mode="linear", align_corners=False)   # :110

preceded, in training mode, by ``recombine_spectrograms`` (:136-145).  ``spectrogram_condition`` returns the
tensor the reference hands to ``channel_expand`` (:111), (B, hidden // 2, T) fp32.  The arithmetic runs in
``csrc/spec_condition.cu`` behind ``eegspec_spec_condition_f32``; the (B, hidden // 2, 33, 15991) intermediate
of the reference is never materialised.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _native as nat
from .transform import _ptr, _require_cuda, _stream_ptr


def spectrogram_condition(spectrogram: torch.Tensor, weight: torch.Tensor, bias: torch.Tensor, out_length: int,
                          partner: Optional[torch.Tensor] = None, alpha: float = 0.5) -> torch.Tensor:
    """``spectrogram``: (B, C, F, Tc) fp32 on the device (what ``compute_stft`` returns); ``weight`` / ``bias``:
    the parameters of ``model.spectrogram_upsample1`` ((C, H2, 3, 3) and (H2,)); ``out_length``: ``x.shape[-1]`` of
    the denoiser input.  ``partner`` (training mode): int tensor (B,), record ``b`` is mixed with ``partner[b]``
    as ``alpha * spec[b] + (1 - alpha) * spec[partner[b]]`` -- the reference draws it with ``torch.randperm``
    inside every class (:139-143); pass ``None`` for evaluation mode."""
    _require_cuda()
    if spectrogram.dim() != 4:
        raise ValueError("spectrogram must be (B, C, F, T)")
    if not spectrogram.is_cuda:
        raise ValueError("spectrogram_condition expects device tensors (the image compute_stft returns)")
    spec = spectrogram.to(torch.float32).contiguous()
    dev = spec.device
    w = weight.detach().to(device=dev, dtype=torch.float32).contiguous()
    b = bias.detach().to(device=dev, dtype=torch.float32).contiguous()
    batch, n_ch, n_freqs, n_cols = spec.shape
    if w.dim() != 4 or w.shape[0] != n_ch or tuple(w.shape[2:]) != (3, 3) or b.shape != (w.shape[1],):
        raise ValueError("weight must be the (C, H2, 3, 3) ConvTranspose2d parameter and bias (H2,)")
    p = nat.CondParams(n_ch, n_freqs, n_cols, int(w.shape[1]), int(out_length), float(alpha))
    out = torch.empty((batch, w.shape[1], int(out_length)), dtype=torch.float32, device=dev)
    if batch == 0:
        return out
    part = None
    if partner is not None:
        part = partner.to(device=dev, dtype=torch.int32).contiguous()
        if part.shape != (batch,) or int(part.min()) < 0 or int(part.max()) >= batch:
            raise ValueError("partner must hold one record index per record")
    with torch.cuda.device(dev):
        nat.check(nat.lib.eegspec_spec_condition_f32(_ptr(spec), batch, C.byref(p), _ptr(w), _ptr(b), _ptr(part),
                                                     _ptr(out), _stream_ptr(dev)),
                  "spectrogram_condition")
    return out


def stft_condition(signal: torch.Tensor, config, weight: torch.Tensor, bias: torch.Tensor,
                   out_length: Optional[int] = None, partner: Optional[torch.Tensor] = None,
                   alpha: float = 0.5) -> torch.Tensor:
    """``spectrogram_condition(compute_stft(signal, config), weight, bias, T)`` without the image.

    The reference's trainer computes the (B, C, 33, 2000) image (``DiffEEG_train.py:131``) only for the model to sample
    0.76 % of its transposed convolution (``diffEEG.py:108-110``).  Every pixel of that image is
    ``sat(value + delta * fraction)`` of two of the row's 64 normalised STFT knots, so the STFT kernel hands over the
    knots (17 KB per row instead of 264 KB) and the condition kernel evaluates the pixels it needs from them.
    Bit-identical to the two-call form; ``signal``: (B, C, L) fp32 on the device, ``out_length`` defaults to L."""
    from .transform import EPS, OUT_COLS, _stft_keys, stft_plan

    _require_cuda()
    nperseg, noverlap = _stft_keys(config)
    batch, n_ch, length = signal.shape
    if not signal.is_cuda:
        raise ValueError("stft_condition expects a device tensor")
    x = signal.to(torch.float32).contiguous()
    dev = x.device
    w = weight.detach().to(device=dev, dtype=torch.float32).contiguous()
    b = bias.detach().to(device=dev, dtype=torch.float32).contiguous()
    if w.dim() != 4 or w.shape[0] != n_ch or tuple(w.shape[2:]) != (3, 3) or b.shape != (w.shape[1],):
        raise ValueError("weight must be the (C, H2, 3, 3) ConvTranspose2d parameter and bias (H2,)")
    t_out = int(out_length) if out_length is not None else length
    out = torch.empty((batch, w.shape[1], t_out), dtype=torch.float32, device=dev)
    if batch == 0:
        return out
    plan = stft_plan(length, nperseg, noverlap)
    n_freqs, n_frames = plan["n_freqs"], plan["n_frames"]
    p = nat.CondParams(n_ch, n_freqs, OUT_COLS, int(w.shape[1]), t_out, float(alpha))
    part = None
    if partner is not None:
        part = partner.to(device=dev, dtype=torch.int32).contiguous()
        if part.shape != (batch,) or int(part.min()) < 0 or int(part.max()) >= batch:
            raise ValueError("partner must hold one record index per record")
    with torch.cuda.device(dev):
        knots = torch.empty((batch, n_ch, n_freqs, n_frames, 2), dtype=torch.float32, device=dev)
        nat.check(nat.lib.eegspec_stft_knots_f32(_ptr(x), batch * n_ch, length, length, nperseg, noverlap, OUT_COLS, EPS,
                                                 _ptr(knots), _stream_ptr(dev)), "stft_condition (knots)")
        nat.check(nat.lib.eegspec_spec_condition_knots_f32(_ptr(knots), n_frames, batch, C.byref(p), _ptr(w), _ptr(b),
                                                           _ptr(part), _ptr(out), _stream_ptr(dev)),
                  "stft_condition")
    return out
