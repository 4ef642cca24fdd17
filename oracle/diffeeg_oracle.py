"""Row (f)2 oracle: the spectrogram-condition front end of the DiffEEG denoiser.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Reference: ``/root/reference/root/src/models/diffEEG.py`` -- constructor :47-53
(``ConvTranspose2d(n_channels -> hidden/2, kernel (3, 3), stride (1, 8), padding (1, 2))``),
forward :108-110 (``relu`` -> ``flatten(start_dim=2)`` -> ``F.interpolate(size=T, mode="linear",
align_corners=False)``) and the within-class mixing of the training mode :136-145.

PINNED: ``tests/golden/diffeeg_golden.npz`` holds (a) the tensor the reference's own ``DiffEEG.forward``
hands to ``channel_expand`` (captured with a forward-pre-hook on the AST-lifted module, seeded weights;
the reference runs this under ``torch.amp.autocast`` -> bfloat16 on CPU, so that vector pins structure --
layout, kernel orientation, padding, interpolation indices -- to ~1e-2) and (b) the output of the lifted
module's own ``spectrogram_upsample1`` layer object in fp32 followed by the three forward lines, which
this restatement must reproduce to fp32 rounding.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn.functional as F

STRIDE = (1, 8)      # diffEEG.py:51
PADDING = (1, 2)     # diffEEG.py:52


def recombine(spec: torch.Tensor, partner: torch.Tensor | None, alpha: float = 0.5) -> torch.Tensor:
    """diffEEG.py:136-145 with the random permutation made explicit: ``partner[b]`` is the record
    ``b`` is mixed with (``partner[b] == b`` leaves it alone, which is what the reference does for a
    class with a single member)."""
    if partner is None:
        return spec
    return alpha * spec + (1.0 - alpha) * spec[partner.long()]


def spectrogram_condition_oracle(spec, weight, bias, out_length: int, partner=None, alpha: float = 0.5):
    """(B, C, F, Tc) image, ConvTranspose2d weight (C, H2, 3, 3) and bias (H2) -> (B, H2, out_length) fp32."""
    spec = torch.as_tensor(spec, dtype=torch.float32)
    weight = torch.as_tensor(weight, dtype=torch.float32)
    bias = torch.as_tensor(bias, dtype=torch.float32)
    if partner is not None:
        spec = recombine(spec, torch.as_tensor(partner), alpha)
    y = F.conv_transpose2d(spec, weight, bias, stride=STRIDE, padding=PADDING)      # :47-53, :108
    y = F.relu(y).flatten(start_dim=2)                                              # :109
    return F.interpolate(y, size=out_length, mode="linear", align_corners=False)    # :110


def source_positions(in_len: int, out_len: int):
    """Index / weight of ``F.interpolate(mode="linear", align_corners=False)`` for fp32 tensors: ATen
    computes ``scale = float(in) / float(out)`` and ``src = scale * (j + 0.5) - 0.5`` in fp32."""
    scale = np.float32(in_len) / np.float32(out_len)
    j = np.arange(out_len, dtype=np.float32)
    src = (scale * (j + np.float32(0.5)) - np.float32(0.5)).astype(np.float32)
    src = np.maximum(src, np.float32(0))
    i0 = src.astype(np.int64)
    i1 = i0 + (i0 < in_len - 1)
    lam1 = (src - i0.astype(np.float32)).astype(np.float32)
    return i0, i1, lam1
