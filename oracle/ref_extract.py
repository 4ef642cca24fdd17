"""Run the reference's own ``compute_stft`` without importing its module.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

``import utils.DiffEEG_utils`` is impossible in this image (``ipywidgets``,
``matplotlib``, ... are missing and ``utils/cfg_utils.py`` creates directories under
``/data2`` at import time), so the ``FunctionDef`` of ``compute_stft``
(``/root/reference/root/src/utils/DiffEEG_utils.py:27-76``) is lifted out of the module's
AST and executed with exactly the three names it needs: ``np``, ``torch`` and
``scipy.signal.stft``.  No reference text is stored in this repository; the file is
read where it lies.  ``/root/reference`` exists only in the build container, so this
module is used only by ``tests/golden/make_golden.py`` and by tests that skip when
the reference tree is absent.
"""
from __future__ import annotations

import ast
import os
from types import SimpleNamespace

REFERENCE_ROOT = os.environ.get("EEGSPEC_REFERENCE_ROOT", "/root/reference")
_DIFFEEG_UTILS = os.path.join(REFERENCE_ROOT, "root", "src", "utils", "DiffEEG_utils.py")


def reference_available() -> bool:
    return os.path.isfile(_DIFFEEG_UTILS)


def _lift(func_name: str, nested_in: str | None = None):
    import numpy as np
    import torch
    from scipy.signal import stft

    with open(_DIFFEEG_UTILS, "r", encoding="utf-8") as fh:
        tree = ast.parse(fh.read(), filename=_DIFFEEG_UTILS)

    scope = tree.body
    if nested_in is not None:
        outer = next(n for n in scope if isinstance(n, ast.FunctionDef) and n.name == nested_in)
        scope = list(ast.walk(outer))
    node = next(n for n in scope if isinstance(n, ast.FunctionDef) and n.name == func_name)
    module = ast.Module(body=[node], type_ignores=[])
    ast.fix_missing_locations(module)
    env = {"np": np, "torch": torch, "stft": stft}
    exec(compile(module, _DIFFEEG_UTILS, "exec"), env)  # noqa: S102 - the reference's own code
    return env[func_name]


def load_reference_compute_stft():
    """The reference ``compute_stft(signal, config)`` (DiffEEG_utils.py:27-76)."""
    return _lift("compute_stft")


def load_reference_compute_stft_single():
    """The plotting twin nested in ``plot_spectrogram_comparison`` (DiffEEG_utils.py:504-509)."""
    return _lift("compute_stft_single", nested_in="plot_spectrogram_comparison")


def make_config(n_fft: int = 64, hop_length: int = 32, window: str = "hann"):
    """Any object with ``.diffEEG_trainer[...]`` works (DiffEEG_utils.py:49-51); the
    defaults are the shipped values (``root/src/utils/cfg_utils.py:168-171``)."""
    return SimpleNamespace(
        diffEEG_trainer={
            "stft_n_fft": n_fft,
            "stft_hop_length": hop_length,
            "stft_window": window,
        }
    )


# --------------------------------------------------------------------------- _EEGTransformer
_DATASET = os.path.join(REFERENCE_ROOT, "root", "src", "data", "dataset.py")
_CFG_UTILS = os.path.join(REFERENCE_ROOT, "root", "src", "utils", "cfg_utils.py")


def reference_cfg_channels():
    """``CFG.feats`` / ``CFG.channel_feats`` / ``CFG.EEG_PTS`` read from the literals of
    ``class CFG`` (/root/reference/root/src/utils/cfg_utils.py:70-96) without importing it."""
    with open(_CFG_UTILS, "r", encoding="utf-8") as fh:
        tree = ast.parse(fh.read(), filename=_CFG_UTILS)
    cfg = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "CFG")
    got = {}
    for node in cfg.body:
        if isinstance(node, ast.Assign) and len(node.targets) == 1 and isinstance(node.targets[0], ast.Name):
            name = node.targets[0].id
            if name in ("feats", "channel_feats", "EEG_FREQ", "EEG_WLEN"):
                got[name] = ast.literal_eval(node.value)
    return SimpleNamespace(feats=got["feats"], channel_feats=got["channel_feats"],
                           EEG_PTS=int(got["EEG_FREQ"] * got["EEG_WLEN"]))


def load_reference_eeg_transformer():
    """The reference ``_EEGTransformer`` class (/root/reference/root/src/data/dataset.py:48-147),
    lifted out of its module's AST and executed with the names it uses (``np``, SciPy's
    ``butter`` / ``lfilter``, ``Optional`` and a ``CFG`` holding the reference's own channel lists)."""
    from typing import Optional

    import numpy as np
    from scipy.signal import butter, lfilter

    with open(_DATASET, "r", encoding="utf-8") as fh:
        tree = ast.parse(fh.read(), filename=_DATASET)
    node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "_EEGTransformer")
    module = ast.Module(body=[node], type_ignores=[])
    ast.fix_missing_locations(module)
    env = {"np": np, "butter": butter, "lfilter": lfilter, "Optional": Optional, "CFG": reference_cfg_channels()}
    exec(compile(module, _DATASET, "exec"), env)  # noqa: S102 - the reference's own code
    return env["_EEGTransformer"]
