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
