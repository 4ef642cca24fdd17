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


# --------------------------------------------------------------------------- generic lift
_DATA_UTILS = os.path.join(REFERENCE_ROOT, "root", "src", "utils", "data_utils.py")
_CONFIG_YML = os.path.join(REFERENCE_ROOT, "root", "config", "config.yml")


def lift_functions(path: str, names, env: dict) -> dict:
    """Execute the top-level ``def``s called ``names`` of the reference file ``path`` inside
    ``env`` (the free names they use) and return ``env``.  Nothing else of the module runs."""
    with open(path, "r", encoding="utf-8") as fh:
        tree = ast.parse(fh.read(), filename=path)
    body = [n for n in tree.body if isinstance(n, (ast.FunctionDef, ast.ClassDef)) and n.name in set(names)]
    missing = set(names) - {n.name for n in body}
    if missing:
        raise KeyError(f"{path}: no top-level definition of {sorted(missing)}")
    module = ast.Module(body=body, type_ignores=[])
    ast.fix_missing_locations(module)
    exec(compile(module, path, "exec"), env)  # noqa: S102 - the reference's own code
    return env


def reference_yaml_config() -> dict:
    """``root/config/config.yml`` as ``load_config`` reads it (config_loader.py:16-28; the
    ``${root_dir}`` substitution only touches path strings, which nothing here uses)."""
    import yaml

    with open(_CONFIG_YML, "r", encoding="utf-8") as fh:
        return yaml.safe_load(fh)


def load_reference_get_eeg_window(use_magic_8: bool = False):
    """The reference ``_get_eeg_window(file)`` (/root/reference/root/src/utils/cfg_utils.py:354-389):
    parquet -> middle 10 000-sample crop -> per-channel NaN fill.  ``CFG`` is a stub holding the
    reference's own literals (channel lists, ``cast_eegs = True``, the magic-8 switch)."""
    from pathlib import Path

    import numpy as np
    import pandas as pd

    lit = reference_cfg_channels()
    cfg = SimpleNamespace(feats=lit.feats, channel_feats=lit.channel_feats, EEG_PTS=lit.EEG_PTS, cast_eegs=True,
                          dataset={"eeg": {"apply_chris_magic_ch8": bool(use_magic_8)}})
    env = {"np": np, "pd": pd, "Path": Path, "CFG": cfg, "EEG_PTS": cfg.EEG_PTS}
    return lift_functions(_CFG_UTILS, ["_get_eeg_window"], env)["_get_eeg_window"], cfg


def load_reference_data_utils(names, feature_names=None):
    """Functions of ``root/src/utils/data_utils.py`` (e.g. ``calculate_differential_signals`` :183-215,
    ``butter_bandpass_filter`` :224-226, ``normalize`` :230-233, ``denoise_filter`` :236-240) with the
    module-level names they read: ``CFG`` = the YAML dict (:41) and ``feature_to_index`` (:42; built from
    ``feature_names``, default the YAML's own ``eeg_features``)."""
    import numpy as np
    from scipy.ndimage import gaussian_filter
    from scipy.signal import butter, filtfilt, iirnotch, lfilter

    cfg = reference_yaml_config()
    feats = list(feature_names) if feature_names is not None else cfg["eeg_features"]
    env = {"np": np, "butter": butter, "lfilter": lfilter, "filtfilt": filtfilt, "iirnotch": iirnotch,
           "gaussian_filter": gaussian_filter, "CFG": cfg,
           "feature_to_index": {x: y for x, y in zip(feats, range(len(feats)))}}
    try:
        from skimage.transform import resize
        env["resize"] = resize
    except Exception:   # skimage is not in this image; only resample_spectrogram needs it
        pass
    return lift_functions(_DATA_UTILS, names, env)


# --------------------------------------------------------------------------- DiffEEG (row f2)
_DIFFEEG_MODEL = os.path.join(REFERENCE_ROOT, "root", "src", "models", "diffEEG.py")


def load_reference_diffeeg(n_channels: int = 19, hidden_channels: int = 32, n_classes: int = 6, dropout: float = 0.1):
    """The reference ``DiffEEG`` module (/root/reference/root/src/models/diffEEG.py:20-145) on CPU, lifted with
    its ``GTU`` helper; ``config`` is a stub with the keys its constructor reads (:23-31)."""
    import random

    import numpy as np
    import torch
    import torch.nn as nn
    import torch.nn.functional as F
    import torch.utils.checkpoint as checkpoint

    env = {"torch": torch, "nn": nn, "F": F, "np": np, "random": random, "checkpoint": checkpoint}
    lift_functions(_DIFFEEG_MODEL, ["GTU", "DiffEEG"], env)
    cfg = SimpleNamespace(device=torch.device("cpu"),
                          diffEEG_trainer={"n_classes": n_classes, "n_channels": n_channels,
                                           "hidden_channels": hidden_channels, "dropout": dropout})
    return env["DiffEEG"](cfg)


# --------------------------------------------------------------------------- classifier-side dataset (rows f3, f4)
_NOTEBOOK = os.path.join(REFERENCE_ROOT, "root", "jupyter_notebooks", "XAI_Multimodality.py")


def load_reference_combined_dataset(eeg_frames=None, spectrogram_frames=None):
    """An instance of the reference's ``CombinedDataset`` (/root/reference/root/jupyter_notebooks/
    XAI_Multimodality.py:1114-1307) with its own ``CFG`` class (:93-260), both lifted from the notebook export.
    ``__init__`` is bypassed (it seeds RNGs and wants a metadata frame); the two loaders the methods call,
    ``load_train_eeg_frame`` / ``load_train_spectr_frame`` (:480-503), are replaced by look-ups in the given
    ``{id: DataFrame}`` dictionaries.  ``process_eeg(row)`` (:1148-1164) and ``process_spectrogram(row)``
    (:1166-1204) then run exactly as in the notebook.  ``resize`` (scikit-image, not installed here) is the
    identity for an image that already has the target shape -- the only case the fixtures exercise."""
    import random

    import numpy as np
    import pandas as pd
    import torch
    from scipy.ndimage import gaussian_filter
    from scipy.signal import butter, filtfilt, iirnotch, lfilter
    from torch.utils.data import Dataset

    def resize(sig, target_shape, mode="reflect", anti_aliasing=True):
        if tuple(sig.shape) != tuple(target_shape):
            raise NotImplementedError("skimage.transform.resize is not installed: only the identity case runs here")
        return sig

    def labels_to_probabilities(labels, num_classes):   # :630-632
        return torch.eye(num_classes)[labels]

    env = {"np": np, "pd": pd, "torch": torch, "random": random, "os": os, "Dataset": Dataset, "butter": butter,
           "lfilter": lfilter, "filtfilt": filtfilt, "iirnotch": iirnotch, "gaussian_filter": gaussian_filter,
           "resize": resize, "labels_to_probabilities": labels_to_probabilities,
           "load_train_eeg_frame": lambda i: (eeg_frames or {})[int(i)],
           "load_train_spectr_frame": lambda i: (spectrogram_frames or {})[int(i)]}
    with open(_NOTEBOOK, "r", encoding="utf-8") as fh:
        tree = ast.parse(fh.read(), filename=_NOTEBOOK)
    wanted = {"CFG": None, "CombinedDataset": None}
    for node in tree.body:                      # the FIRST definition of each (the export repeats some cells)
        if isinstance(node, ast.ClassDef) and node.name in wanted and wanted[node.name] is None:
            wanted[node.name] = node
    module = ast.Module(body=[wanted["CFG"], wanted["CombinedDataset"]], type_ignores=[])
    ast.fix_missing_locations(module)
    exec(compile(module, _NOTEBOOK, "exec"), env)  # noqa: S102 - the reference's own code
    cls = env["CombinedDataset"]
    ds = cls.__new__(cls)
    ds.cfg = env["CFG"]
    ds.training_flag, ds.augmentations, ds.plot = False, None, False
    ds.feature_to_index = ds.cfg.feature_to_index                         # :1126-1127
    ds.differential_channels_start_index = len(ds.feature_to_index)
    return ds
