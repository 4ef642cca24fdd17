"""B200-native EEG -> spectrogram transform (hand-written sm_100a CUDA behind a C ABI).

Drop-in for the one data-parallel hot path of
KC-decoder/Multimodal-Brain-Pattern-Identification_XAI:

* ``compute_stft(signal, config)``      -- /root/reference/root/src/utils/DiffEEG_utils.py:27-76
* ``compute_stft_single(signal, config)`` -- DiffEEG_utils.py:504-509
* ``eeg_to_spectrogram(eeg)``           -- the 4-chain x 128-mel image BASELINE.json names
* ``eeg_to_spectrogram_sliding(eeg)``   -- the same image for every overlapping window of one long record
* ``EEGTransformer(...).transform(x)``   -- ``_EEGTransformer``, /root/reference/root/src/data/dataset.py:48-104
* ``install()``                         -- rebinds ``compute_stft`` inside the reference's modules

The directory name is not a Python identifier; ``import b200spec`` (repo root) is the same
package under an importable name.  Attributes resolve lazily so that ``b200spec.build`` can be
imported before the CUDA library exists; everything else raises ImportError without it.
"""
import importlib as _importlib

_API = {
    "compute_stft": "transform",
    "compute_stft_single": "transform",
    "stft_image": "transform",
    "stft_plan": "transform",
    "interp_bracket": "transform",
    "eeg_to_spectrogram": "transform",
    "spectrogram_from_eeg": "transform",
    "eeg_to_spectrogram_sliding": "transform",
    "sliding_window_count": "transform",
    "mel_params": "transform",
    "melspec_geometry": "transform",
    "install": "transform",
    "launch_count": "transform",
    "spectrogram_condition": "diffeeg",
    "stft_condition": "diffeeg",
    "process_eeg": "classifier",
    "process_spectrogram": "classifier",
    "spectrogram_window": "classifier",
    "EEGTransformer": "eegtransformer",
    "xform_params": "eegtransformer",
    "eeg_transform_stft": "eegtransformer",
    "shard_bounds": "sharding",
    "shard_slice": "sharding",
    "gather_spectrograms": "sharding",
    "bind_host_to_gpu": "sharding",
    "EEG_COLUMNS": "montage",
    "MAP_FEATURES": "montage",
    "CHAIN_ORDER": "montage",
    "pair_indices": "montage",
    "crop_offset": "montage",
}
__all__ = sorted(_API)


def __getattr__(name):
    if name in _API:
        return getattr(_importlib.import_module("." + _API[name], __name__), name)
    raise AttributeError(f"module {__name__!r} has no attribute {name!r}")


def __dir__():
    return sorted(list(globals()) + list(_API))
