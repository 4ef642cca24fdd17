"""Electrode order and bipolar montage of the hot path (integer facts, bit-exact).

* ``EEG_COLUMNS``  -- channel order of a record: /root/reference/root/config/config.yml:34-54
* ``MAP_FEATURES`` -- the 18 differential pairs:   /root/reference/root/config/config.yml:529-547
  (consumed by ``calculate_differential_signals``, /root/reference/root/src/utils/data_utils.py:183-215)
* ``MAGIC8``       -- the 8 "magic" differences:   /root/reference/root/src/data/dataset.py:106-123

The first 16 pairs, four at a time, are the chains LL, LP, RP, RR of the spectrogram image.
"""
from __future__ import annotations

EEG_COLUMNS = ("Fp1", "F3", "C3", "P3", "F7", "T3", "T5", "O1", "Fz", "Cz", "Pz",
               "Fp2", "F4", "C4", "P4", "F8", "T4", "T6", "O2", "EKG")

CHAINS = {
    "LL": ("Fp1", "F7", "T3", "T5", "O1"),
    "LP": ("Fp1", "F3", "C3", "P3", "O1"),
    "RP": ("Fp2", "F8", "T4", "T6", "O2"),
    "RR": ("Fp2", "F4", "C4", "P4", "O2"),
}
CHAIN_ORDER = ("LL", "LP", "RP", "RR")

MAP_FEATURES = tuple(
    (CHAINS[c][i], CHAINS[c][i + 1]) for c in CHAIN_ORDER for i in range(4)
) + (("Fz", "Cz"), ("Cz", "Pz"))

MAGIC8 = (("Fp1", "T3"), ("T3", "O1"), ("Fp1", "C3"), ("C3", "O1"),
          ("Fp2", "C4"), ("C4", "O2"), ("Fp2", "T4"), ("T4", "O2"))


def channel_index(columns=EEG_COLUMNS) -> dict:
    return {name: i for i, name in enumerate(columns)}


def pair_indices(pairs=MAP_FEATURES[:16], columns=EEG_COLUMNS):
    """[(minuend channel, subtrahend channel), ...] as integer indices into ``columns``."""
    idx = channel_index(columns)
    return [(idx[a], idx[b]) for a, b in pairs]


def crop_offset(n_pts: int, window: int = 10_000) -> int:
    """Start of the middle window: /root/reference/root/src/utils/cfg_utils.py:369-371."""
    return (n_pts - window) // 2
