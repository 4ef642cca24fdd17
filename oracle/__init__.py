"""CPU oracle for the EEG -> spectrogram hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product: only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and only as the checker / the timed CPU baseline.  The
product path (the package next to this directory) never imports it and has no CPU
fallback.

Parity status
-------------
* mode R (``compute_stft``): PINNED.  ``oracle/stft_oracle.py`` is checked against
  golden vectors produced by executing the reference's own function
  (``/root/reference/root/src/utils/DiffEEG_utils.py:27-76``, AST-extracted by
  ``oracle/ref_extract.py``; fixtures + generator script in ``tests/golden/``).
* row f1 (``_EEGTransformer``): PINNED the same way (``oracle/eegtransform_oracle.py``
  against the lifted class, ``tests/golden/eegtransform_golden.npz``).
* mode N front half (crop, per-series NaN fill, montage differences): PINNED.
  ``crop_offset`` / ``fill_missing`` / ``montage_differences`` of
  ``oracle/melspec_oracle.py`` reproduce, bit for bit, golden vectors produced by the
  reference's own ``_get_eeg_window`` (cfg_utils.py:354-389, run on parquet files) and
  ``calculate_differential_signals`` (data_utils.py:183-215)
  (``tests/golden/fronthalf_golden.npz``).
* mode N back half (mel / dB image): PARITY UNPINNED by the reference -- it does
  not exist in ``/root/reference`` (no librosa anywhere); it is specified only by
  ``BASELINE.json``.  ``oracle/melspec_oracle.py`` restates the published librosa
  algorithm (librosa is not installed; 0.10.x semantics).  Narrowed, not closed, by
  a second END-TO-END implementation from library code only (torchaudio
  ``MelSpectrogram`` + ``amplitude_to_DB``; ``tests/golden/melspec_xcheck.npz``,
  generator ``tests/golden/make_golden.py::xcheck_image``): oracle and CUDA path are
  both compared with it (agreement 4e-6 on the (dB+40)/40 image).
"""
