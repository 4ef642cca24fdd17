"""CPU oracle for the EEG -> spectrogram hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is part of the product: only ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs of
``bench.py`` may import it, and only as the checker / the timed CPU baseline.  The
product path (the package next to this directory) never imports it and has no CPU
fallback.

Parity status
-------------
* mode R (``compute_stft``): PINNED.  ``oracle/stft_port.py`` is checked against
  golden vectors produced by executing the reference's own function
  (``/root/reference/root/src/utils/DiffEEG_utils.py:27-76``, AST-extracted by
  ``oracle/ref_extract.py``; fixtures + generator script in ``tests/golden/``).
* mode N (montage -> mel -> dB image): PARITY UNPINNED by the reference.  The mel
  back half does not exist in ``/root/reference`` (no librosa anywhere); it is
  specified only by ``BASELINE.json``.  ``oracle/melspec_restate.py`` restates the
  published librosa algorithm (librosa is not installed; 0.10.x semantics) and is
  cross-checked against ``torchaudio``'s Slaney filterbank and ``torch.stft``.
  The front half (channel order, pair list, crop, NaN fill) is pinned to the
  reference lines cited in that file.
"""
