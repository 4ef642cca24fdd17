"""Timed CPU baseline: the oracle port of the reference path on the host cores.

TEST/BENCH INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Imported by ``bench.py`` for its
``cpu_baseline`` leg and for ``--impl reference``.  The reference is Python and cannot travel to
the GPU box, so ``kind`` is ``"port"``: ``compute_stft_port`` makes the same SciPy / NumPy calls,
the same number of times, as ``/root/reference/root/src/utils/DiffEEG_utils.py:27-76``; the mel
image uses the librosa restatement (librosa itself is not installed).

The reference is a single-threaded Python loop; to give it "all the host threads it can use"
records are farmed out to one worker process per core (BLAS/OpenMP pinned to one thread each),
which is how its authors parallelise it too (DataLoader workers).
"""
from __future__ import annotations

import os
import time

import numpy as np

for _v in ("OMP_NUM_THREADS", "MKL_NUM_THREADS", "OPENBLAS_NUM_THREADS"):
    os.environ.setdefault(_v, "1")


def host_cores() -> int:
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:  # pragma: no cover
        return max(1, os.cpu_count() or 1)


def synth_records(mode: str, n: int, seed: int) -> np.ndarray:
    """Same distribution as bench.py's device-side generator (random walk + noise, clipped)."""
    rng = np.random.default_rng(seed)
    if mode == "xf":   # raw EEG, time-major (L, C) like the reference's parquet windows
        return (rng.standard_normal((n, 10000, 20), dtype=np.float32) * np.float32(40.0))
    shape = (n, 20, 10000) if mode in ("n", "r50") else (n, 19, 2000)
    x = np.cumsum(rng.standard_normal(shape, dtype=np.float32), axis=-1) * np.float32(0.3)
    x += rng.standard_normal(shape, dtype=np.float32)
    return np.clip(x, -1024, 1024).astype(np.float32)


class _Cfg:
    diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}


def _work(task):
    mode, n, seed = task
    from oracle import melspec_oracle, stft_oracle

    x = synth_records(mode, n, seed)
    t0 = time.perf_counter()
    if mode == "xf":
        from oracle import eegtransform_oracle

        out = eegtransform_oracle.eeg_transform_batch_oracle(x, n_feats=19, downsample=5)
    elif mode == "n":
        out = melspec_oracle.eeg_to_spectrogram_batch_oracle(x)
    else:
        out = stft_oracle.compute_stft_port(x, _Cfg)
    return time.perf_counter() - t0, float(out.mean())


def _warm(_):
    from oracle import melspec_oracle, stft_oracle  # noqa: F401  (imports dominate spawn cost)
    import scipy.signal  # noqa: F401

    return os.getpid()


class CpuArm:
    """A pool of worker processes; ``run(mode, n_records)`` returns records/s over all of them."""

    def __init__(self, cores: int | None = None):
        import multiprocessing as mp

        self.cores = cores or host_cores()
        self.pool = mp.get_context("spawn").Pool(self.cores) if self.cores > 1 else None
        if self.pool is not None:
            self.pool.map(_warm, range(self.cores * 2))

    def run(self, mode: str, n_records: int, seed: int = 0) -> dict:
        per = max(1, n_records // self.cores)
        tasks = [(mode, per, seed + i) for i in range(self.cores)]
        t0 = time.perf_counter()
        res = self.pool.map(_work, tasks, chunksize=1) if self.pool else [_work(t) for t in tasks]
        wall = time.perf_counter() - t0
        done = per * self.cores
        return {"records": done, "wall_s": wall, "records_per_s": done / wall,
                "per_core_records_per_s": done / sum(r[0] for r in res)}

    def close(self):
        if self.pool is not None:
            self.pool.close()
            self.pool.join()
            self.pool = None
