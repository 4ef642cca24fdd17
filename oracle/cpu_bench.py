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
    if mode == "f2":   # compute_stft images of the DiffEEG micro-batch
        return rng.random((n, 19, 33, 2000), dtype=np.float32)
    if mode == "f3":   # (400, 300) Kaggle-spectrogram windows
        return rng.random((n, 400, 300), dtype=np.float32) * np.float32(40.0)
    if mode == "f4":   # parquet rows (n, 20)
        return (np.cumsum(rng.standard_normal((n, 10000, 20), dtype=np.float32), axis=1) * np.float32(0.5)
                + rng.standard_normal((n, 10000, 20), dtype=np.float32) * np.float32(5.0))
    shape = (n, 20, 10000) if mode in ("n", "r50") else (n, 19, 2000)
    x = np.cumsum(rng.standard_normal(shape, dtype=np.float32), axis=-1) * np.float32(0.3)
    x += rng.standard_normal(shape, dtype=np.float32)
    return np.clip(x, -1024, 1024).astype(np.float32)


class _Cfg:
    diffEEG_trainer = {"stft_n_fft": 64, "stft_hop_length": 32, "stft_window": "hann"}


def _compute(mode: str, x: np.ndarray) -> float:
    from oracle import melspec_oracle, stft_oracle

    if mode == "xf":
        from oracle import eegtransform_oracle

        out = eegtransform_oracle.eeg_transform_batch_oracle(x, n_feats=19, downsample=5)
    elif mode == "f2":
        import torch

        from oracle import diffeeg_oracle

        torch.set_num_threads(1)
        g = torch.Generator().manual_seed(0)
        w, b = torch.randn((19, 16, 3, 3), generator=g) * 0.1, torch.zeros(16)
        out = diffeeg_oracle.spectrogram_condition_oracle(x, w, b, 2000).numpy()
    elif mode == "f3":
        from oracle import classifier_oracle

        out = np.stack([classifier_oracle.process_spectrogram_oracle(r) for r in x])
    elif mode == "f4":
        from oracle import classifier_oracle

        out = np.stack([classifier_oracle.process_eeg_oracle(r) for r in x])
    elif mode == "n":
        out = melspec_oracle.eeg_to_spectrogram_batch_oracle(x)
    else:
        out = stft_oracle.compute_stft_port(x, _Cfg)
    return float(out.mean())


def _worker(conn):
    """One worker process: inputs are synthesised when the task arrives and the clock only runs
    between "go" and the last byte of the result -- input generation is never timed."""
    from oracle import melspec_oracle, stft_oracle  # noqa: F401  (imports dominate spawn cost)
    import scipy.signal  # noqa: F401

    conn.send(("up", os.getpid()))
    cache = {}
    while True:
        msg = conn.recv()
        if msg[0] == "quit":
            return
        _, mode, n, seed = msg
        key = (mode, n, seed)
        if key not in cache:
            cache.clear()                       # one resident input set per worker
            cache[key] = synth_records(mode, n, seed)
        conn.send(("ready",))
        conn.recv()                             # "go"
        t0 = time.perf_counter()
        chk = _compute(mode, cache[key])
        conn.send(("done", time.perf_counter() - t0, chk))


class CpuArm:
    """One worker process per host core.  ``run(mode, n)`` hands every worker ``n / cores`` records,
    waits until all of them hold their inputs, then times "go" -> all results back: the wall clock
    covers the transform only (synthesis and task dispatch happen before it starts)."""

    def __init__(self, cores: int | None = None):
        import multiprocessing as mp

        self.cores = cores or host_cores()
        ctx = mp.get_context("spawn")
        self.workers = []
        for _ in range(self.cores):
            parent, child = ctx.Pipe()
            proc = ctx.Process(target=_worker, args=(child,), daemon=True)
            proc.start()
            self.workers.append((proc, parent))
        for _, conn in self.workers:
            assert conn.recv()[0] == "up"

    def run(self, mode: str, n_records: int, seed: int = 0) -> dict:
        per = max(1, n_records // self.cores)
        for i, (_, conn) in enumerate(self.workers):
            conn.send(("task", mode, per, seed + i))
        for _, conn in self.workers:
            assert conn.recv()[0] == "ready"
        t0 = time.perf_counter()
        for _, conn in self.workers:
            conn.send(("go",))
        res = [conn.recv() for _, conn in self.workers]
        wall = time.perf_counter() - t0
        done = per * self.cores
        return {"records": done, "per_worker": per, "wall_s": wall, "records_per_s": done / wall,
                "per_core_records_per_s": done / sum(r[1] for r in res)}

    def close(self):
        for proc, conn in self.workers:
            try:
                conn.send(("quit",))
            except (BrokenPipeError, OSError):
                pass
        for proc, _ in self.workers:
            proc.join(timeout=5)
        self.workers = []
