"""GPU parity, mode R: the CUDA path (through the C ABI) against the reference's own outputs
(golden), against the oracle port on seeded inputs, and size-independent properties at the
BASELINE shapes.  Tolerance: |gpu - reference| <= 1e-4 on the normalised [0,1] image
(north_star allows 1e-3); integer facts (shape, frame count, NaN pattern) are exact."""
import sys
import types

import numpy as np
import pytest
import torch

from conftest import Cfg, eeg_like
from oracle import stft_oracle as so

import b200spec

pytestmark = pytest.mark.gpu
ATOL = 1e-4

CASES = ["diffeeg_64_32", "hms50s_64_32", "nfft128", "hop16", "one_window", "nfft16",
         "nfft256_odd_len", "flat_rows", "nan_sample", "decimating", "gather"]


def run(x, nperseg=64, noverlap=32):
    y = b200spec.compute_stft(torch.from_numpy(np.ascontiguousarray(x)).cuda(), Cfg(nperseg, noverlap))
    torch.cuda.synchronize()
    return y.cpu().numpy()


@pytest.mark.parametrize("name", CASES)
def test_against_reference_golden(stft_golden, name):
    x, want = stft_golden[f"{name}/x"], stft_golden[f"{name}/y"]
    nperseg, noverlap = (int(v) for v in stft_golden[f"{name}/cfg"])
    got = run(x, nperseg, noverlap)
    assert got.shape == want.shape and got.dtype == np.float32
    np.testing.assert_array_equal(np.isnan(got), np.isnan(want))       # NaN rows: exact pattern
    err = np.nanmax(np.abs(got - want)) if not np.isnan(want).all() else 0.0
    assert err <= ATOL, f"{name}: max abs err {err:.3e}"
    fin = ~np.isnan(got)
    assert got[fin].min() >= 0.0 and got[fin].max() <= 1.0


@pytest.mark.parametrize("shape,cfg", [
    ((64, 19, 2000), (64, 32)),      # the DiffEEG micro-batch the reference runs
    ((4, 8, 2000), (64, 32)),        # magic-8 variant
    ((1, 20, 10000), (64, 32)),      # BASELINE config 1
    ((3, 5, 2000), (128, 64)),
    ((2, 3, 2000), (256, 128)),
    ((2, 2, 4096), (512, 256)),
    ((1, 2, 4096), (1024, 512)),
    ((2, 3, 1000), (32, 16)),
    ((1, 4, 3000), (64, 17)),        # odd hop
    ((1, 4, 777), (64, 0)),          # no overlap, length not a multiple of anything
])
def test_against_oracle_port(shape, cfg):
    rng = np.random.default_rng(abs(hash((shape, cfg))) % (2 ** 32))
    x = eeg_like(rng, shape)
    want = so.compute_stft_port(x[:8], Cfg(*cfg))          # the port is slow: first 8 records
    got = run(x, *cfg)
    assert got.shape == (shape[0], shape[1], cfg[0] // 2 + 1, 2000)
    err = np.abs(got[:8] - want).max()
    assert err <= ATOL, f"max abs err {err:.3e}"


@pytest.mark.parametrize("kind", ["randn", "walk", "clipped", "tiny", "huge"])
def test_input_distributions(kind):
    rng = np.random.default_rng(11)
    x = rng.standard_normal((2, 4, 2000)).astype(np.float32)
    if kind == "walk":
        x = np.cumsum(x, -1).astype(np.float32)
    elif kind == "clipped":
        x = np.clip(x * 2000, -1024, 1024).astype(np.float32)
    elif kind == "tiny":
        x = (x * 1e-4).astype(np.float32)
    elif kind == "huge":
        x = (x * 1e4).astype(np.float32)
    want = so.compute_stft_port(x, Cfg())
    err = np.abs(run(x) - want).max()
    assert err <= ATOL, f"{kind}: max abs err {err:.3e}"


def test_long_continuous_record_config4():
    """BASELINE config 4: one hour at 200 Hz, 22 501 frames squeezed into 2000 columns."""
    rng = np.random.default_rng(4)
    x = eeg_like(rng, (1, 2, 720000), walk=0.05)
    assert b200spec.stft_plan(720000, 64, 32)["n_frames"] == 22501
    want = so.compute_stft_port(x, Cfg())
    err = np.abs(run(x) - want).max()
    assert err <= ATOL, f"max abs err {err:.3e}"


def test_long_record_column_chunks_nan_rows_and_determinism():
    """Decimating plans run as column chunks with atomics for the row extrema: a NaN sample poisons exactly the
    frequency rows it reaches (all of them here: a NaN channel), other channels are untouched, results repeat."""
    rng = np.random.default_rng(41)
    x = eeg_like(rng, (2, 3, 300000), walk=0.05)                 # 9 376 frames > 2 x 2000 + 2: the decimating plan
    assert b200spec.stft_plan(300000, 64, 32)["gather"] == 1
    x[1, 2, 123456] = np.nan
    want = so.compute_stft_port(x, Cfg())
    got = run(x)
    assert np.isnan(got[1, 2]).any()
    np.testing.assert_array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(want)
    assert np.abs(got[ok] - want[ok]).max() <= ATOL
    assert np.array_equal(run(x), got, equal_nan=True)
    assert got[ok].min() >= 0.0 and got[ok].max() <= 1.0


def test_all_nan_channel_and_inf_free_rows():
    x = eeg_like(np.random.default_rng(5), (1, 3, 2000))
    x[0, 1] = np.nan
    got = run(x)
    assert np.isnan(got[0, 1]).all() and not np.isnan(got[0, 0]).any() and not np.isnan(got[0, 2]).any()


def test_layout_device_and_noncontiguous_input():
    rng = np.random.default_rng(6)
    base = eeg_like(rng, (4, 2000, 6))                         # (B, L, C): the dataset's layout
    sig = torch.from_numpy(base).cuda().permute(0, 2, 1)       # what __getitem__ hands over
    assert not sig.is_contiguous()
    out = b200spec.compute_stft(sig, Cfg())
    assert out.device == sig.device and out.dtype == torch.float32 and out.shape == (4, 6, 33, 2000)
    ref = b200spec.compute_stft(sig.contiguous(), Cfg())
    assert torch.equal(out, ref)
    dbl = b200spec.compute_stft(sig.double(), Cfg())           # non-fp32 input is cast like .float()
    assert torch.equal(dbl, ref)


def test_rows_are_independent_and_deterministic():
    rng = np.random.default_rng(8)
    x = torch.from_numpy(eeg_like(rng, (16, 19, 2000))).cuda()
    full = b200spec.compute_stft(x, Cfg())
    again = b200spec.compute_stft(x, Cfg())
    assert torch.equal(full, again)
    assert torch.equal(b200spec.compute_stft(x[5:6], Cfg()), full[5:6])      # small-grid column split
    perm = torch.randperm(16, device="cuda")
    assert torch.equal(b200spec.compute_stft(x[perm], Cfg()), full[perm])


def test_runs_on_the_callers_stream():
    x = torch.from_numpy(eeg_like(np.random.default_rng(9), (8, 19, 2000))).cuda()
    want = b200spec.compute_stft(x, Cfg())
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        got = b200spec.compute_stft(x, Cfg())
    s.synchronize()
    assert torch.equal(got, want)


def test_host_front_door_matches_device_path():
    x = torch.from_numpy(eeg_like(np.random.default_rng(10), (700, 2, 2000)))   # > one pipeline chunk
    out = b200spec.compute_stft(x, Cfg())
    assert out.device.type == "cpu" and out.shape == (700, 2, 33, 2000)
    assert torch.equal(out, b200spec.compute_stft(x.cuda(), Cfg()).cpu())


def test_compute_stft_single(stft_golden):
    x, want = stft_golden["single/x"], stft_golden["single/y"]
    got = b200spec.compute_stft_single(torch.from_numpy(x).cuda(), Cfg())
    assert got.shape == (33, 64)
    assert np.abs(got.cpu().numpy() - want).max() <= 1e-5 * max(1.0, want.max())
    cpu = b200spec.compute_stft_single(torch.from_numpy(x), Cfg())
    assert cpu.device.type == "cpu" and torch.equal(cpu, got.cpu())


def test_error_behaviour_matches_reference_stack():
    x = torch.zeros(1, 1, 2000, device="cuda")
    with pytest.raises(ValueError, match="noverlap must be less than nperseg"):
        b200spec.compute_stft(x, Cfg(64, 64))                  # SciPy's error, same text
    with pytest.raises(ValueError):
        b200spec.compute_stft(x[0], Cfg())                     # rank: `B, C, L = signal.shape`
    with pytest.raises(NotImplementedError):
        b200spec.compute_stft(x, Cfg(window="hamming"))
    with pytest.raises(NotImplementedError):
        b200spec.compute_stft(x, Cfg(48, 24))                  # not a power of two: refused, no fallback
    assert b200spec.compute_stft(torch.zeros(0, 19, 2000, device="cuda"), Cfg()).shape == (0, 19, 33, 2000)


def test_install_rebinds_the_reference_symbols():
    utils_pkg, train_pkg = types.ModuleType("utils"), types.ModuleType("training")
    m1, m2 = types.ModuleType("utils.DiffEEG_utils"), types.ModuleType("training.DiffEEG_train")
    m1.compute_stft = m2.compute_stft = lambda *a: None
    saved = {k: sys.modules.get(k) for k in ("utils", "training", m1.__name__, m2.__name__)}
    sys.modules.update({"utils": utils_pkg, "training": train_pkg, m1.__name__: m1, m2.__name__: m2})
    try:
        assert b200spec.install() == [m1.__name__, m2.__name__]
        assert m1.compute_stft is b200spec.compute_stft and m2.compute_stft is b200spec.compute_stft
        out = m2.compute_stft(torch.zeros(1, 2, 2000, device="cuda"), Cfg())
        assert out.shape == (1, 2, 33, 2000) and float(out.abs().max()) == 0.0   # flat rows -> 0
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_full_size_properties_config2():
    """BASELINE config 2 shape in mode R: (1024, 20, 10000) -> (1024, 20, 33, 2000), 5.4 GB."""
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn((1024, 20, 10000), generator=g, device="cuda")
    x = (x.cumsum(-1) * 0.3 + torch.randn((1024, 20, 10000), generator=g, device="cuda")).clamp_(-1024, 1024)
    out = b200spec.compute_stft(x, Cfg())
    assert out.shape == (1024, 20, 33, 2000)
    assert not torch.isnan(out).any()
    lo, hi = out.amin(dim=-1), out.amax(dim=-1)
    assert float(lo.min()) >= 0.0 and float(lo.max()) <= 1e-6          # every row touches 0 ...
    assert float(hi.max()) <= 1.0 and float(hi.min()) >= 1.0 - 1e-5    # ... and 1
    # endpoints are knots: column 0 / 1999 equal frame 0 / n-1, so a 3-record slice must agree
    sub = b200spec.compute_stft(x[500:503].contiguous(), Cfg())
    assert torch.equal(sub, out[500:503])
    want = so.compute_stft_port(x[777:778, :3].cpu().numpy(), Cfg())
    assert np.abs(out[777:778, :3].cpu().numpy() - want).max() <= ATOL
