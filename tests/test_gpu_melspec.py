"""GPU parity, mode N: montage -> mel -> dB image against the oracle restatement.

Parity of the mel back half is UNPINNED by the reference (it has no such code); the
tolerance is the north_star's: |gpu - oracle| <= 1e-3 on the (dB+40)/40 image.  Integer
facts (shape, crop offset, pair indices, frame count, chain order) are exact."""
import numpy as np
import pytest
import torch

from conftest import eeg_like
from oracle import melspec_oracle as mo

import b200spec

pytestmark = pytest.mark.gpu
ATOL = 1e-3


def run(x, **kw):
    y = b200spec.eeg_to_spectrogram(torch.from_numpy(np.ascontiguousarray(x)).cuda(), **kw)
    torch.cuda.synchronize()
    return y.cpu().numpy()


ENGINES = [pytest.param(3, id="chain"), pytest.param(2, id="tcgen05"), pytest.param(1, id="simt")]   # 0 (default) == 3 here


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("case,key,mode", [("clean", "y", "constant"), ("clean", "y_reflect", "reflect"),
                                            ("nan", "y", "constant"), ("range", "y", "constant")])
def test_against_golden(melspec_golden, case, key, mode, engine):
    want = melspec_golden[f"{case}/{key}"]
    got = run(melspec_golden[f"{case}/x"], pad_mode=mode, engine=engine)
    assert got.shape == want.shape == (1, 128, 256, 4) and got.dtype == np.float32
    err = np.abs(got - want).max()
    assert err <= ATOL, f"{case}/{mode}: max abs err {err:.3e}"


@pytest.mark.parametrize("engine", ENGINES)
@pytest.mark.parametrize("case", ["clean", "nan", "range"])
def test_against_independent_torchaudio_image(melspec_golden, melspec_xcheck, case, engine):
    """The CUDA path against the image torchaudio's library code produces (no oracle code involved)."""
    got = run(melspec_golden[f"{case}/x"], engine=engine)
    err = np.abs(got - melspec_xcheck[f"{case}/y"]).max()
    assert err <= ATOL, f"{case}: max abs err {err:.3e}"


@pytest.mark.parametrize("engine", ENGINES)
def test_against_oracle_batch(engine):
    rng = np.random.default_rng(21)
    x = eeg_like(rng, (5, 20, 10000))
    want = mo.eeg_to_spectrogram_batch_oracle(x[:3])
    got = run(x, engine=engine)
    assert got.shape == (5, 128, 256, 4)
    err = np.abs(got[:3] - want).max()
    assert err <= ATOL, f"max abs err {err:.3e}"


@pytest.mark.parametrize("scale", [1e-6, 1e-3, 1.0, 1e3, 1e6])
def test_tensor_core_engine_is_scale_invariant(scale):
    """The fp16 operand split is protected by an exact power-of-two pre-scale: amplitudes from
    nanovolts to volts give the image the oracle gives (dB referencing removes the scale)."""
    rng = np.random.default_rng(26)
    x = eeg_like(rng, (1, 20, 10000)) * np.float32(scale)
    want = mo.eeg_to_spectrogram_batch_oracle(x)
    for engine in (3, 2, 1):
        err = np.abs(run(x, engine=engine) - want).max()
        assert err <= ATOL, f"engine {engine} scale {scale}: max abs err {err:.3e}"


def test_engines_agree_at_size():
    rng = np.random.default_rng(27)
    x = eeg_like(rng, (40, 20, 10000))
    a, b, c = run(x, engine=2), run(x, engine=1), run(x, engine=3)
    assert np.abs(a - b).max() <= 2e-4 and np.abs(c - b).max() <= 2e-4


def test_clip_level_and_last_frame_maximum():
    """A loud burst in the last 40 samples puts the pair maximum into frame 256 (beyond the kept
    width) and most of the image more than 80 dB below it: exercises the frames that only feed
    the maximum and the deferred clip (the speculative no-clip pass must be corrected).  A
    second record has the burst in the middle (clip without the extra frame)."""
    rng = np.random.default_rng(29)
    x = eeg_like(rng, (3, 20, 10000)) * np.float32(1e-2)
    x[0, :, -40:] += (rng.standard_normal((20, 40)) * 3e3).astype(np.float32)
    x[1, :, 5000:5100] += (rng.standard_normal((20, 100)) * 3e3).astype(np.float32)
    want = mo.eeg_to_spectrogram_batch_oracle(x)
    assert (want[0] <= -1.0 + 1e-6).mean() > 0.3 and (want[1] <= -1.0 + 1e-6).mean() > 0.3   # clipped pixels exist
    for engine in (3, 1):
        err = np.abs(run(x, engine=engine) - want).max()
        assert err <= ATOL, f"engine {engine}: max abs err {err:.3e}"


def test_single_record_time_major_and_crop():
    rng = np.random.default_rng(22)
    x = eeg_like(rng, (20, 18000))                           # 90 s: middle 50 s is used
    want = mo.eeg_to_spectrogram_oracle(x)
    got = run(x)
    assert got.shape == (128, 256, 4)
    assert np.abs(got - want).max() <= ATOL
    tm = b200spec.spectrogram_from_eeg(torch.from_numpy(np.ascontiguousarray(x.T)).cuda())
    assert np.abs(tm.cpu().numpy() - got).max() == 0.0       # parquet layout: electrode rows gathered, same engine
    mid = run(x[:, 4000:14000])
    assert np.abs(mid - got).max() == 0.0                    # crop offset is exact integer math


def test_time_major_batch_equals_channel_major():
    """The parquet layout (B, L, 20) goes through the row gather and the default engine: bit-identical
    to the channel-major call, for a batch that spans more than one gather pass' worth of clusters."""
    rng = np.random.default_rng(26)
    x = torch.from_numpy(eeg_like(rng, (90, 20, 10400))).cuda()
    x[3, 5, 700:900] = float("nan")
    cm = b200spec.eeg_to_spectrogram(x)
    tm = b200spec.eeg_to_spectrogram(x.transpose(1, 2).contiguous(), time_major=True)
    assert torch.equal(cm, tm)
    odd = x[:, :19].contiguous()                               # 19 channels: unaligned time-major rows (scalar gather)
    assert torch.equal(b200spec.eeg_to_spectrogram(odd), b200spec.eeg_to_spectrogram(odd.transpose(1, 2).contiguous(), time_major=True))


def test_chain_order_and_pair_wiring():
    """Silence every electrode of one chain: that chain -- and only that chain -- goes to 1.0."""
    rng = np.random.default_rng(23)
    x = eeg_like(rng, (1, 20, 10000))
    cols = {n: i for i, n in enumerate(mo.EEG_COLUMNS)}
    for name in ("Fp2", "F4", "C4", "P4", "O2"):             # RR = chain 3
        x[0, cols[name]] = 0.0
    got = run(x)[0]
    assert np.all(got[:, :, 3] == 1.0)                       # all-zero pair: dB 0 everywhere -> (0+40)/40
    assert got[:, :, :3].max() < 1.0 or (got[:, :, :3] < 1.0).mean() > 0.99
    want = mo.eeg_to_spectrogram_oracle(x[0])
    assert np.abs(got - want).max() <= ATOL


def test_custom_montage_and_shorter_chains():
    rng = np.random.default_rng(24)
    x = eeg_like(rng, (2, 20, 10000))
    pairs = [(0, 5), (5, 7), (11, 16), (16, 18)]             # two chains of two pairs
    got = run(x, pairs=pairs, pairs_per_chain=2)
    assert got.shape == (2, 128, 256, 2)
    fb = mo.mel_filterbank()
    want = np.zeros((128, 256, 2), np.float32)
    for c in range(2):
        for k in range(2):
            a, b = pairs[2 * c + k]
            db = mo.pair_melspec_db(mo.fill_missing(x[0, a] - x[0, b]), fb)
            want[:, :, c] += (db[:, :256] + 40) / 40
        want[:, :, c] /= 2
    assert np.abs(got[0] - want).max() <= ATOL


def test_host_front_door_and_determinism():
    rng = np.random.default_rng(25)
    x = torch.from_numpy(eeg_like(rng, (130, 20, 10000)))    # three pipeline chunks
    out = b200spec.eeg_to_spectrogram(x)
    assert out.device.type == "cpu" and out.shape == (130, 128, 256, 4)
    dev = b200spec.eeg_to_spectrogram(x.cuda())
    assert torch.equal(out, dev.cpu())
    assert torch.equal(dev, b200spec.eeg_to_spectrogram(x.cuda()))
    # the host call copies only the electrode rows the montage reads (0-7 and 11-18 of EEG_COLUMNS):
    # Fz, Cz, Pz and EKG never travel, so poisoning them changes nothing
    y = x.clone()
    y[:, [8, 9, 10, 19]] = float("nan")
    assert torch.equal(out, b200spec.eeg_to_spectrogram(y.pin_memory()))
    import ctypes
    from b200spec import _native as nat
    h2d, d2h = ctypes.c_int64(), ctypes.c_int64()
    p = b200spec.mel_params()
    nat.check(nat.lib.eegspec_melspec_host_bytes(130, 20, 10000, ctypes.byref(p), ctypes.byref(h2d), ctypes.byref(d2h)), "bytes")
    assert h2d.value == 130 * 16 * 10000 * 4 and d2h.value == out.numel() * 4
    tm = b200spec.eeg_to_spectrogram(x[:3].transpose(1, 2).contiguous(), time_major=True)   # parquet layout: one flat copy
    assert torch.equal(tm, out[:3])                              # gathered into electrode rows, same engine


def test_errors():
    with pytest.raises(ValueError):
        b200spec.eeg_to_spectrogram(torch.zeros(1, 20, 9999, device="cuda"))       # shorter than the crop
    with pytest.raises(ValueError):
        b200spec.eeg_to_spectrogram(torch.zeros(1, 10, 10000, device="cuda"))      # pair index out of range
    with pytest.raises(NotImplementedError):
        b200spec.eeg_to_spectrogram(torch.zeros(1, 20, 10000, device="cuda"), win_length=64)


def test_full_size_properties_config2():
    """BASELINE config 2: 1 024 records of 50 s -> (1024, 128, 256, 4)."""
    g = torch.Generator(device="cuda").manual_seed(2)
    x = torch.randn((1024, 20, 10000), generator=g, device="cuda")
    x = (x.cumsum(-1) * 0.3 + torch.randn((1024, 20, 10000), generator=g, device="cuda")).clamp_(-1024, 1024)
    out = b200spec.eeg_to_spectrogram(x)
    assert out.shape == (1024, 128, 256, 4) and not torch.isnan(out).any()
    assert float(out.max()) <= 1.0 + 1e-6 and float(out.min()) >= -1.0 - 1e-6      # dB in [-80, 0]
    sub = b200spec.eeg_to_spectrogram(x[300:302].contiguous())
    assert torch.equal(sub, out[300:302])                                          # records independent
    want = mo.eeg_to_spectrogram_oracle(x[911].cpu().numpy())
    assert np.abs(out[911].cpu().numpy() - want).max() <= ATOL
