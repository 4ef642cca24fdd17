"""Generate the golden vectors under tests/golden/ (run in the BUILD container only).

    python tests/golden/make_golden.py

* ``stft_golden.npz``   -- inputs and outputs of the REFERENCE's own ``compute_stft`` /
  ``compute_stft_single`` (``/root/reference/root/src/utils/DiffEEG_utils.py:27-76, 504-509``),
  executed here through ``oracle/ref_extract.py`` (AST lift, nothing copied).  These pin the
  mode-R oracle and the CUDA path.
* ``melspec_golden.npz`` -- mode N has no reference implementation (parity UNPINNED, see
  ``oracle/melspec_oracle.py``); the vectors come from the oracle restatement and guard the
  CUDA path and the oracle itself against regressions.  The montage facts inside are read
  from the reference's ``config.yml``.

* ``eegtransform_golden.npz`` -- inputs and outputs of the REFERENCE's own
  ``_EEGTransformer.transform`` (``dataset.py:73-104``), same mechanism; pins row (f)1.

Versions used to generate the committed files are stored inside them.
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import melspec_oracle as mo  # noqa: E402
from oracle import ref_extract as rx  # noqa: E402


def eeg_like(rng, shape, walk=0.3, noise=1.0):
    x = np.cumsum(rng.standard_normal(shape), axis=-1) * walk
    t = np.arange(shape[-1]) / 200.0
    x = x + 10.0 * np.sin(2 * np.pi * 10.0 * t) + noise * rng.standard_normal(shape)
    return np.clip(x, -1024, 1024).astype(np.float32)


STFT_CASES = {
    # name: (B, C, L, nperseg, noverlap, kind)
    "diffeeg_64_32": (2, 2, 2000, 64, 32, "eeg"),
    "hms50s_64_32": (1, 2, 10000, 64, 32, "eeg"),
    "nfft128": (1, 1, 2000, 128, 64, "eeg"),
    "hop16": (1, 1, 2000, 64, 48, "randn"),
    "one_window": (1, 1, 64, 64, 32, "randn"),
    "nfft16": (1, 2, 2000, 16, 8, "randn"),
    "nfft256_odd_len": (1, 1, 2001, 256, 100, "eeg"),
    "flat_rows": (1, 3, 2000, 64, 32, "flat"),
    "nan_sample": (1, 2, 2000, 64, 32, "nan"),
    "decimating": (1, 1, 70000, 64, 32, "eeg"),
    "gather": (1, 1, 200000, 64, 32, "eeg"),
}


def make_stft():
    import scipy
    import torch

    ref = rx.load_reference_compute_stft()
    ref_single = rx.load_reference_compute_stft_single()
    rng = np.random.default_rng(20261104)
    out = {}
    for name, (b, c, length, nperseg, noverlap, kind) in STFT_CASES.items():
        if kind == "randn":
            x = rng.standard_normal((b, c, length)).astype(np.float32)
        else:
            x = eeg_like(rng, (b, c, length))
        if kind == "flat":
            x[0, 0] = 0.0            # all-zero channel
            x[0, 1] = 37.5           # constant channel
        if kind == "nan":
            x[0, 0, 777] = np.nan    # poisons every row of channel 0, leaves channel 1 alone
        cfg = rx.make_config(nperseg, noverlap)
        y = ref(torch.from_numpy(x), cfg).numpy()
        assert y.dtype == np.float32
        out[f"{name}/x"] = x
        out[f"{name}/y"] = y
        out[f"{name}/cfg"] = np.array([nperseg, noverlap], dtype=np.int64)
    x = eeg_like(rng, (2000,))
    out["single/x"] = x
    out["single/y"] = ref_single(torch.from_numpy(x), rx.make_config(64, 32)).astype(np.float32)
    out["versions"] = np.array([f"scipy {scipy.__version__}", f"numpy {np.__version__}",
                                f"torch {torch.__version__}"])
    np.savez_compressed(os.path.join(HERE, "stft_golden.npz"), **out)
    print("stft_golden.npz:", {k: v.shape for k, v in out.items() if k.endswith("/y")})


def make_melspec():
    import yaml

    rng = np.random.default_rng(20261105)
    out = {}
    x = eeg_like(rng, (1, 20, 10000))
    x[:, 19] = (np.arange(10000) % 200 < 3) * 500.0            # EKG-like spike train
    out["clean/x"] = x
    out["clean/y"] = mo.eeg_to_spectrogram_batch_oracle(x)
    out["clean/y_reflect"] = mo.eeg_to_spectrogram_batch_oracle(x, pad_mode="reflect")

    xn = eeg_like(rng, (1, 20, 10400))                           # longer than the crop window
    holes = rng.random((10400,)) < 0.005
    xn[0, 4, holes] = np.nan                                     # F7: touches LL pairs 0 and 1
    xn[0, 13, 100:900] = np.nan                                  # C4: a run of NaNs
    xn[0, 17] = np.nan                                           # T6 entirely missing
    out["nan/x"] = xn
    out["nan/y"] = mo.eeg_to_spectrogram_batch_oracle(xn)

    xq = eeg_like(rng, (1, 20, 10000)) * np.float32(1e-3)
    xq[0, 0, 5000:5040] += 900.0                                 # electrode pop: > 80 dB range
    xq[0, 11] = 0.0
    xq[0, 12] = 0.0                                              # Fp2-F4 identically zero
    out["range/x"] = xq
    out["range/y"] = mo.eeg_to_spectrogram_batch_oracle(xq)

    with open(os.path.join(rx.REFERENCE_ROOT, "root", "config", "config.yml")) as fh:
        cfg = yaml.safe_load(fh)
    out["ref_eeg_columns"] = np.array(cfg["EEG_COLUMNS"])
    out["ref_map_features"] = np.array(cfg["map_features"])
    out["versions"] = np.array([f"numpy {np.__version__}"])
    np.savez_compressed(os.path.join(HERE, "melspec_golden.npz"), **out)
    print("melspec_golden.npz:", {k: v.shape for k, v in out.items() if "/y" in k})


XFORM_CASES = {
    # name: (L, C, transformer kwargs, kind)
    "default19": (10000, 20, dict(n_feats=19, downsample=5), "eeg"),
    "magic8": (10000, 8, dict(n_feats=8, apply_chris_magic_ch8=True, downsample=5), "eeg"),
    "nan_clip": (4000, 20, dict(n_feats=19, downsample=5), "nan"),
    "no_downsample": (2000, 20, dict(n_feats=19), "eeg"),
    "no_filter_ds4": (3001, 20, dict(n_feats=19, apply_butter_lowpass_filter=False, downsample=4), "eeg"),
    "raw_scale": (2403, 20, dict(n_feats=19, normalize=False, downsample=3), "eeg"),
    # appended later: the draws above are unchanged
    "mu_law": (10000, 20, dict(n_feats=19, apply_mu_law_encoding=True, downsample=5), "eeg"),
    "mu_law_nofilter_nan": (3000, 20, dict(n_feats=19, apply_butter_lowpass_filter=False, apply_mu_law_encoding=True, downsample=2), "nan"),
    # the reference's real input: the (10000, 19) window of _get_eeg_window (cfg_utils.py:366-373), no EKG column
    "window19": (10000, 19, dict(n_feats=19, downsample=5), "nan"),
    "window19_ds4": (3000, 19, dict(n_feats=19, downsample=4), "eeg"),
    "window19_odd_length": (2001, 19, dict(n_feats=19, downsample=5), "eeg"),
}


def make_eegtransform():
    """Inputs and outputs of the REFERENCE's own ``_EEGTransformer.transform``
    (/root/reference/root/src/data/dataset.py:73-104) through ``oracle/ref_extract.py``."""
    import scipy

    ref_cls = rx.load_reference_eeg_transformer()
    rng = np.random.default_rng(20261106)
    out = {}
    for name, (length, n_ch, kw, kind) in XFORM_CASES.items():
        x = (np.cumsum(rng.standard_normal((length, n_ch)), axis=0) * 3.0
             + 40.0 * rng.standard_normal((length, n_ch))).astype(np.float32)
        if kind == "nan":
            x[100:140, 3] = np.nan
            x[rng.random((length,)) < 0.01, 7] = np.nan
            x[:, 11] = np.nan                                   # a dead electrode
            x[2000:2050, 0] = 5000.0                            # beyond the clip level
            x[2100:2150, 1] = -np.inf
        out[f"{name}/x"] = x
        out[f"{name}/y"] = np.asarray(ref_cls(**kw).transform(x), dtype=np.float64)
    out["versions"] = np.array([f"numpy {np.__version__}", f"scipy {scipy.__version__}"])
    path = os.path.join(HERE, "eegtransform_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


def make_fronthalf():
    """``fronthalf_golden.npz``: inputs and outputs of the REFERENCE's own ``_get_eeg_window``
    (/root/reference/root/src/utils/cfg_utils.py:354-389: parquet -> middle crop -> per-channel NaN
    fill) and ``calculate_differential_signals`` (/root/reference/root/src/utils/data_utils.py:183-215),
    lifted by ``oracle/ref_extract.py`` and run on temporary parquet files.  These pin
    ``crop_offset`` / ``fill_missing`` / ``montage_differences`` of the mode-N oracle."""
    import tempfile

    import pandas as pd
    import pyarrow

    rng = np.random.default_rng(20261107)
    out = {}
    for name, n_pts in (("odd", 10001), ("longodd", 13579)):
        x = np.round(eeg_like(rng, (n_pts, 20)))                # integer microvolts: the fixture compresses
        x[rng.random((n_pts,)) < 0.01, 2] = np.nan              # scattered holes
        x[n_pts // 2 - 40:n_pts // 2 + 40, 5] = np.nan          # a run inside the crop window
        x[:300, 9] = np.nan                                     # a run that the crop may cut off
        x[:, 14] = np.nan                                       # a dead electrode
        out[f"window/{name}/x"] = x
        with tempfile.TemporaryDirectory() as tmp:
            path = os.path.join(tmp, "eeg.parquet")
            pd.DataFrame(x, columns=mo.EEG_COLUMNS).to_parquet(path)
            for magic in (False, True):
                fn, cfg = rx.load_reference_get_eeg_window(use_magic_8=magic)
                with np.errstate(all="ignore"):
                    import warnings
                    with warnings.catch_warnings():
                        warnings.simplefilter("ignore")
                        win = fn(path)
                out[f"window/{name}/{'magic8' if magic else 'full19'}"] = win
                out[f"channels/{'magic8' if magic else 'full19'}"] = np.array(cfg.feats if magic else cfg.channel_feats)
    x = np.round(eeg_like(rng, (20, 600))).astype(np.float64)
    x[3, 100:110] = np.nan
    out["diff/x"] = x
    env = rx.load_reference_data_utils(["calculate_differential_signals"], feature_names=mo.EEG_COLUMNS)
    out["diff/eeg_columns"] = env["calculate_differential_signals"](x)              # (20 + 18, 600)
    env = rx.load_reference_data_utils(["calculate_differential_signals"])            # the YAML's own 19 eeg_features
    out["diff/yaml_features"] = env["calculate_differential_signals"](x[:19])         # (19 + 18, 600)
    out["diff/yaml_feature_names"] = np.array(rx.reference_yaml_config()["eeg_features"])
    out["versions"] = np.array([f"numpy {np.__version__}", f"pandas {pd.__version__}", f"pyarrow {pyarrow.__version__}"])
    path = os.path.join(HERE, "fronthalf_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


def xcheck_image(x):
    """(20, L) EEG -> (128, 256, 4) with LIBRARY code only for the back half: torchaudio's
    MelSpectrogram (torch.stft + melscale_fbanks) and amplitude_to_DB.  Nothing of oracle/ is used
    except the montage table; the front half is three NumPy lines (difference, nanmean fill)."""
    import torch
    import torchaudio

    mel = torchaudio.transforms.MelSpectrogram(
        sample_rate=200, n_fft=1024, win_length=128, hop_length=39, f_min=0.0, f_max=20.0, n_mels=128,
        window_fn=torch.hann_window, power=2.0, center=True, pad_mode="constant", norm="slaney",
        mel_scale="slaney").double()
    off = (x.shape[1] - 10000) // 2
    w = x[:, off:off + 10000]
    img = np.zeros((128, 256, 4), np.float32)
    pairs = mo.pair_indices(16)
    for k in range(4):
        for kk in range(4):
            a, b = pairs[4 * k + kk]
            y = w[a] - w[b]
            if np.isnan(y).all():
                y = np.zeros_like(y)
            elif np.isnan(y).any():
                y = np.where(np.isnan(y), np.nanmean(y), y)
            m = mel(torch.from_numpy(y.astype(np.float32)).double()).float()         # (128, 257)
            ref = float(m.max())
            db = torchaudio.functional.amplitude_to_DB(m[None], 10.0, 1e-10, float(np.log10(max(1e-10, ref))), 80.0)[0]
            img[:, :, k] += ((db[:, :256] + 40.0) / 40.0).numpy()
        img[:, :, k] /= 4.0
    return img


def make_melspec_xcheck():
    """``melspec_xcheck.npz``: a SECOND, independent mode-N image from library code (torchaudio), on the
    inputs of ``melspec_golden.npz``.  It narrows -- it does not close -- "parity unpinned": torchaudio
    is not librosa, but its Slaney bank, centred constant-pad STFT and dB conversion are independent of
    everything written in this repository."""
    import torch
    import torchaudio

    gold = np.load(os.path.join(HERE, "melspec_golden.npz"))
    out = {}
    for case in ("clean", "nan", "range"):
        out[f"{case}/y"] = xcheck_image(gold[f"{case}/x"][0])[None]
    out["versions"] = np.array([f"torch {torch.__version__}", f"torchaudio {torchaudio.__version__}", f"numpy {np.__version__}"])
    path = os.path.join(HERE, "melspec_xcheck.npz")
    np.savez_compressed(path, **out)
    for case in ("clean", "nan", "range"):
        print(case, "xcheck vs oracle golden: max abs diff", float(np.abs(out[f"{case}/y"] - gold[f"{case}/y"]).max()))
    print("wrote", path, os.path.getsize(path), "bytes")


def diffeeg_inputs(case: str):
    """Deterministic inputs of the row-(f)2 cases (regenerated by the tests: the images are too big to commit)."""
    cfg = {"ship19": (3, 19, 32, 7), "ckpt8": (2, 8, 128, 8), "train19": (6, 19, 32, 9)}[case]
    b, c, hidden, seed = cfg
    rng = np.random.default_rng(seed)
    spec = rng.random((b, c, 33, 2000), dtype=np.float32)          # compute_stft images live in [0, 1]
    spec[0, 0] = 0.0                                               # a flat channel
    return spec, c, hidden, seed


def make_diffeeg():
    """``diffeeg_golden.npz``: the spectrogram condition the REFERENCE's own ``DiffEEG`` computes
    (/root/reference/root/src/models/diffEEG.py:47-53, 108-110, 136-145), lifted by ``oracle/ref_extract.py``:
    ``<case>/hooked`` = the input of ``channel_expand`` during the reference's ``forward`` (bf16 autocast on CPU),
    ``<case>/fp32`` = its own ``spectrogram_upsample1`` layer in fp32 + the three forward lines."""
    import torch
    import torch.nn.functional as F

    out = {}
    for case in ("ship19", "ckpt8", "train19"):
        spec_np, c, hidden, seed = diffeeg_inputs(case)
        torch.manual_seed(seed)
        m = rx.load_reference_diffeeg(n_channels=c, hidden_channels=hidden)
        spec = torch.from_numpy(spec_np)
        b = spec.shape[0]
        w, bias = m.spectrogram_upsample1.weight.detach(), m.spectrogram_upsample1.bias.detach()
        out[f"{case}/weight"], out[f"{case}/bias"] = w.numpy(), bias.numpy()
        labels = torch.tensor([0, 1, 0, 0, 1, 2][:b])
        onehot = F.one_hot(labels, 6).float()
        cap = {}
        handle = m.channel_expand.register_forward_pre_hook(lambda mod, inp: cap.__setitem__("x", inp[0].detach().float().clone()))
        if case == "train19":
            m.train()
            torch.manual_seed(1234)
            mixed = m.recombine_spectrograms(spec, labels)          # the reference's own mixing
            torch.manual_seed(1234)                                 # replay its randperm calls to learn the partners
            partner = torch.arange(b)
            for cls in range(m.n_classes):
                idx = (labels == cls).nonzero(as_tuple=True)[0]
                if len(idx) > 1:
                    partner[idx] = idx[torch.randperm(len(idx))]
            assert torch.equal(mixed, 0.5 * spec + 0.5 * spec[partner])
            out[f"{case}/partner"] = partner.numpy().astype(np.int32)
            spec_in = mixed
            m.eval()
        else:
            spec_in = spec
        import warnings
        with torch.no_grad(), warnings.catch_warnings():
            warnings.simplefilter("ignore")
            m(torch.zeros(b, c, 2000), onehot, torch.arange(b, dtype=torch.float32) + 3, spec_in)
            own = F.interpolate(F.relu(m.spectrogram_upsample1(spec_in)).flatten(start_dim=2), size=2000, mode="linear",
                                align_corners=False)
        handle.remove()
        out[f"{case}/hooked"] = cap["x"].numpy()
        out[f"{case}/fp32"] = own.numpy()
    out["versions"] = np.array([f"torch {torch.__version__}", f"numpy {np.__version__}"])
    path = os.path.join(HERE, "diffeeg_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


CLASSIFIER_EEG_CASES = {"clean": (10000, "clean"), "nan": (10400, "nan"), "long": (13000, "clean"), "odd": (10001, "nan")}


def classifier_eeg_input(gold, name):
    x = gold[f"eeg/{name}/x_i16"].astype(np.float32)
    x[gold[f"eeg/{name}/x_i16"] == -32768] = np.nan
    return x
CLASSIFIER_SPEC_CASES = {"offset40": (320, 40), "no_offset_short": (250, None), "offset_near_end": (310, 250)}


def make_classifier():
    """``classifier_golden.npz``: inputs and outputs of the REFERENCE's own ``CombinedDataset.process_eeg`` and
    ``process_spectrogram`` (/root/reference/root/jupyter_notebooks/XAI_Multimodality.py:1148-1204), lifted from
    the notebook export by ``oracle/ref_extract.py``; pins rows (f)3 and (f)4."""
    import pandas as pd
    import scipy

    rng = np.random.default_rng(20261108)
    out = {}
    for name, (n_pts, kind) in CLASSIFIER_EEG_CASES.items():
        x = np.round(eeg_like(rng, (20, n_pts), walk=0.5, noise=6.0)).T.copy()       # (n, 20) integer microvolts
        if kind == "nan":
            x[n_pts // 2:n_pts // 2 + 25, 3] = np.nan          # everything after it is NaN after the first filter
            x[rng.random(n_pts) < 0.002, 7] = np.nan
            x[n_pts - 1, 12] = np.nan                          # the very last sample
        ds = rx.load_reference_combined_dataset({1: pd.DataFrame(x, columns=mo.EEG_COLUMNS)}, {})
        out[f"eeg/{name}/x_i16"] = np.where(np.isnan(x), -32768, x).astype(np.int16)   # NaN -> -32768 (see classifier_eeg_input)
        out[f"eeg/{name}/y"] = ds.process_eeg({"eeg_id": 1}).numpy()
    for name, (n_rows, offset) in CLASSIFIER_SPEC_CASES.items():
        raw = rng.gamma(2.0, 8.0, size=(n_rows, 401)).astype(np.float16).astype(np.float32)   # positive power, fp16-exact
        raw[:, 0] = np.arange(n_rows) * 2 + 1                                             # the 'time' column
        raw[rng.random((n_rows, 401)) < 0.003] = np.nan
        raw[40:60, 200] = np.nan
        ds = rx.load_reference_combined_dataset({}, {2: pd.DataFrame(raw)})
        row = {"spectrogram_id": 2}
        if offset is not None:
            row["spectrogram_label_offset_seconds"] = offset
        y = ds.process_spectrogram(pd.Series(row) if offset is not None else row).numpy()
        assert y.shape == (3, 400, 300) and np.array_equal(y[0], y[1]) and np.array_equal(y[0], y[2])
        out[f"spec/{name}/raw_f16"] = raw.astype(np.float16)
        out[f"spec/{name}/offset"] = np.array(-1 if offset is None else offset)
        out[f"spec/{name}/y_plane"] = y[0]                      # the three planes are copies (np.tile, :1190)
    out["versions"] = np.array([f"numpy {np.__version__}", f"scipy {scipy.__version__}", f"pandas {pd.__version__}"])
    path = os.path.join(HERE, "classifier_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    if not rx.reference_available():
        raise SystemExit("/root/reference is not mounted: golden vectors can only be made in the build container")
    only = set(sys.argv[1:])
    for fn in (make_stft, make_melspec, make_eegtransform, make_fronthalf, make_melspec_xcheck, make_diffeeg, make_classifier):
        if not only or fn.__name__ in only:
            fn()
