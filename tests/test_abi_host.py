"""CPU: the C-ABI library loads without a GPU, exports every declared symbol, and its
host-side integer logic (framing plan, brackets, geometry, argument errors) is exact."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT
from oracle import melspec_oracle as mo
from oracle import stft_oracle as so

import b200spec
from b200spec import _native as nat


def declared_functions():
    with open(os.path.join(ROOT, "include", "eegspec.h")) as fh:
        text = re.sub(r"/\*.*?\*/", "", fh.read(), flags=re.S)
    return sorted(set(re.findall(r"\b(eegspec_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    names = declared_functions()
    assert len(names) >= 14
    lib = C.CDLL(nat.LIB_PATH)
    for name in names:
        assert hasattr(lib, name), name
    assert sorted(nat.SIGNATURES) == names
    assert nat.lib.eegspec_abi_version() == 1


def test_struct_layout_matches_header():
    assert C.sizeof(nat.StftPlan) == 40
    assert C.sizeof(nat.MelParams) == 16 * 4 + 2 * 32 * 4 + 2 * 4
    # eegspec_xform_params: int n_out, 2 x int[32], int, 2 floats, int (+4 pad), 2 x double[5], 3 ints, float
    assert C.sizeof(nat.XformParams) == 4 + 2 * 32 * 4 + 4 + 8 + 4 + 4 + 2 * 5 * 8 + 3 * 4 + 4
    assert C.sizeof(nat.CondParams) == 6 * 4
    # eegspec_cls_eeg_params: (int + pad, 2 x double[13]) twice, int, 2 x int[32], 2 ints, float (+ tail pad to 8)
    assert C.sizeof(nat.ClsEegParams) == 2 * (8 + 2 * 13 * 8) + 4 + 2 * 32 * 4 + 2 * 4 + 4
    assert C.sizeof(nat.ClsSpecParams) == 2 * 4 + 8 * 8 + 3 * 4 + 4


def test_struct_sizes_match_the_compiled_header(tmp_path):
    """sizeof of every parameter struct as gcc sees include/eegspec.h == the ctypes mirror."""
    import subprocess

    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include "eegspec.h"\nint main(void){printf("%zu %zu %zu %zu %zu %zu\\n",'
                   'sizeof(eegspec_stft_plan),sizeof(eegspec_mel_params),sizeof(eegspec_xform_params),'
                   'sizeof(eegspec_cond_params),sizeof(eegspec_cls_eeg_params),sizeof(eegspec_cls_spec_params));return 0;}\n')
    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(v) for v in subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()]
    want = [C.sizeof(t) for t in (nat.StftPlan, nat.MelParams, nat.XformParams, nat.CondParams, nat.ClsEegParams,
                                  nat.ClsSpecParams)]
    assert got == want


@pytest.mark.parametrize("length", [64, 100, 2000, 2001, 10000, 70000, 200000, 720000])
@pytest.mark.parametrize("nperseg,noverlap", [(64, 32), (128, 64), (64, 48), (256, 128), (16, 3), (1024, 512)])
def test_plan_is_bit_exact(length, nperseg, noverlap):
    if length < nperseg:
        with pytest.raises(ValueError):
            b200spec.stft_plan(length, nperseg, noverlap)
        return
    want = so.frame_plan(length, nperseg, noverlap)
    got = b200spec.stft_plan(length, nperseg, noverlap)
    for key in ("hop", "half", "nadd", "n_frames", "n_freqs"):
        assert got[key] == want[key], key
    assert got["n_slots"] == (4000 if got["gather"] else want["n_frames"])


@pytest.mark.parametrize("n_frames", [2, 64, 314, 1999, 2189, 22501])
def test_brackets_are_bit_exact(n_frames):
    want = so.interp_brackets(n_frames)
    got = np.array([b200spec.interp_bracket(j, n_frames) for j in range(2000)])
    np.testing.assert_array_equal(got, want)


def test_status_codes_and_messages():
    plan = nat.StftPlan()
    assert nat.lib.eegspec_stft_plan_make(2000, 64, 64, 2000, C.byref(plan)) == nat.ENOVERLAP
    assert nat.strerror(nat.ENOVERLAP) == "noverlap must be less than nperseg."   # SciPy's words
    with pytest.raises(ValueError, match="noverlap must be less than nperseg"):
        b200spec.stft_plan(2000, 64, 64)
    assert nat.lib.eegspec_stft_plan_make(10, 64, 32, 2000, C.byref(plan)) == nat.ESHORT
    assert nat.lib.eegspec_stft_plan_make(2000, 64, 32, 2000, None) == nat.EINVAL
    # null pointers are rejected before anything touches a device
    assert nat.lib.eegspec_stft_image_f32(None, 1, 2000, 2000, 64, 32, 2000, 1e-8, None, None, 0, None) == nat.EINVAL
    assert nat.lib.eegspec_stft_logmag_f32(None, 1, 2000, 2000, 64, 32, None, None) == nat.EINVAL
    assert b"unknown" not in nat.lib.eegspec_strerror(nat.EUNSUPPORTED)


def test_mel_defaults_and_geometry():
    p = b200spec.mel_params()
    assert (p.sr, p.n_fft, p.win_length, p.n_mels, p.crop, p.n_chains, p.pairs_per_chain) == \
        (200, 1024, 128, 128, 10000, 4, 4)
    pairs = [(p.pair_a[i], p.pair_b[i]) for i in range(16)]
    assert pairs == [tuple(r) for r in mo.pair_indices(16)]
    assert pairs == b200spec.pair_indices()
    assert b200spec.melspec_geometry() == (257, 256)
    assert nat.lib.eegspec_melspec_workspace(1, C.byref(p)) >= 16 * 128 * 256 * 4
    with pytest.raises(NotImplementedError):
        b200spec.melspec_geometry(b200spec.mel_params(win_length=256))
    with pytest.raises(ValueError):
        b200spec.melspec_geometry(b200spec.mel_params(fmax=500.0))


def test_host_door_byte_count():
    """The mel host door copies in only the electrode rows some montage pair reads: 16 of the 20
    EEG_COLUMNS rows (Fz, Cz, Pz and EKG stay home); a time-major record travels whole."""
    p = b200spec.mel_params()
    used = {p.pair_a[i] for i in range(16)} | {p.pair_b[i] for i in range(16)}
    assert len(used) == 16 and not used & {8, 9, 10, 19}
    h2d, d2h = C.c_int64(), C.c_int64()
    assert nat.lib.eegspec_melspec_host_bytes(7, 20, 10000, C.byref(p), C.byref(h2d), C.byref(d2h)) == 0
    assert (h2d.value, d2h.value) == (7 * 16 * 10000 * 4, 7 * 128 * 256 * 4 * 4)
    tm = b200spec.mel_params(time_major=True)
    assert nat.lib.eegspec_melspec_host_bytes(7, 20, 10000, C.byref(tm), C.byref(h2d), None) == 0
    assert h2d.value == 7 * 20 * 10000 * 4
    assert nat.lib.eegspec_melspec_host_bytes(7, 0, 10000, C.byref(p), C.byref(h2d), None) == nat.EINVAL


def test_workspace_covers_the_time_major_gather():
    """Time-major (parquet-layout) records are gathered into electrode rows inside the caller's
    workspace: 16 rows x crop floats per record for up to 1 024 records per pass."""
    cm, tm = b200spec.mel_params(), b200spec.mel_params(time_major=True)
    for batch in (1, 64, 1024, 5000):
        need_cm = nat.lib.eegspec_melspec_workspace(batch, C.byref(cm))
        need_tm = nat.lib.eegspec_melspec_workspace(batch, C.byref(tm))
        assert need_tm >= need_cm > 0
        assert need_tm >= min(batch, 1024) * 16 * 10000 * 4


def test_montage_constants_match_oracle():
    from b200spec import montage

    assert list(montage.EEG_COLUMNS) == mo.EEG_COLUMNS
    assert list(montage.MAP_FEATURES) == mo.MAP_FEATURES
    assert montage.crop_offset(18000) == mo.crop_offset(18000) == 4000


def test_compute_paths_refuse_to_run_without_cuda():
    import torch

    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from conftest import Cfg

    with pytest.raises(RuntimeError, match="no CPU path"):
        b200spec.compute_stft(torch.zeros(1, 1, 2000), Cfg())
    with pytest.raises(RuntimeError, match="no CPU path"):
        b200spec.eeg_to_spectrogram(torch.zeros(1, 20, 10000))
    with pytest.raises(ValueError):
        b200spec.compute_stft(torch.zeros(2000), Cfg())           # wrong rank, like the reference
    with pytest.raises(NotImplementedError):
        b200spec.compute_stft(torch.zeros(1, 1, 2000), Cfg(window="hamming"))


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "multimodal-brain-pattern-identification_xai_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, f)) as fh:
                    src = fh.read()
                assert "import oracle" not in src and "from oracle" not in src, f
