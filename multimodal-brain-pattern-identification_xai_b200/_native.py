"""ctypes binding of ``lib/libeegspec.so`` (the C ABI in ``include/eegspec.h``).

There is no fallback of any kind: if the shared object is missing this module raises at
import, and every compute call raises when CUDA is unavailable.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG_DIR, "lib", "libeegspec.so")
MAX_PAIRS = 32

OK = 0
EINVAL, ENOVERLAP, EUNSUPPORTED, ESHORT, EWORKSPACE, EALIGN, ESTATICLIMIT = -1, -2, -3, -4, -5, -6, -7
CLS_SPEC_STATIC_BYTES = 8 << 20     # EEGSPEC_CLS_SPEC_STATIC_BYTES


class StftPlan(C.Structure):
    _fields_ = [
        ("hop", C.c_int32), ("half", C.c_int32), ("nadd", C.c_int32), ("n_freqs", C.c_int32),
        ("n_frames", C.c_int64), ("n_slots", C.c_int64), ("gather", C.c_int32),
        ("reserved", C.c_int32),
    ]


class MelParams(C.Structure):
    _fields_ = [
        ("sr", C.c_int32), ("n_fft", C.c_int32), ("win_length", C.c_int32), ("hop", C.c_int32),
        ("n_mels", C.c_int32), ("fmin", C.c_float), ("fmax", C.c_float), ("crop", C.c_int32),
        ("n_time_div", C.c_int32), ("amin", C.c_float), ("top_db", C.c_float),
        ("db_offset", C.c_float), ("db_scale", C.c_float), ("pad_mode", C.c_int32),
        ("n_chains", C.c_int32), ("pairs_per_chain", C.c_int32),
        ("pair_a", C.c_int32 * MAX_PAIRS), ("pair_b", C.c_int32 * MAX_PAIRS),
        ("time_major", C.c_int32), ("engine", C.c_int32),
    ]


MAX_XF_CH = 32


class XformParams(C.Structure):
    _fields_ = [
        ("n_out", C.c_int32), ("sel_a", C.c_int32 * MAX_XF_CH), ("sel_b", C.c_int32 * MAX_XF_CH),
        ("normalize", C.c_int32), ("clip_abs", C.c_float), ("scale", C.c_float), ("filter", C.c_int32),
        ("b", C.c_double * 5), ("a", C.c_double * 5), ("downsample", C.c_int32),
        ("out_channel_major", C.c_int32), ("mu_law", C.c_int32), ("mu", C.c_float),
    ]


class CondParams(C.Structure):
    _fields_ = [("n_channels", C.c_int32), ("n_freqs", C.c_int32), ("n_cols", C.c_int32),
                ("out_channels", C.c_int32), ("out_length", C.c_int32), ("mix_alpha", C.c_float)]


CLS_MAX_TAPS = 13


class ClsEegParams(C.Structure):
    _fields_ = [("n_taps1", C.c_int32), ("b1", C.c_double * CLS_MAX_TAPS), ("a1", C.c_double * CLS_MAX_TAPS),
                ("n_taps2", C.c_int32), ("b2", C.c_double * CLS_MAX_TAPS), ("a2", C.c_double * CLS_MAX_TAPS),
                ("n_pairs", C.c_int32), ("pair_a", C.c_int32 * MAX_PAIRS), ("pair_b", C.c_int32 * MAX_PAIRS),
                ("n_features", C.c_int32), ("fixed_length", C.c_int32), ("eps", C.c_float)]


class ClsSpecParams(C.Structure):
    _fields_ = [("rows", C.c_int32), ("cols", C.c_int32), ("notch_b", C.c_double * 3), ("notch_a", C.c_double * 3),
                ("notch_zi", C.c_double * 2), ("gauss_sigma", C.c_float), ("gauss_truncate", C.c_float),
                ("eps", C.c_float)]


# name -> (restype, argtypes); mirrors include/eegspec.h one to one.
SIGNATURES = {
    "eegspec_abi_version": (C.c_int, []),
    "eegspec_strerror": (C.c_char_p, [C.c_int]),
    "eegspec_stft_plan_make": (C.c_int, [C.c_int64, C.c_int, C.c_int, C.c_int64, C.POINTER(StftPlan)]),
    "eegspec_interp_bracket": (C.c_int64, [C.c_int64, C.c_int64, C.c_int64]),
    "eegspec_stft_image_workspace": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int64]),
    "eegspec_stft_image_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                         C.c_int64, C.c_float, C.c_void_p, C.c_void_p, C.c_size_t,
                                         C.c_void_p]),
    "eegspec_stft_logmag_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                          C.c_void_p, C.c_void_p]),
    "eegspec_mel_params_default": (None, [C.POINTER(MelParams)]),
    "eegspec_melspec_geometry": (C.c_int, [C.POINTER(MelParams), C.POINTER(C.c_int32),
                                           C.POINTER(C.c_int32)]),
    "eegspec_melspec_workspace": (C.c_size_t, [C.c_int64, C.POINTER(MelParams)]),
    "eegspec_melspec_workspace_default": (C.c_size_t, [C.c_int64, C.POINTER(MelParams)]),
    "eegspec_melspec_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64, C.c_int64,
                                      C.POINTER(MelParams), C.c_void_p, C.c_void_p, C.c_size_t,
                                      C.c_void_p]),
    "eegspec_melspec_sliding_f32": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.c_int64, C.c_int64,
                                              C.POINTER(MelParams), C.c_void_p, C.c_void_p, C.c_size_t,
                                              C.c_void_p]),
    "eegspec_stft_image_host_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_int,
                                              C.c_int64, C.c_float, C.c_void_p, C.c_int]),
    "eegspec_melspec_host_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int64,
                                           C.POINTER(MelParams), C.c_void_p, C.c_int]),
    "eegspec_release": (C.c_int, [C.c_int]),
    "eegspec_melspec_host_bytes": (C.c_int, [C.c_int64, C.c_int32, C.c_int64, C.POINTER(MelParams),
                                             C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "eegspec_xform_params_default": (None, [C.POINTER(XformParams)]),
    "eegspec_xform_out_length": (C.c_int64, [C.c_int64, C.POINTER(XformParams)]),
    "eegspec_eeg_transform_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int64,
                                            C.POINTER(XformParams), C.c_void_p, C.c_void_p]),
    "eegspec_eeg_transform_stft_workspace": (C.c_size_t, [C.c_int64, C.c_int64, C.POINTER(XformParams), C.c_int,
                                                          C.c_int, C.c_int64]),
    "eegspec_eeg_transform_stft_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int64,
                                                 C.POINTER(XformParams), C.c_int, C.c_int, C.c_int64, C.c_float,
                                                 C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]),
    "eegspec_spec_condition_f32": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(CondParams), C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p, C.c_void_p]),
    "eegspec_cls_eeg_params_default": (None, [C.POINTER(ClsEegParams)]),
    "eegspec_classifier_eeg_workspace": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int32]),
    "eegspec_classifier_eeg_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int64,
                                             C.POINTER(ClsEegParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t,
                                             C.c_void_p]),
    "eegspec_cls_spec_params_default": (None, [C.POINTER(ClsSpecParams)]),
    "eegspec_classifier_spec_workspace": (C.c_size_t, [C.c_int64, C.POINTER(ClsSpecParams)]),
    "eegspec_classifier_spec_f32": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(ClsSpecParams), C.c_void_p, C.c_void_p,
                                              C.c_void_p, C.c_size_t, C.c_void_p]),
    "eegspec_classifier_spec_static_f32": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(ClsSpecParams), C.c_void_p,
                                                     C.c_void_p, C.c_void_p]),
    "eegspec_stft_knots_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, C.c_int64,
                                         C.c_float, C.c_void_p, C.c_void_p]),
    "eegspec_spec_condition_knots_f32": (C.c_int, [C.c_void_p, C.c_int32, C.c_int64, C.POINTER(CondParams), C.c_void_p,
                                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "eegspec_launch_count": (C.c_int64, []),
}


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
            "`python __graft_entry__.py` (or b200spec.build.build()); there is no CPU fallback."
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)      # AttributeError if the .so lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    if lib.eegspec_abi_version() != 1:
        raise ImportError("libeegspec.so ABI version mismatch; rebuild it")
    return lib


lib = _load()


def strerror(status: int) -> str:
    return lib.eegspec_strerror(int(status)).decode()


def check(status: int, what: str) -> None:
    """Map a C status to the exception the reference's own stack would raise."""
    if status == OK:
        return
    msg = f"{what}: {strerror(status)} (status {status})"
    if status == ENOVERLAP:
        raise ValueError(strerror(status))            # SciPy's message, verbatim
    if status == EUNSUPPORTED:
        raise NotImplementedError(msg)
    if status < 0:
        raise ValueError(msg)
    raise RuntimeError(msg)
