"""Host-side mirror of the reference's EEG -> spectrogram interface.

``compute_stft(signal, config)`` keeps the exact signature, argument meaning, output layout
and error behaviour of ``/root/reference/root/src/utils/DiffEEG_utils.py:27-76``; the body is
one asynchronous launch of the fused sm_100a kernel through the C ABI (``include/eegspec.h``).
PyTorch only supplies device memory and the current CUDA stream.
"""
from __future__ import annotations

import ctypes as C
import sys
from typing import Optional, Sequence

import torch

from . import _native as nat
from . import montage as _montage

OUT_COLS = 2000   # literal in the reference (DiffEEG_utils.py:57-58)
EPS = 1e-8        # literal in the reference (DiffEEG_utils.py:74)


def _require_cuda() -> None:
    if not torch.cuda.is_available():
        raise RuntimeError("eegspec has no CPU path: a CUDA device (B200, sm_100a) is required")


def _stream_ptr(device: torch.device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t: Optional[torch.Tensor]) -> C.c_void_p:
    return C.c_void_p(t.data_ptr() if t is not None else 0)


def _stft_keys(config) -> tuple[int, int]:
    tr = config.diffEEG_trainer                          # ref :49-51 reads exactly these keys
    nperseg, noverlap, window = tr["stft_n_fft"], tr["stft_hop_length"], tr["stft_window"]
    if window != "hann":
        raise NotImplementedError(f"stft_window={window!r}: only 'hann' is implemented (no fallback)")
    return int(nperseg), int(noverlap)                   # NB: hop_length IS SciPy's noverlap (ref :50)


def stft_plan(length: int, nperseg: int, noverlap: int, out_cols: int = OUT_COLS) -> dict:
    """Frame count / padding of ``scipy.signal.stft`` as the reference calls it (integers)."""
    plan = nat.StftPlan()
    nat.check(nat.lib.eegspec_stft_plan_make(length, nperseg, noverlap, out_cols, C.byref(plan)),
              "stft_plan")
    return {f: getattr(plan, f) for f, _ in nat.StftPlan._fields_ if f != "reserved"}


def interp_bracket(j: int, n_frames: int, out_cols: int = OUT_COLS) -> int:
    return int(nat.lib.eegspec_interp_bracket(j, n_frames, out_cols))


def _as_f32(signal: torch.Tensor) -> torch.Tensor:
    if not isinstance(signal, torch.Tensor):
        raise TypeError("signal must be a torch.Tensor")
    return signal if signal.dtype == torch.float32 else signal.float()


def stft_image(signal: torch.Tensor, nperseg: int = 64, noverlap: int = 32,
               out_cols: int = OUT_COLS, eps: float = EPS) -> torch.Tensor:
    """(B, C, L) -> (B, C, nperseg//2+1, out_cols) fp32 on ``signal.device``."""
    batch_size, n_channels, length = signal.shape        # ValueError on wrong rank, like ref :37
    _require_cuda()
    x = _as_f32(signal)
    rows = batch_size * n_channels
    n_freqs = nperseg // 2 + 1
    if rows == 0:
        return torch.empty((batch_size, n_channels, n_freqs, out_cols), dtype=torch.float32,
                           device=signal.device)
    if not x.is_cuda:
        # Host tensors (what a DataLoader hands the trainer): stream through the device.
        x = x.contiguous()
        out = torch.empty((batch_size, n_channels, n_freqs, out_cols), dtype=torch.float32,
                          pin_memory=True)
        nat.check(nat.lib.eegspec_stft_image_host_f32(_ptr(x), rows, length, nperseg, noverlap,
                                                      out_cols, eps, _ptr(out),
                                                      torch.cuda.current_device()),
                  "compute_stft")
        return out
    x = x.contiguous()
    with torch.cuda.device(x.device):
        out = torch.empty((batch_size, n_channels, n_freqs, out_cols), dtype=torch.float32,
                          device=x.device)
        ws_bytes = nat.lib.eegspec_stft_image_workspace(rows, length, nperseg, noverlap, out_cols)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device) if ws_bytes else None
        nat.check(nat.lib.eegspec_stft_image_f32(_ptr(x), rows, length, length, nperseg, noverlap,
                                                 out_cols, eps, _ptr(out), _ptr(ws), ws_bytes,
                                                 _stream_ptr(x.device)),
                  "compute_stft")
    return out


def compute_stft(signal, config):
    """Drop-in for the reference's ``compute_stft`` (DiffEEG_utils.py:27-76).

    Args:
        signal (torch.Tensor): EEG of shape (batch_size, n_channels, length)
        config: any object with ``config.diffEEG_trainer[...]`` holding ``stft_n_fft``,
            ``stft_hop_length`` (used as SciPy's *noverlap*, exactly like the reference) and
            ``stft_window`` ('hann').

    Returns:
        torch.Tensor: (batch_size, n_channels, stft_n_fft//2+1, 2000) fp32 on ``signal.device``,
        every (record, channel, frequency) row min-max normalised to [0, 1].
    """
    nperseg, noverlap = _stft_keys(config)
    return stft_image(signal, nperseg, noverlap, OUT_COLS, EPS)


def compute_stft_single(signal, config):
    """The plotting twin (DiffEEG_utils.py:504-509): ``log1p|STFT|`` of one 1-D series,
    (F, n_frames) fp32; a leading batch of rows is accepted as well -> (R, F, n_frames)."""
    nperseg, noverlap = _stft_keys(config)
    _require_cuda()
    x = _as_f32(signal)
    squeeze = x.dim() == 1
    if x.dim() not in (1, 2):
        raise ValueError("compute_stft_single expects a 1-D series (or a 2-D stack of rows)")
    dev_in = x.device
    x = x.reshape(-1, x.shape[-1]).contiguous()
    if not x.is_cuda:
        x = x.cuda()
    rows, length = x.shape
    plan = stft_plan(length, nperseg, noverlap)
    with torch.cuda.device(x.device):
        out = torch.empty((rows, plan["n_freqs"], plan["n_frames"]), dtype=torch.float32,
                          device=x.device)
        nat.check(nat.lib.eegspec_stft_logmag_f32(_ptr(x), rows, length, length, nperseg, noverlap,
                                                  _ptr(out), _stream_ptr(x.device)),
                  "compute_stft_single")
    out = out[0] if squeeze else out
    return out if dev_in.type == "cuda" else out.cpu()


# --------------------------------------------------------------------------------- mode N
def mel_params(pad_mode: str = "constant", time_major: bool = False,
               pairs: Optional[Sequence[tuple[int, int]]] = None, pairs_per_chain: int = 4,
               **overrides) -> nat.MelParams:
    """Defaults = the north_star recipe; every field of ``eegspec_mel_params`` is overridable."""
    p = nat.MelParams()
    nat.lib.eegspec_mel_params_default(C.byref(p))
    if pad_mode not in ("constant", "reflect"):
        raise NotImplementedError(f"pad_mode={pad_mode!r}")
    p.pad_mode = 0 if pad_mode == "constant" else 1
    p.time_major = int(bool(time_major))
    if pairs is not None:
        pairs = list(pairs)
        if len(pairs) % pairs_per_chain or len(pairs) > nat.MAX_PAIRS or not pairs:
            raise ValueError("pairs must fill whole chains and number at most 32")
        p.n_chains, p.pairs_per_chain = len(pairs) // pairs_per_chain, pairs_per_chain
        for i, (a, b) in enumerate(pairs):
            p.pair_a[i], p.pair_b[i] = int(a), int(b)
    for key, val in overrides.items():
        if not hasattr(p, key):
            raise TypeError(f"unknown mel parameter {key!r}")
        setattr(p, key, val)
    return p


def melspec_geometry(params: Optional[nat.MelParams] = None) -> tuple[int, int]:
    params = params if params is not None else mel_params()
    n_frames, n_time = C.c_int32(), C.c_int32()
    nat.check(nat.lib.eegspec_melspec_geometry(C.byref(params), C.byref(n_frames), C.byref(n_time)),
              "melspec_geometry")
    return n_frames.value, n_time.value


def eeg_to_spectrogram(eeg: torch.Tensor, params: Optional[nat.MelParams] = None,
                       out: Optional[torch.Tensor] = None, **kwargs) -> torch.Tensor:
    """Raw EEG -> the 4-chain log-mel image of the multimodal model's spectrogram branch.

    ``eeg``: (B, 20, L>=10000) fp32, channels in ``montage.EEG_COLUMNS`` order (or (20, L) for one
    record; (B, L, 20) with ``time_major=True``, the parquet layout).  Returns
    (B, 128, 256, 4) fp32 -- (mel, time, chain) with chains LL, LP, RP, RR -- on ``eeg.device``.
    ``out`` (device input only): a contiguous fp32 tensor of that shape to write into, e.g. a slice of
    a dataset-sized buffer (the 107 k-segment set keeps its 56 GB of images resident that way).
    """
    _require_cuda()
    params = params if params is not None else mel_params(**kwargs)
    x = _as_f32(eeg)
    single = x.dim() == 2
    if single:
        x = x.unsqueeze(0)
    if x.dim() != 3:
        raise ValueError("eeg must be (B, n_ch, L) or (n_ch, L)")
    batch = x.shape[0]
    n_ch, length = (x.shape[2], x.shape[1]) if params.time_major else (x.shape[1], x.shape[2])
    _, n_time = melspec_geometry(params)
    shape = (batch, params.n_mels, n_time, params.n_chains)
    x = x.contiguous()
    if batch == 0:
        return torch.empty(shape, dtype=torch.float32, device=eeg.device)
    if not x.is_cuda:
        if out is not None:
            raise ValueError("out= is only supported for device inputs")
        out = torch.empty(shape, dtype=torch.float32, pin_memory=True)
        nat.check(nat.lib.eegspec_melspec_host_f32(_ptr(x), batch, n_ch, length, C.byref(params),
                                                   _ptr(out), torch.cuda.current_device()),
                  "eeg_to_spectrogram")
    else:
        with torch.cuda.device(x.device):
            if out is None:
                out = torch.empty(shape, dtype=torch.float32, device=x.device)
            elif (tuple(out.shape) != shape or out.dtype != torch.float32 or out.device != x.device
                  or not out.is_contiguous()):
                raise ValueError(f"out must be a contiguous fp32 tensor of shape {shape} on {x.device}")
            # The default engine needs no scratch for channel-major records (only the row gather of
            # time-major ones does): ask for the generic engines' 0.5 GiB only when the library says so.
            ws_bytes = nat.lib.eegspec_melspec_workspace_default(batch, C.byref(params))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device) if ws_bytes else None
            rc = nat.lib.eegspec_melspec_f32(_ptr(x), batch, n_ch, length, n_ch * length, C.byref(params),
                                             _ptr(out), _ptr(ws), ws_bytes, _stream_ptr(x.device))
            if rc == nat.EWORKSPACE:
                ws_bytes = nat.lib.eegspec_melspec_workspace(batch, C.byref(params))
                ws = torch.empty(ws_bytes, dtype=torch.uint8, device=x.device)
                rc = nat.lib.eegspec_melspec_f32(_ptr(x), batch, n_ch, length, n_ch * length, C.byref(params),
                                                 _ptr(out), _ptr(ws), ws_bytes, _stream_ptr(x.device))
            nat.check(rc, "eeg_to_spectrogram")
    return out[0] if single else out


def sliding_window_count(total_length: int, window: int = 10000, stride: int = 400) -> int:
    """Number of whole ``window``-sample windows at every ``stride`` samples (0 if the record is shorter)."""
    return 0 if total_length < window else (total_length - window) // stride + 1


def eeg_to_spectrogram_sliding(eeg: torch.Tensor, window: int = 10000, stride: int = 400,
                               params: Optional[nat.MelParams] = None, **kwargs) -> torch.Tensor:
    """One long record -> the log-mel image of every overlapping window (BASELINE configs[3]).

    ``eeg``: (20, L) fp32 channel-major (or (L, 20) with ``time_major=True``), e.g. a 1-hour
    recording, L = 720 000.  Window ``w`` is samples ``[w*stride, w*stride + window)`` -- the 50-s
    window ``cfg_utils.py:369-371`` cuts, taken every ``stride`` samples (400 = 2 s, the HMS
    label-offset granularity).  Returns (n_windows, 128, 256, 4); row ``w`` is bit-identical to
    ``eeg_to_spectrogram`` of the materialised slice.  The windows are read in place by the kernel:
    the record is moved (and, for a host tensor, copied to the device) once, not once per window.
    """
    _require_cuda()
    params = params if params is not None else mel_params(**kwargs)
    if window != params.crop:
        params = _copy_params(params)
        params.crop = int(window)
    x = _as_f32(eeg)
    if x.dim() != 2:
        raise ValueError("eeg must be one (n_ch, L) record (or (L, n_ch) with time_major=True)")
    if stride <= 0:
        raise ValueError("stride must be positive")
    n_ch, total = (x.shape[1], x.shape[0]) if params.time_major else (x.shape[0], x.shape[1])
    n_win = sliding_window_count(total, window, stride)
    _, n_time = melspec_geometry(params)
    shape = (n_win, params.n_mels, n_time, params.n_chains)
    if n_win == 0:
        return torch.empty(shape, dtype=torch.float32, device=eeg.device)
    on_host = not x.is_cuda
    xd = x.contiguous().cuda(non_blocking=True) if on_host else x.contiguous()
    with torch.cuda.device(xd.device):
        out = torch.empty(shape, dtype=torch.float32, device=xd.device)
        ws_bytes = nat.lib.eegspec_melspec_workspace_default(n_win, C.byref(params))
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=xd.device) if ws_bytes else None
        rc = nat.lib.eegspec_melspec_sliding_f32(_ptr(xd), n_ch, total, stride, n_win, C.byref(params),
                                                 _ptr(out), _ptr(ws), ws_bytes, _stream_ptr(xd.device))
        if rc == nat.EWORKSPACE:        # a stride / geometry the default engine does not serve
            ws_bytes = nat.lib.eegspec_melspec_workspace(n_win, C.byref(params))
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=xd.device)
            rc = nat.lib.eegspec_melspec_sliding_f32(_ptr(xd), n_ch, total, stride, n_win, C.byref(params),
                                                     _ptr(out), _ptr(ws), ws_bytes, _stream_ptr(xd.device))
        nat.check(rc, "eeg_to_spectrogram_sliding")
    if on_host:
        host = torch.empty(shape, dtype=torch.float32, pin_memory=True)
        host.copy_(out)
        return host
    return out


def _copy_params(params: nat.MelParams) -> nat.MelParams:
    q = nat.MelParams()
    C.memmove(C.byref(q), C.byref(params), C.sizeof(nat.MelParams))
    return q


def spectrogram_from_eeg(eeg: torch.Tensor, **kwargs) -> torch.Tensor:
    """One parquet-ordered record (L, 20) -> (128, 256, 4); the name the public HMS recipe uses."""
    if eeg.dim() != 2:
        raise ValueError("spectrogram_from_eeg expects one (L, n_channels) record")
    return eeg_to_spectrogram(eeg.unsqueeze(0), time_major=True, **kwargs)[0]


# --------------------------------------------------------------------------------- drop-in
def install(verbose: bool = False) -> list[str]:
    """Rebind ``compute_stft`` in the reference's modules if they are imported.

    ``DiffEEG_train.py:22`` binds the function by name at import time, so both
    ``utils.DiffEEG_utils`` and ``training.DiffEEG_train`` have to be patched."""
    patched = []
    for mod_name in ("utils.DiffEEG_utils", "training.DiffEEG_train"):
        mod = sys.modules.get(mod_name)
        if mod is not None and hasattr(mod, "compute_stft"):
            setattr(mod, "compute_stft", compute_stft)
            patched.append(mod_name)
    if verbose:
        print(f"eegspec: compute_stft rebound in {patched or 'no loaded module'}")
    return patched


def launch_count() -> int:
    return int(nat.lib.eegspec_launch_count())


PAIRS_16 = _montage.pair_indices()
