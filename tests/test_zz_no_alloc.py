"""No-allocation entry point ``eegspec_classifier_spec_static_f32`` (collected last on purpose:
a failure here must not hide the parity tests).

It has no counterpart in the reference; it serves callers that cannot call cudaMalloc.  Same kernels as
``eegspec_classifier_spec_f32`` with the scratch in a device array built into the library, so the results must be
the same bits as the workspace entry point's (the test still states the tolerance the row (f)3 tests use, 1e-4 on
the [0, 1] image).  Two calls in a row with different inputs share the array: the second must not see the first."""
import ctypes as C

import numpy as np
import pytest
import torch

from b200spec import _native as nat

pytestmark = pytest.mark.gpu
ATOL = 1e-4
ROWS, COLS = 400, 300


def _params():
    p = nat.ClsSpecParams()
    nat.lib.eegspec_cls_spec_params_default(C.byref(p))
    p.rows, p.cols = ROWS, COLS
    return p


def _windows(batch, seed):
    rng = np.random.default_rng(seed)
    win = rng.gamma(2.0, 8.0, size=(batch, ROWS, COLS)).astype(np.float32)
    win[0, 3, 7] = np.nan
    if batch > 2:
        win[2, 100:110, :50] = np.nan
    win[:, 300:] = 0.0
    return torch.from_numpy(win).cuda()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _workspace_call(win, p):
    batch = win.shape[0]
    out = torch.empty((batch, 3, ROWS, COLS), dtype=torch.float32, device="cuda")
    status = torch.empty(batch, dtype=torch.int32, device="cuda")
    ws_bytes = nat.lib.eegspec_classifier_spec_workspace(batch, C.byref(p))
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    rc = nat.lib.eegspec_classifier_spec_f32(win.data_ptr(), batch, C.byref(p), out.data_ptr(), status.data_ptr(),
                                             ws.data_ptr(), ws_bytes, _stream())
    assert rc == 0, nat.strerror(rc)
    torch.cuda.synchronize()
    return out, status


def test_static_scratch_entry_matches_the_workspace_entry_on_two_calls_in_a_row():
    p = _params()
    max_batch = 1
    while nat.lib.eegspec_classifier_spec_workspace(max_batch + 1, C.byref(p)) <= nat.CLS_SPEC_STATIC_BYTES:
        max_batch += 1
    assert max_batch == 17                                   # the limit include/eegspec.h states
    # inputs and outputs of both calls exist before the first call: nothing is allocated between or inside them
    cases = [(_windows(5, 21), torch.full((5, 3, ROWS, COLS), -7.0, device="cuda"), torch.full((5,), -1, dtype=torch.int32, device="cuda")),
             (_windows(max_batch, 22), torch.full((max_batch, 3, ROWS, COLS), -7.0, device="cuda"),
              torch.full((max_batch,), -1, dtype=torch.int32, device="cuda"))]
    want = [_workspace_call(win, p) for win, _, _ in cases]
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for win, out, status in cases:                           # two calls in a row, same stream, no sync in between
        rc = nat.lib.eegspec_classifier_spec_static_f32(win.data_ptr(), win.shape[0], C.byref(p), out.data_ptr(),
                                                        status.data_ptr(), _stream())
        assert rc == 0, nat.strerror(rc)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for (win, out, status), (w_out, w_status) in zip(cases, want):
        assert torch.equal(status, w_status) and int(status.sum()) == 0
        err = float((out - w_out).abs().max())
        assert err <= ATOL, f"batch {win.shape[0]}: max abs difference {err:.3e}"
        assert torch.equal(out, w_out)                       # same kernels, same order: the same bits
        assert float(out.min()) >= 0.0 and float(out.max()) <= 1.0


def test_static_scratch_entry_refuses_what_does_not_fit():
    p = _params()
    win = torch.zeros((18, ROWS, COLS), device="cuda")
    out = torch.full((18, 3, ROWS, COLS), -7.0, device="cuda")
    rc = nat.lib.eegspec_classifier_spec_static_f32(win.data_ptr(), 18, C.byref(p), out.data_ptr(), None, _stream())
    assert rc == nat.ESTATICLIMIT
    assert "built-in scratch" in nat.strerror(rc)
    rc = nat.lib.eegspec_classifier_spec_static_f32(win.data_ptr(), 1 << 40, C.byref(p), out.data_ptr(), None, _stream())
    assert rc == nat.ESTATICLIMIT
    torch.cuda.synchronize()
    assert float(out.max()) == -7.0                          # nothing was launched
    assert nat.lib.eegspec_classifier_spec_static_f32(None, 1, C.byref(p), out.data_ptr(), None, _stream()) == nat.EINVAL
