"""mdd_ctc_loss_static (include/mddctc.h): the CTC loss with its scratch in the library's static device array.

Collected last on purpose.  The entry point is called through the C ABI with ctypes
and compared with the existing entry point mdd_ctc_loss (caller-supplied workspace) on the same inputs, on two calls in
a row of different shapes, so that the second call runs over what the first one left in the shared array.
"""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-5      # the tolerances of tests/test_gpu_parity.py
GRAD_ATOL = 1e-4
MDD_E_STATIC_LIMIT = -5


def _case(torch, T, B, Cc, Lmax, seed, peaky=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(T, B, Cc, generator=g)
    if peaky:                                                    # sends some utterances through the later tiers
        x[:, :, 0] += 6.0
        x *= 4.0
    lp = torch.log_softmax(x, -1)
    il = torch.randint(T // 2, T + 1, (B,), generator=g)
    tl = torch.randint(1, Lmax + 1, (B,), generator=g)
    il[0], tl[0] = T, Lmax
    tg = torch.randint(1, Cc, (B, Lmax), generator=g)
    return lp.cuda(), tg.cuda(), il.cuda(), tl.cuda()


def _call(torch, L, static, lp, tg, il, tl, algo=0):
    T, B, Cc = lp.shape
    Lmax = tg.shape[1]
    nll = torch.full((B,), -7.0, dtype=torch.float32, device="cuda")
    grad = torch.full((T, B, Cc), -7.0, dtype=torch.float32, device="cuda")
    status = torch.full((B,), -7, dtype=torch.int32, device="cuda")
    total = torch.full((1,), -7.0, dtype=torch.float32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    head = (lp.data_ptr(), T, B, Cc, lp.stride(0), lp.stride(1), tg.data_ptr(), Lmax, il.data_ptr(), tl.data_ptr(), Lmax,
            0, 0, None, 1.0, nll.data_ptr(), grad.data_ptr(), total.data_ptr(), status.data_ptr(), algo)
    if static:
        rc = L.mdd_ctc_loss_static(*head, st)
    else:
        need = C.c_size_t()
        assert L.mdd_ctc_loss_workspace_bytes(T, B, Cc, Lmax, algo, C.byref(need)) == 0
        ws = torch.empty(need.value, dtype=torch.uint8, device="cuda")
        rc = L.mdd_ctc_loss(*head, ws.data_ptr(), need.value, st)
    torch.cuda.synchronize()
    return rc, nll.cpu(), grad.cpu(), status.cpu(), total.cpu()


def test_ctc_loss_static_matches_ctc_loss_twice_in_a_row():
    import torch
    import mddctc
    L = mddctc._lib.lib()
    # two calls in a row, different shapes and tiers; then the first shape again over the second call's leftovers
    cases = [(_case(torch, 300, 48, 43, 40, 1), 0), (_case(torch, 120, 9, 40, 12, 2, peaky=True), 0),
             (_case(torch, 300, 48, 43, 40, 1), 0), (_case(torch, 200, 7, 43, 30, 3), 1)]
    for args, algo in cases:
        rc_s, nll_s, grad_s, st_s, tot_s = _call(torch, L, True, *args, algo=algo)
        rc_w, nll_w, grad_w, st_w, tot_w = _call(torch, L, False, *args, algo=algo)
        assert rc_s == 0 and rc_w == 0, (rc_s, rc_w, L.mdd_error_string(rc_s))
        assert np.array_equal(st_s.numpy(), st_w.numpy())
        fin = torch.isfinite(nll_w)
        assert bool((torch.isfinite(nll_s) == fin).all()) and int(fin.sum()) > 0
        np.testing.assert_allclose(nll_s[fin].numpy(), nll_w[fin].numpy(), rtol=LOSS_RTOL)
        np.testing.assert_allclose(tot_s.numpy(), tot_w.numpy(), rtol=LOSS_RTOL)
        assert float((grad_s[:, fin] - grad_w[:, fin]).abs().max()) <= GRAD_ATOL
        # and against torch's own CPU loss on the first utterance, so that both being wrong together is caught
        lp, tg, il, tl = (a.cpu() for a in args)
        ref = torch.nn.functional.ctc_loss(lp[:, :1].double(), tg[:1, :int(tl[0])], il[:1], tl[:1], reduction="none")
        if bool(torch.isfinite(ref).all()):
            np.testing.assert_allclose(float(nll_s[0]), float(ref[0]), rtol=LOSS_RTOL)


def test_ctc_loss_static_rejects_shapes_beyond_its_array():
    import torch
    import mddctc
    L = mddctc._lib.lib()
    T, B, Cc, Lmax = 500, 1024, 43, 60
    need = C.c_size_t()
    assert L.mdd_ctc_loss_workspace_bytes(T, B, Cc, Lmax, 0, C.byref(need)) == 0
    assert need.value > (16 << 20)
    one = torch.zeros(1, device="cuda")
    rc = L.mdd_ctc_loss_static(one.data_ptr(), T, B, Cc, B * Cc, Cc, one.data_ptr(), Lmax, one.data_ptr(), one.data_ptr(),
                               Lmax, 0, 0, None, 1.0, one.data_ptr(), None, None, None, 0,
                               torch.cuda.current_stream().cuda_stream)
    assert rc == MDD_E_STATIC_LIMIT
    L.mdd_error_string.restype = C.c_char_p
    assert b"MDD_CTC_STATIC_WS_BYTES" in L.mdd_error_string(rc)
    torch.cuda.synchronize()                                     # nothing was launched, nothing faulted
