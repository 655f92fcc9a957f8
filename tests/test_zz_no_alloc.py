"""GPU: cgb200_dot_static, the entry point that allocates nothing (include/cgb200.h; no counterpart
in the reference), against the existing entry point
dot_product, through the C ABI.  Collected last on purpose.  Passes
alone in a fresh process -- the new entry point needs no library context -- and after the others."""
import ctypes as C

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

EPS = np.finfo(np.float64).eps


def _static_dot(L, ta, tb, out, stream):
    import torch

    rc = L.cgb200_dot_static(C.c_void_p(ta.data_ptr()), C.c_void_p(tb.data_ptr()), ta.numel(),
                             C.c_void_p(out.data_ptr()), C.c_void_p(stream))
    torch.cuda.synchronize()
    return rc


def test_dot_static_matches_dot_product_on_two_calls_in_a_row(pkg):
    import torch

    L = pkg.lib()                      # (no gpu_lib fixture: the first call below runs before cgb200_init)
    assert L.cgb200_device_available() == 1, "gpu test on a box without a CUDA device"
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    out = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    # two calls in a row, different lengths and data: the second finds the first one's scratch contents
    for n, sa, sb in [((1 << 21) + 12345, 31, 32), (777_777, 33, 34), (5, 35, 36), (0, 1, 2)]:
        a, b = H.rhs(n, sa), H.rhs(n, sb)
        ta, tb = torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()
        out.fill_(float("nan"))
        assert _static_dot(L, ta, tb, out, stream) == 0, pkg.last_error()
        got = float(out.item())
        va = pkg.MvSparse(n, n, H.dptr(a), None, None)
        vb = pkg.MvSparse(n, n, H.dptr(b), None, None)
        want = L.dot_product(C.byref(va), C.byref(vb)) if n > 0 else 0.0   # the existing entry point
        scale = float(np.dot(np.abs(a), np.abs(b)))
        assert abs(got - want) <= 4 * max(n, 1) * EPS * scale, (n, got, want)
        assert abs(got - H.oracle().oracle_dot(n, H.dptr(a), H.dptr(b))) <= 4 * max(n, 1) * EPS * scale
        # same inputs again: same bits (fixed chunk order), whatever the scratch held
        assert _static_dot(L, ta, tb, out, stream) == 0, pkg.last_error()
        assert float(out.item()) == got


def test_dot_static_refuses_what_its_scratch_cannot_hold(pkg):
    import torch

    L = pkg.lib()
    t = torch.zeros(8, dtype=torch.float64, device="cuda")
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    p = C.c_void_p(t.data_ptr())
    assert L.cgb200_dot_static(p, p, pkg.DOT_STATIC_MAX_N + 1, C.c_void_p(out.data_ptr()), None) == -1
    assert "CGB200_DOT_STATIC_MAX_N" in pkg.last_error()
    assert L.cgb200_dot_static(None, p, 8, C.c_void_p(out.data_ptr()), None) == -1
    assert L.cgb200_dot_static(p, p, -1, C.c_void_p(out.data_ptr()), None) == -1
    # the largest accepted size runs (2 GiB per vector; the same buffer twice)
    big = torch.ones(pkg.DOT_STATIC_MAX_N, dtype=torch.float64, device="cuda")
    assert L.cgb200_dot_static(C.c_void_p(big.data_ptr()), C.c_void_p(big.data_ptr()), big.numel(),
                               C.c_void_p(out.data_ptr()), None) == 0, pkg.last_error()
    torch.cuda.synchronize()
    assert float(out.item()) == float(pkg.DOT_STATIC_MAX_N)
