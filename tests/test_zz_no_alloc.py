"""bagpu_select_kth_static (include/ba_b200.h): the allocation-free variant of the exact order-statistic selection.

The existing entry points compute the same number inside the build: bagpu_solve records the value std::nth_element leaves at
errors_.begin() + n / 2 as median_maha of every iteration (reference src/BundleAdjuster.cpp:1356-1358).  The test takes the
residuals of the initial state from bagpu_build + bagpu_get_proj_residuals, forms the selection input |r|^2 * weight, hands
it to the new entry point in device memory (torch is only the owner of those buffers) and compares with median_maha of
iteration 0 of bagpu_solve on the same handle — two different problems, two calls in a row, so that the second call finds the
first call's state in the static scratch.  The input rebuilt on the host can differ from the device's in the last bit (the
device contracts r0*r0 + r1*r1 into an FMA), hence 1e-12 relative against the existing entry point; against np.partition of
the very same input the order statistic is bit-exact.

Collected last (file name), and self-contained: it passes alone in a fresh process and after the other GPU tests.
"""
import ctypes as C
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_N = 1 << 31   # BAGPU_SELECT_STATIC_MAX_N


def _lib():
    from ba_b200.capi import lib
    L = lib()
    L.bagpu_select_kth_static.restype = C.c_int
    L.bagpu_select_kth_static.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p]
    L.bagpu_last_error.restype = C.c_char_p
    return L


def _existing(name, scale):
    """Selection input of the initial state and the median the existing entry points report for it."""
    from ba_b200 import synth
    from ba_b200.capi import BundleAdjusterGPU
    p = synth.make(name, scale)
    s = BundleAdjusterGPU(p.lm_dim, p.pose_dim)
    o = s.default_options()
    o.error_change_threshold = 0
    o.param_change_threshold = 0
    s.set_options(o)
    s.load(p)
    s.build()
    r = s.get_proj_residuals()["residual"]              # residuals of the initial state
    w0 = np.asarray(p.arrays()["proj_weight"], np.float64)
    x = (r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1]) * w0
    s.solve(1)
    median = s.get_trace()[0]["median_maha"]
    s.close()
    return np.ascontiguousarray(x), median


def _select(L, torch, x, k, out):
    d_x = torch.from_numpy(x).cuda()
    torch.cuda.synchronize()
    rc = L.bagpu_select_kth_static(d_x.data_ptr(), x.size, k, out.data_ptr(), None)
    assert rc == 0, L.bagpu_last_error().decode()
    torch.cuda.synchronize()
    return float(out.cpu()[0])


def test_static_select_matches_existing_entry_points_twice_in_a_row():
    import torch
    assert torch.cuda.is_available()
    L = _lib()
    out = torch.full((1,), -1.0, dtype=torch.float64, device="cuda")
    cases = [_existing("c1", 1.0), _existing("c4", 0.02)]
    for rep in range(2):                                 # the whole pair twice: state of the previous call is always there
        for x, median in cases:
            assert x.size > 1000 and np.all(x >= 0)
            got = _select(L, torch, x, x.size // 2, out)
            assert got == np.partition(x, x.size // 2)[x.size // 2]             # exact on the same input
            assert abs(got - median) <= 1e-12 * abs(median), (got, median)     # the existing entry points' value


def test_static_select_other_ranks_and_no_allocation():
    import torch
    L = _lib()
    rng = np.random.default_rng(7)
    x = np.abs(rng.standard_normal(300001)) * np.exp(rng.uniform(-30, 30, 300001))
    x[::7] = x[3]                                        # ties
    x[5] = 0.0
    out = torch.zeros(1, dtype=torch.float64, device="cuda")
    d_x = torch.from_numpy(x).cuda()
    srt = np.sort(x)
    L.bagpu_select_kth_static(d_x.data_ptr(), x.size, 0, out.data_ptr(), None)   # first call on this device: module is loaded
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for k in (0, 1, x.size // 2, x.size - 2, x.size - 1):
        rc = L.bagpu_select_kth_static(d_x.data_ptr(), x.size, k, out.data_ptr(), None)
        assert rc == 0, L.bagpu_last_error().decode()
        torch.cuda.synchronize()
        assert float(out.cpu()[0]) == srt[k], k
    assert torch.cuda.mem_get_info()[0] >= free0          # nothing was allocated on the device by the calls
    # on a side stream, result read after that stream has drained
    st = torch.cuda.Stream()
    rc = L.bagpu_select_kth_static(d_x.data_ptr(), x.size, 12345, out.data_ptr(), C.c_void_p(st.cuda_stream))
    assert rc == 0
    st.synchronize()
    assert float(out.cpu()[0]) == srt[12345]


def test_static_select_rejects_bad_sizes():
    import torch
    L = _lib()
    buf = torch.zeros(8, dtype=torch.float64, device="cuda")
    for n, k in ((MAX_N + 1, 0), (0, 0), (8, 8)):
        rc = L.bagpu_select_kth_static(buf.data_ptr(), n, k, buf.data_ptr(), None)
        assert rc == 1                                   # BAGPU_ERR_INVALID
        assert b"bagpu_select_kth_static" in L.bagpu_last_error()
    assert b"BAGPU_SELECT_STATIC_MAX_N" in (L.bagpu_select_kth_static(buf.data_ptr(), MAX_N + 1, 0, buf.data_ptr(), None), L.bagpu_last_error())[1]
    assert L.bagpu_select_kth_static(None, 8, 0, buf.data_ptr(), None) == 1
