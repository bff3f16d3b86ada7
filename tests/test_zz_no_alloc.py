"""wxb_vec_multidot_static (include/wxb200.h): the no-allocation variant of the multi-dot, whose reduction scratch is a
file-scope device array inside the library.  Collected last on purpose.

Checked through the C ABI against the existing entry point wxb_vec_multidot on the same inputs: both run the same two
kernels in the same fixed order, so the results are equal bit for bit (the tolerance the multi-dot tests use against
numpy, 1e-13 of sum |x||y|, is asserted as well), on two calls in a row with different inputs (the shared scratch keeps
the previous call's contents), and the size limit is rejected with WXB_EINVAL."""

import ctypes

import numpy
import pytest

pytestmark = pytest.mark.gpu


def test_multidot_static_matches_multidot_on_two_calls_in_a_row():
    import torch

    from wxfactory_b200 import capi

    lib = capi.load()
    torch.cuda.set_device(0)
    n = 200_001
    scratch = torch.empty(int(lib.wxb_vec_multidot_scratch(capi.WXB_VEC_MAX_ROWS, 2)), dtype=torch.float64, device="cuda")
    for call, (m, k) in enumerate([(22, 2), (5, 1)]):
        rng = numpy.random.default_rng(100 + call)
        V = rng.standard_normal((m + 2, n))
        Vd = torch.from_numpy(V).cuda()
        W = Vd[m : m + k]
        want = torch.full((m * k,), float("nan"), dtype=torch.float64, device="cuda")
        got = torch.full((m * k,), float("nan"), dtype=torch.float64, device="cuda")
        assert lib.wxb_vec_multidot(n, m, Vd.data_ptr(), Vd.stride(0), k, W.data_ptr(), W.stride(0), want.data_ptr(), scratch.data_ptr(), None) == 0
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        assert lib.wxb_vec_multidot_static(n, m, Vd.data_ptr(), Vd.stride(0), k, W.data_ptr(), W.stride(0), got.data_ptr(), None) == 0
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        g, w = got.cpu().numpy().reshape(m, k), want.cpu().numpy().reshape(m, k)
        assert numpy.array_equal(g, w)
        ref = V[:m] @ V[m : m + k].T
        assert numpy.all(numpy.abs(g - ref) <= 1e-13 * (numpy.abs(V[:m]) @ numpy.abs(V[m : m + k]).T))
    # beyond the static scratch: rejected, nothing launched
    big = torch.zeros(capi.WXB_VEC_MAX_ROWS + 1, 64, dtype=torch.float64, device="cuda")
    out = torch.zeros(2 * (capi.WXB_VEC_MAX_ROWS + 1), dtype=torch.float64, device="cuda")
    assert lib.wxb_vec_multidot_static(64, capi.WXB_VEC_MAX_ROWS + 1, big.data_ptr(), 64, 1, big.data_ptr(), 64, out.data_ptr(), None) == -1
    assert lib.wxb_vec_multidot_static(0, 1, big.data_ptr(), 64, 1, big.data_ptr(), 64, out.data_ptr(), None) == -1
