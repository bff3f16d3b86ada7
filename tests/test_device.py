"""B200Device mirrors the reference's Device / CudaDevice surface (wx_factory/device/device.py:16-68,125-220)."""

import numpy
import pytest

from wxfactory_b200 import B200Device

SURFACE = ["synchronize", "array", "pinned", "to_host", "timestamp", "elapsed", "has_128_bits_float", "get_default"]


def test_surface_matches_the_reference_device():
    for name in SURFACE:
        assert callable(getattr(B200Device, name)), name


def test_no_gpu_means_value_error_not_a_cpu_fallback():
    import torch

    if torch.cuda.is_available():
        pytest.skip("covered on the GPU box")
    with pytest.raises(ValueError, match="no GPU devices were detected"):
        B200Device()


@pytest.mark.gpu
def test_round_trip_timestamps_and_rhs_on_device_arrays(golden):
    from wxfactory_b200 import RhsB200

    dev = B200Device()
    a = numpy.arange(12.0).reshape(3, 4)
    d = dev.array(a)
    assert d.is_cuda and numpy.array_equal(dev.to_host(d), a)
    p = dev.pinned((5, 2), dtype=numpy.float64)
    assert p.shape == (5, 2) and p.dtype == numpy.float64
    prob, ex = golden("dcmip21")
    rhs = RhsB200(prob, device=dev.index)
    t0 = dev.timestamp()
    out = rhs(dev.array(ex["Q0"]))
    t1 = dev.timestamp()
    dev.synchronize()
    el = dev.elapsed([t0, t1])
    assert len(el) == 2 and el[0] > 0 and abs(el[0] - el[1]) < 1e-12
    want = ex["rhs0"]
    assert numpy.linalg.norm(dev.to_host(out) - want) <= 1e-12 * numpy.linalg.norm(want)
    assert not dev.has_128_bits_float()
    rhs.close()
