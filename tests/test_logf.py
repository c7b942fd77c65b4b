"""orb_logf.cuh (the logf inside the relocalisation matcher's MapPoint::PredictScale, src/MapPoint.cc:575-590) must be
bit-identical to the host libm's logf, which is what the reference binary calls.  The host copy of the inline function is
tapped through borb_debug_logf, the device copy through borb_debug_logf_device.  (All 2^31 positive floats were checked once
against glibc 2.39 when the table was read out of libm: 0 differences; this is the sample that travels.)"""
import ctypes as C
import ctypes.util

import numpy as np
import pytest


def _inputs():
    rng = np.random.default_rng(17)
    bits = rng.integers(0, 0x7f800000, 3_000_000, dtype=np.uint32)  # every positive finite float, subnormals included
    ratios = rng.uniform(0.3, 12.0, 1_000_000).astype(np.float32)   # what mfMaxDistance / dist3D looks like
    edge = np.array([0.0, -0.0, 1.0, np.inf, -1.0, np.nan, 1e-45, 1.1754942e-38, 1.17549435e-38, 3.4028235e38,
                     0.99999994, 1.0000001, 0.7, 1.4, 1.2, 1.2 ** 7], np.float32)
    return np.concatenate([bits.view(np.float32), ratios, edge])


def _libm_logf(x: np.ndarray) -> np.ndarray:
    libm = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.logf.restype = C.c_float
    libm.logf.argtypes = [C.c_float]
    return np.array([libm.logf(v) for v in x.tolist()], np.float32)


def test_host_copy_matches_libm():
    from orb_slam2_test_b200 import extractor
    x = _inputs()[::8]
    got = extractor.debug_logf(x)
    want = _libm_logf(x)
    assert np.array_equal(got.view(np.uint32) & np.where(np.isnan(want), 0, 0xffffffff).astype(np.uint32),
                          want.view(np.uint32) & np.where(np.isnan(want), 0, 0xffffffff).astype(np.uint32))
    assert np.array_equal(np.isnan(got), np.isnan(want))


@pytest.mark.gpu
def test_device_copy_matches_the_host_copy_and_libm():
    from orb_slam2_test_b200 import extractor
    x = _inputs()
    dev = extractor.debug_logf(x, device=True)
    host = extractor.debug_logf(x)
    ok = np.isnan(host) & np.isnan(dev) | (host.view(np.uint32) == dev.view(np.uint32))
    assert ok.all(), f"{(~ok).sum()} of {len(x)} inputs differ between the device and the host copy"
    sample = x[::16]
    want = _libm_logf(sample)
    got = dev[::16]
    ok = np.isnan(want) & np.isnan(got) | (want.view(np.uint32) == got.view(np.uint32))
    assert ok.all()
