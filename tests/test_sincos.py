"""orb_sincos.cuh (the routine the descriptor kernel uses for cos/sin of the keypoint angle) must be
bit-identical to the host libm's sincosf, which is what the reference binary calls
(src/ORBextractor.cc:121-122 compiled at -O3 merges cos+sin into one sincosf).  The same inline
function is compiled for the host into liborb_b200.so and tapped through borb_debug_sincos."""
import ctypes as C
import ctypes.util

import numpy as np


def test_sincos_matches_libm():
    from orb_slam2_test_b200 import extractor
    lib = extractor.load_library()
    libm = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.sincosf.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    rng = np.random.default_rng(9)
    factor = np.float32(np.pi / np.float32(180.0))
    # angles the path can produce: fastAtan2 degrees in [0, 360] times factorPI; plus the
    # quadrant boundaries and tiny values
    deg = np.concatenate([rng.uniform(0, 360, 150_000).astype(np.float32),
                          np.arange(0, 361, 0.25, dtype=np.float32),
                          np.float32([0, 1e-6, 1e-4, 44.999, 45, 45.001, 89.9999, 90, 135, 180, 270, 359.9999, 360])])
    rad = (deg * factor).astype(np.float32)
    s1, c1, s2, c2 = C.c_float(), C.c_float(), C.c_float(), C.c_float()
    bad = 0
    for r in rad.tolist():
        lib.borb_debug_sincos(r, C.byref(s1), C.byref(c1))
        libm.sincosf(r, C.byref(s2), C.byref(c2))
        bad += (s1.value != s2.value) or (c1.value != c2.value)
    assert bad == 0, f"{bad} of {len(rad)} angles differ from libm sincosf"
