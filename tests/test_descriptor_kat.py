"""Known-answer test of the FLOAT FORM of the descriptor's pattern rotation (Findings 3-4).

computeOrbDescriptor (/root/reference/src/ORBextractor.cc:128-130) samples the blurred image at
    cvRound(pattern[idx].x*b + pattern[idx].y*a),  cvRound(pattern[idx].x*a - pattern[idx].y*b).
The reference builds with -O3 -march=native, so its object holds the CONTRACTED form fma(x, b, rn(y*a)) -- one
rounding less than the expression as written.  The two forms differ in the last bit of the float often, and in the
rounded integer when the value sits within an ulp of k + 0.5.  This test searches (angle, pattern point) pairs where
exactly that happens, builds for each a patch on which the two pixels the two forms read give opposite answers, and
holds every implementation to the reference's own object code (oracle/_ref/librefkat.so: the reference's
src/ORBextractor.cc included as text and compiled with its flags):
  - the CPU restatement (oracle/orb_oracle.cpp) equals the reference object,
  - the as-written (unfused) form, emulated here, does NOT -- so the test is decisive,
  - the CUDA device function of the extract kernels (through borb_debug_describe) equals the reference object.
"""
import ctypes as C
import json
from pathlib import Path

import numpy as np
import pytest

GOLD = Path(__file__).resolve().parent / "golden" / "desc_kat.json"


def _sincosf(rad: np.float32):
    libm = C.CDLL("libm.so.6")
    libm.sincosf.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    s, c = C.c_float(), C.c_float()
    libm.sincosf(C.c_float(float(rad)), C.byref(s), C.byref(c))
    return np.float32(s.value), np.float32(c.value)


def _coords(pat, a, b, fused: bool):
    """(row, col) of every pattern point for cos = a, sin = b (float32), in the fused or the as-written form"""
    x, y = pat[:, 0].astype(np.float64), pat[:, 1].astype(np.float64)
    a64, b64 = np.float64(a), np.float64(b)
    ya, yb = (y * a64).astype(np.float32).astype(np.float64), (y * b64).astype(np.float32).astype(np.float64)
    if fused:  # x*b is exact in double, the sum of two <= 28-bit terms is exact too: ONE rounding to float
        row, col = (x * b64 + ya).astype(np.float32), (x * a64 - yb).astype(np.float32)
    else:
        xb, xa = (x * b64).astype(np.float32).astype(np.float64), (x * a64).astype(np.float32).astype(np.float64)
        row, col = (xb + ya).astype(np.float32), (xa - yb).astype(np.float32)
    return np.rint(row).astype(np.int64), np.rint(col).astype(np.int64)


def _describe_emulated(img, cx, cy, pat, a, b, fused):
    r, c = _coords(pat, a, b, fused)
    v = img[cy + r, cx + c].astype(np.int32)
    bits = (v[0::2] < v[1::2]).astype(np.uint8)
    return np.packbits(bits.reshape(32, 8), axis=1, bitorder="little").ravel()


def find_cases(pat, want=24):
    """angles (float32 degrees) and pattern indices where the two forms round to different pixels.  Such angles sit
    next to a rounding tie, so they are looked for there: for pattern point (x, y) and half-integer h the row
    x sin t + y cos t = R sin(t + phi) equals h at t = asin(h / R) - phi (and pi - asin - phi), the column
    x cos t - y sin t = R cos(t + phi) at t = +-acos(h / R) - phi; the float32 angles around each solution are tried."""
    factor_pi = np.float32(np.pi / 180.0)  # (float)(CV_PI/180.f)
    cases, seen = [], set()
    for idx in range(0, 512, 7):
        x, y = float(pat[idx, 0]), float(pat[idx, 1])
        R, phi = np.hypot(x, y), np.arctan2(y, x)
        for k in range(-18, 18):
            h = k + 0.5
            if abs(h) >= R:
                continue
            sols = [np.arcsin(h / R) - phi, np.pi - np.arcsin(h / R) - phi, np.arccos(h / R) - phi, -np.arccos(h / R) - phi]
            for t in sols:
                deg = np.float32(np.degrees(t) % 360.0)
                lo = deg
                for _ in range(6):
                    lo = np.nextafter(lo, np.float32(-1))
                for _ in range(13):
                    if 0.0 <= lo < 360.0 and float(lo) not in seen:
                        s_, c_ = _sincosf(np.float32(lo * factor_pi))
                        one = pat[idx:idx + 1]
                        if _coords(one, c_, s_, True) != _coords(one, c_, s_, False):
                            seen.add(float(lo))
                            cases.append((float(lo), idx))
                    lo = np.nextafter(lo, np.float32(400))
        if len(cases) >= 4 * want:
            break
    return cases[::max(1, len(cases) // want)][:want]


def build_patch(pat, deg, idx):
    """65 x 65 patch, keypoint at (32, 32): the pixel the fused form reads for pattern point idx and the one the unfused
    form reads get opposite extremes, everything else mid-grey, so that bit idx // 2 flips between the forms"""
    factor_pi = np.float32(np.pi / 180.0)
    s, c = _sincosf(np.float32(np.float32(deg) * factor_pi))
    rf, cf = _coords(pat, c, s, True)
    ru, cu = _coords(pat, c, s, False)
    img = np.full((65, 65), 128, np.uint8)
    lo_first = idx % 2 == 0  # even index = t0 of (t0 < t1)
    img[32 + rf[idx], 32 + cf[idx]] = 0 if lo_first else 255
    img[32 + ru[idx], 32 + cu[idx]] = 255 if lo_first else 0
    return img, c, s


@pytest.fixture(scope="module")
def cases(oracle_built):
    from oracle import pyoracle
    pat = pyoracle.pattern().astype(np.int64)
    if GOLD.exists():
        got = [tuple(x) for x in json.loads(GOLD.read_text())["cases"]]
    else:
        got = find_cases(pat)
        GOLD.write_text(json.dumps({"what": "(angle in degrees as float32, pattern point index) pairs where fma(x, b, rn(y*a)) and "
                                            "rn(rn(x*b) + rn(y*a)) round to different pixels; found by tests/test_descriptor_kat.py::find_cases",
                                    "cases": got}, indent=1))
    assert len(got) >= 16
    return pat, got


def test_restatement_is_the_fused_form_and_the_test_is_decisive(cases):
    from oracle import pyoracle
    pat, got = cases
    kat = pyoracle.HERE / "_ref" / "librefkat.so"
    ref = C.CDLL(str(kat)) if kat.exists() else None
    flips = 0
    for deg, idx in got:
        img, c, s = build_patch(pat, deg, idx)
        ours = pyoracle.describe(img, 32, 32, np.float32(deg))
        fused = _describe_emulated(img, 32, 32, pat, c, s, True)
        unfused = _describe_emulated(img, 32, 32, pat, c, s, False)
        assert np.array_equal(ours, fused), f"restatement is not the fused form at {deg} deg, point {idx}"
        assert not np.array_equal(fused, unfused), f"case ({deg}, {idx}) does not separate the forms"
        flips += int(np.unpackbits(fused ^ unfused).sum())
        if ref is not None:  # the reference's own object code
            want = np.zeros(32, np.uint8)
            ref.refkat_describe(img.ctypes.data_as(C.c_void_p), 65, 65, C.c_size_t(img.strides[0]), 32, 32,
                                C.c_float(deg), want.ctypes.data_as(C.c_void_p))
            assert np.array_equal(want, ours), f"reference object differs from the restatement at {deg} deg, point {idx}"
            assert not np.array_equal(want, unfused)
    assert flips >= len(got)


@pytest.mark.gpu
def test_cuda_device_function_is_the_fused_form(cases):
    from oracle import pyoracle
    from orb_slam2_test_b200 import ORBextractor
    pat, got = cases
    ext = ORBextractor(500, 1.2, 8, 20, 7)
    for deg, idx in got:
        img, c, s = build_patch(pat, deg, idx)
        want = pyoracle.describe(img, 32, 32, np.float32(deg))
        g = ext.debug_describe(img, 32, 32, np.float32(deg))
        assert np.array_equal(g, want), f"CUDA K6 differs at {deg} deg, point {idx}"
        assert not np.array_equal(g, _describe_emulated(img, 32, 32, pat, c, s, False))
    # and on textured patches at many angles
    rng = np.random.default_rng(3)
    tex = rng.integers(0, 256, (65, 65), dtype=np.uint8)
    for deg in np.linspace(0, 359.9, 97, dtype=np.float32):
        assert np.array_equal(ext.debug_describe(tex, 32, 32, deg), pyoracle.describe(tex, 32, 32, deg))
