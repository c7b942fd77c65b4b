"""Pin the oracle's five OpenCV primitives bit-exact to the cv2 4.13.0 wheel.

The reference does not vendor or pin OpenCV and has no tests (parity unpinned by the reference
itself); OpenCV 4.13.0 semantics are the project's anchor.  cv2 exists in this image both here
and on the GPU box.
"""
import ctypes as C

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

u8p = C.POINTER(C.c_uint8)


class KP(C.Structure):
    _fields_ = [("x", C.c_int), ("y", C.c_int), ("s", C.c_int)]


@pytest.fixture(scope="module")
def lib(oracle_built):
    L = oracle_built.Oracle.load()
    L.cvp_resize_linear_u8.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, u8p, C.c_size_t, C.c_int, C.c_int]
    L.cvp_border_reflect101.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_int]
    L.cvp_gaussian7_sigma2.argtypes = [u8p, C.c_size_t, u8p, C.c_size_t, C.c_int, C.c_int]
    L.cvp_fast_atan2.restype = C.c_float
    L.cvp_fast_atan2.argtypes = [C.c_float, C.c_float]
    L.cvp_fast9_16.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(KP), C.c_int]
    cv2.setNumThreads(1)
    return L


def _p(a):
    return a.ctypes.data_as(u8p)


def test_cv2_version_is_the_anchor():
    assert cv2.__version__.startswith("4.13"), cv2.__version__


# all seven level transitions of both benchmark shapes + odd cases incl. upscale
RESIZE_CASES = [(1241, 376, 1034, 313), (1034, 313, 862, 261), (862, 261, 718, 218), (718, 218, 598, 181),
                (598, 181, 499, 151), (499, 151, 416, 126), (416, 126, 346, 105),
                (640, 480, 533, 400), (533, 400, 444, 333), (444, 333, 370, 278), (370, 278, 309, 231),
                (309, 231, 257, 193), (257, 193, 214, 161), (214, 161, 179, 134),
                (100, 90, 179, 134), (57, 43, 57, 20), (33, 9, 20, 9)]


@pytest.mark.parametrize("sw,sh,dw,dh", RESIZE_CASES)
def test_resize_linear(lib, sw, sh, dw, dh):
    rng = np.random.default_rng(sw * 7 + dh)
    src = rng.integers(0, 256, (sh, sw), dtype=np.uint8)
    ref = cv2.resize(src, (dw, dh), interpolation=cv2.INTER_LINEAR)
    out = np.zeros((dh, dw), np.uint8)
    lib.cvp_resize_linear_u8(_p(src), sw, sw, sh, _p(out), dw, dw, dh)
    assert np.array_equal(ref, out)


@pytest.mark.parametrize("w,h", [(100, 50), (346, 105), (21, 20), (1241, 376)])
def test_border_reflect101(lib, w, h):
    rng = np.random.default_rng(w)
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    ref = cv2.copyMakeBorder(src, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
    buf = np.zeros((h + 38, w + 38), np.uint8)
    buf[19:19 + h, 19:19 + w] = src
    lib.cvp_border_reflect101(_p(buf), w + 38, w, h, 19)
    assert np.array_equal(ref, buf)


@pytest.mark.parametrize("w,h", [(1241, 376), (640, 480), (179, 134), (37, 21), (7, 7), (5, 9), (3, 3)])
def test_gaussian_blur(lib, w, h):
    rng = np.random.default_rng(h)
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    ref = cv2.GaussianBlur(src, (7, 7), 2, None, 2, cv2.BORDER_REFLECT_101)
    out = np.zeros_like(src)
    lib.cvp_gaussian7_sigma2(_p(src), w, _p(out), w, w, h)
    assert np.array_equal(ref, out)
    inplace = src.copy()
    lib.cvp_gaussian7_sigma2(_p(inplace), w, _p(inplace), w, w, h)
    assert np.array_equal(ref, inplace)


def test_fast_atan2(lib):
    rng = np.random.default_rng(11)
    ys = rng.integers(-2_000_000, 2_000_000, 20000)
    xs = rng.integers(-2_000_000, 2_000_000, 20000)
    pairs = list(zip(ys.tolist(), xs.tolist())) + [(0, 0), (0, -5), (-1, 0), (5, 5), (-5, 5), (1, 0), (0, 1),
                                                   (7, -7), (-7, -7), (3, 4), (4, 3)]
    for y, x in pairs:
        a = np.float32(cv2.fastAtan2(float(y), float(x)))
        b = np.float32(lib.cvp_fast_atan2(float(y), float(x)))
        assert a == b, (y, x, a, b)


def _textures(rng, w, h, kind):
    if kind == 0:
        return rng.integers(0, 256, (h, w), dtype=np.uint8)
    if kind == 1:
        blocks = (rng.integers(0, 2, (h // 4 + 1, w // 4 + 1)) * rng.integers(30, 200)).astype(np.uint8)
        img = np.kron(blocks, np.ones((4, 4), np.uint8))[:h, :w]
        return (img + rng.integers(0, 6, (h, w))).astype(np.uint8)
    return cv2.GaussianBlur(rng.integers(0, 256, (h, w), dtype=np.uint8), (5, 5), 1.2)


def test_fast9_16_with_nms(lib):
    rng = np.random.default_rng(5)
    total = 0
    for trial in range(45):
        w, h = int(rng.integers(7, 80)), int(rng.integers(7, 80))
        img = np.ascontiguousarray(_textures(rng, w, h, trial % 3))
        for th in (20, 7, 1, 100):
            det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True,
                                                 type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            ref = [(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in det.detect(img)]
            buf = (KP * 4096)()
            n = lib.cvp_fast9_16(_p(img), w, w, h, th, buf, 4096)
            got = [(buf[i].x, buf[i].y, buf[i].s) for i in range(n)]
            assert got == ref, (w, h, th)   # same keypoints, same scores, same (row-major) order
            total += n
    assert total > 5000


def test_threshold_fallback_identity(lib):
    """K(20) == {p in K(7): score >= 20}: what lets the GPU use one score plane per cell."""
    rng = np.random.default_rng(9)
    checked = 0
    for trial in range(60):
        w, h = int(rng.integers(20, 44)), int(rng.integers(20, 44))
        img = np.ascontiguousarray(_textures(rng, w, h, trial % 3))
        b7, b20 = (KP * 2048)(), (KP * 2048)()
        n7 = lib.cvp_fast9_16(_p(img), w, w, h, 7, b7, 2048)
        n20 = lib.cvp_fast9_16(_p(img), w, w, h, 20, b20, 2048)
        k7 = [(b7[i].x, b7[i].y, b7[i].s) for i in range(n7)]
        k20 = [(b20[i].x, b20[i].y, b20[i].s) for i in range(n20)]
        assert k20 == [k for k in k7 if k[2] >= 20]
        checked += n7
    assert checked > 2000


# ---- the step before the extractor: cvtColor and remap -------------------

@pytest.mark.parametrize("code,ch,rgb", [("COLOR_BGR2GRAY", 3, False), ("COLOR_RGB2GRAY", 3, True),
                                         ("COLOR_BGRA2GRAY", 4, False), ("COLOR_RGBA2GRAY", 4, True)])
def test_cvt_gray_matches_cv2(code, ch, rgb, oracle_built):
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, (97, 333, ch), dtype=np.uint8)
    assert np.array_equal(oracle_built.cvt_gray(img, rgb), cv2.cvtColor(img, getattr(cv2, code)))


def test_cvt_gray_exhaustive(oracle_built):
    v = np.arange(256, dtype=np.uint8)
    full = np.stack(np.meshgrid(v, v, v, indexing="ij"), -1).reshape(4096, 4096, 3)
    assert np.array_equal(oracle_built.cvt_gray(full, False), cv2.cvtColor(full, cv2.COLOR_BGR2GRAY))


def _euroc_maps(eye):
    from orb_slam2_test_b200 import synth
    K, D, R, P = synth.EUROC[eye]
    m1, m2 = cv2.initUndistortRectifyMap(np.array(K).reshape(3, 3), np.array(D), np.array(R).reshape(3, 3),
                                         np.array(P).reshape(3, 4)[:, :3], (752, 480), cv2.CV_32F)
    return m1, m2


@pytest.mark.parametrize("ch", [1, 3, 4])
def test_remap_matches_cv2(ch, oracle_built):
    rng = np.random.default_rng(12 + ch)
    src = rng.integers(0, 256, (120, 200) if ch == 1 else (120, 200, ch), dtype=np.uint8)
    h, w = 100, 180
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    cases = [((xx * 1.07 - 8 + rng.normal(0, 1.5, (h, w))).astype(np.float32),
              (yy * 1.15 - 6 + rng.normal(0, 1.5, (h, w))).astype(np.float32)),
             (rng.uniform(-5, 205, (h, w)).astype(np.float32), rng.uniform(-5, 125, (h, w)).astype(np.float32)),
             (xx + np.float32(1 / 64), yy + np.float32(3 / 64)),            # ties of the 1/32 quantisation
             (rng.uniform(-7e4, 7e4, (h, w)).astype(np.float32), rng.uniform(-7e4, 7e4, (h, w)).astype(np.float32))]
    for mx, my in cases:
        assert np.array_equal(oracle_built.remap_linear(src, mx, my), cv2.remap(src, mx, my, cv2.INTER_LINEAR))


@pytest.mark.parametrize("eye", [0, 1])
def test_remap_euroc_rectification_matches_cv2(eye, oracle_built):
    """The call of Examples/Stereo/stereo_euroc.cc:136-137 with the maps of :107-108."""
    from orb_slam2_test_b200 import synth
    m1, m2 = _euroc_maps(eye)
    img = synth.make_frame((480, 752), seed=40, eye=eye)
    assert np.array_equal(oracle_built.remap_linear(img, m1, m2), cv2.remap(img, m1, m2, cv2.INTER_LINEAR))
    # the package's own map generator (inputs for the GPU tests) agrees with OpenCV's to float rounding
    x, y = synth.rectify_maps(*synth.EUROC[eye], size=(752, 480))
    assert np.abs(x - m1).max() < 1e-3 and np.abs(y - m2).max() < 1e-3
