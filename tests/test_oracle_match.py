"""Arithmetic pins and properties of the matcher restatement (oracle/match_oracle.cpp) that do not need
the reference tree."""
import ctypes as C

import numpy as np
import pytest

import matchcase


def test_small_gemm_form_is_cv2s():
    """x3Dc = Rcw * x3Dw + tcw (src/ORBmatcher.cc:1537): OpenCV 4.13's float gemm on 3x3 * 3x1 sums the
    individually rounded products left to right and adds C; the restatement and oracle/cvcompat use that form."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(3)
    for _ in range(4000):
        R = rng.standard_normal((3, 3)).astype(np.float32)
        x = (rng.standard_normal((3, 1)) * 30).astype(np.float32)
        t = rng.standard_normal((3, 1)).astype(np.float32)
        want = cv2.gemm(R, x, 1.0, t, 1.0)
        got = np.zeros((3, 1), np.float32)
        for i in range(3):
            s = np.float32(0)
            for k in range(3):
                s = np.float32(s + np.float32(R[i, k] * x[k, 0]))
            got[i, 0] = np.float32(s + t[i, 0])
        assert np.array_equal(got, want)


def test_transposed_gemm_form_is_cv2s():
    """Ow = -Rcw.t() * tcw (src/ORBmatcher.cc:1516, :1676; src/Frame.cc:334) is cv::gemm(R, t, -1, None, 0, GEMM_1_T): with a
    transposed operand OpenCV 4.13 leaves the small-matrix shortcut and sums the products in DOUBLE, left to right, times
    alpha, one rounding -- not the float form of the plain product above (which differs on ~80 % of random inputs).
    oracle/cvcompat (MatT), the restatement and the library's camera_centre use that form."""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(5)
    n_float_form_differs = 0
    for _ in range(4000):
        R = rng.standard_normal((3, 3)).astype(np.float32)
        t = (rng.standard_normal((3, 1)) * 10).astype(np.float32)
        want = cv2.gemm(R, t, -1.0, None, 0.0, flags=cv2.GEMM_1_T)
        got = np.zeros((3, 1), np.float32)
        flt = np.zeros((3, 1), np.float32)
        for i in range(3):
            s = 0.0
            f = np.float32(0)
            for k in range(3):
                s = s + float(R[k, i]) * float(t[k, 0])
                f = np.float32(f + np.float32(np.float32(-R[k, i]) * t[k, 0]))
            got[i, 0] = np.float32(s * -1.0)
            flt[i, 0] = f
        assert np.array_equal(got, want)
        n_float_form_differs += not np.array_equal(flt, want)
    assert n_float_form_differs > 1000


def test_norm_form_is_cv2s():
    """dist3D = cv::norm(PO) (src/ORBmatcher.cc:1712): squares summed in double, left to right, double sqrt"""
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(6)
    for _ in range(4000):
        v = (rng.standard_normal((3, 1)) * 10).astype(np.float32)
        s = 0.0
        for k in range(3):
            s += float(v[k, 0]) * float(v[k, 0])
        assert np.sqrt(s) == cv2.norm(v)


def _three_maxima(sz):
    """ORBmatcher::ComputeThreeMaxima (src/ORBmatcher.cc:1800-1843), line by line"""
    max1 = max2 = max3 = 0
    i1 = i2 = i3 = -1
    for i, s in enumerate(sz):
        if s > max1:
            max3, max2, max1 = max2, max1, s
            i3, i2, i1 = i2, i1, i
        elif s > max2:
            max3, max2 = max2, s
            i3, i2 = i2, i
        elif s > max3:
            max3, i3 = s, i
    if max2 < np.float32(0.1) * np.float32(max1):
        i2 = i3 = -1
    elif max3 < np.float32(0.1) * np.float32(max1):
        i3 = -1
    return i1, i2, i3


def test_three_maxima(oracle_built):
    lib = oracle_built.Oracle.load()
    rng = np.random.default_rng(5)
    for trial in range(2000):
        kind = trial % 4
        sz = rng.integers(0, [3, 40, 400, 10][kind], 30).astype(np.int32)
        if kind == 3:
            sz[rng.integers(0, 30)] = 100  # the 10 % rule
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        lib.mo_three_maxima(sz.ctypes.data_as(C.c_void_p), 30, C.byref(a), C.byref(b), C.byref(c))
        assert (a.value, b.value, c.value) == _three_maxima(sz.tolist())


def test_frame_against_itself(oracle_built):
    """Current = last, same pose, exact descriptors, every map point observed: each map point projects onto
    its own keypoint at distance 0, so the slots hold their own index wherever that keypoint is the first
    zero-distance candidate, every accepted query is counted and nothing is rotated away (all in bin 0)."""
    cur, last, cam, bounds, scale = matchcase.make_case(seed=21, obs_rate=1.0, outlier_rate=0.0)
    me = dict(last)
    me["mp_desc"] = last["desc"].copy()
    me_cur = {k: last[k] for k in ("kps", "desc", "un_xy", "u_right", "cell_start", "cell_idx", "Tcw")}
    m, nm = oracle_built.search_by_projection(me_cur, me, cam, last["bounds"], scale, 7.0)
    has = (last["mp_flags"] & 1).astype(bool)
    own = m == np.arange(len(m))
    assert own[has].mean() > 0.97 and not (m[~has] >= 0).any()
    assert nm == int(has.sum())


def test_outliers_and_missing_points_are_skipped(oracle_built):
    cur, last, cam, bounds, scale = matchcase.make_case(seed=22)
    m, nm = oracle_built.search_by_projection(cur, last, cam, bounds, scale, 15.0)
    used = set(m[m >= 0].tolist())
    fl = last["mp_flags"]
    assert all((fl[i] & 1) and not (fl[i] & 4) for i in used)
    assert nm >= len(used) > 500
