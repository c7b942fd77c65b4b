"""The C-ABI boundary (include/borb.h): the built library loads on a machine without a GPU, exports
every symbol the header declares, and its compute entry points FAIL LOUDLY there (no CPU path)."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
HEADER = (ROOT / "include" / "borb.h").read_text()


def declared_symbols():
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    return sorted(set(re.findall(r"\b(borb_[a-z0-9_]+)\s*\(", body)))


def test_header_declares_the_expected_surface():
    syms = declared_symbols()
    for must in ("borb_create", "borb_destroy", "borb_extract_host", "borb_result", "borb_pyramid",
                 "borb_table", "borb_levels", "borb_scale_factor", "borb_last_error"):
        assert must in syms


def test_library_exports_every_declared_symbol():
    from orb_slam2_test_b200 import extractor
    lib = extractor.load_library()
    for s in declared_symbols():
        assert hasattr(lib, s), f"liborb_b200.so does not export {s}"
    assert sorted(extractor.SYMBOLS) == declared_symbols(), "extractor.SYMBOLS out of date with the header"


def test_no_torch_or_opencv_types_in_the_abi():
    body = re.sub(r"/\*.*?\*/", "", HEADER, flags=re.S)
    for bad in ("torch", "at::", "cv::", "cudaStream", "std::"):
        assert bad not in body


def test_keypoint_record_is_cv_keypoint():
    from orb_slam2_test_b200.extractor import KP_DTYPE
    assert KP_DTYPE.itemsize == 28
    assert [KP_DTYPE.fields[n][1] for n in KP_DTYPE.names] == [0, 4, 8, 12, 16, 20, 24]


def test_no_cpu_fallback_without_a_device():
    """Without a CUDA device borb_create must fail with BORB_ERR_CUDA, never compute on the host."""
    from conftest import HAVE_GPU
    if HAVE_GPU:
        pytest.skip("a GPU is present")
    from orb_slam2_test_b200 import BorbError, ORBextractor
    with pytest.raises(BorbError) as ei:
        ORBextractor(1000, 1.2, 8, 20, 7)
    assert ei.value.code == -2 and "no CPU path" in str(ei.value)


def test_bad_arguments_are_rejected_before_touching_the_device():
    from orb_slam2_test_b200 import extractor
    lib = extractor.load_library()
    h = C.c_void_p()
    for args in ((0, 1.2, 8, 20, 7), (1000, 1.0, 8, 20, 7), (1000, 1.2, 0, 20, 7), (1000, 1.2, 99, 20, 7),
                 (1000, 1.2, 8, 0, 7), (1000, 1.2, 8, 20, 255)):
        rc = lib.borb_create(*args, -1, 1, C.byref(h))
        assert rc == -1, args
        assert not h.value
    assert lib.borb_create(1000, 1.2, 8, 20, 7, -1, 1, None) == -1
    assert lib.borb_levels(None) == 0


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under the package may import, link or dlopen it."""
    pkg = ROOT / "orb_slam2_test_b200"
    for p in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cu")) + list(pkg.rglob("*.cuh")) + list(pkg.rglob("*.cc")) + \
            list(pkg.rglob("*.h")) + list(pkg.rglob("Makefile")):
        txt = p.read_text()
        assert "pyoracle" not in txt and "liborb_oracle" not in txt and "libreforb" not in txt, p
        assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), p
        assert not re.search(r'#include\s+"[^"]*oracle/', txt), p


def test_host_fast_score_tap_matches_oracle_fast(oracle_built):
    """borb_debug_fast_score evaluates on the host the very function the FAST kernel runs per pixel;
    pin it to the oracle's cv::FAST restatement (score > 0 <=> corner at the threshold, same score)."""
    from orb_slam2_test_b200 import extractor
    lib = extractor.load_library()
    L = oracle_built.Oracle.load()

    class KP(C.Structure):
        _fields_ = [("x", C.c_int), ("y", C.c_int), ("s", C.c_int)]
    u8p = C.POINTER(C.c_uint8)
    L.cvp_fast9_16.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(KP), C.c_int]
    L.cvp_fast_corner_score.argtypes = [u8p, C.c_size_t, C.c_int]
    L.cvp_fast_corner_score.restype = C.c_int
    rng = np.random.default_rng(3)
    base = (np.add.outer(np.arange(64) // 5, np.arange(64) // 6) % 2 * 70 + 80)
    img = np.clip(base + rng.integers(-14, 15, (64, 64)), 0, 255).astype(np.uint8)
    n_corner = 0
    for t in (7, 20):
        for y in range(3, 61):
            for x in range(3, 61):
                p = img.ctypes.data + y * 64 + x
                s = lib.borb_debug_fast_score(p, 64, t)
                if s:
                    n_corner += 1
                    assert s >= t
                    assert s == L.cvp_fast_corner_score(C.cast(p, u8p), 64, t), (x, y, t)
        # the set of corners agrees with cv::FAST without NMS?  cvp_fast9_16 applies NMS, so check
        # its survivors are corners with the same score
        out = (KP * 4096)()
        n = L.cvp_fast9_16(img.ctypes.data_as(u8p), 64, 64, 64, t, out, 4096)
        assert n > 0
        for i in range(n):
            p = img.ctypes.data + out[i].y * 64 + out[i].x
            assert lib.borb_debug_fast_score(p, 64, t) == out[i].s
    assert n_corner > 100
