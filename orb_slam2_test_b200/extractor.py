"""Python host mirror of the reference's ORBextractor interface over the C ABI (include/borb.h).

The reference class (include/ORBextractor.h:46-144 of the reference tree) is C++; the C++ drop-in
lives in orb_slam2_test_b200/shim/.  This module is the same surface for Python callers, tests and
bench.py: same constructor arguments, same getter names, `__call__(image, mask)` in place of
`operator()`, `mvImagePyramid` as a list of level views.  Everything runs in liborb_b200.so on the
GPU -- there is no CPU path, and importing this on a machine without the built library raises.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
# BORB_LIB: alternative build of the same library (kernel A/B experiments); default = in-tree build
LIB_PATH = Path(os.environ["BORB_LIB"]) if os.environ.get("BORB_LIB") else _HERE / "liborb_b200.so"

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

BORB_WANT_PYRAMID = 1

_vp = C.c_void_p
_lib = None

# every symbol include/borb.h declares (tests check the list against the header)
SYMBOLS = [
    "borb_create", "borb_destroy", "borb_last_error", "borb_version", "borb_levels",
    "borb_scale_factor", "borb_table", "borb_features_per_level", "borb_umax",
    "borb_extract_host", "borb_extract_device", "borb_fetch", "borb_sync", "borb_result",
    "borb_pyramid", "borb_level_size", "borb_launches_per_call", "borb_last_device_ms",
    "borb_debug_candidates", "borb_debug_blurred", "borb_debug_sincos", "borb_debug_fast_score", "borb_debug_describe", "borb_debug_set_match_rounds",
    "borb_submit_host", "borb_wait", "borb_host_alloc", "borb_host_free", "borb_timer_start",
    "borb_timer_stop", "borb_timer_ms", "borb_stage_times", "borb_set_stage_timing", "borb_last_stage_ms", "borb_counts",
    "borb_stereo_match", "borb_stereo_result", "borb_set_input_format", "borb_set_rectify_map",
    "borb_last_prestep_ms", "borb_set_camera", "borb_frame_grid",
    "borb_search_by_projection", "borb_search_local_points", "borb_search_for_initialization", "borb_match_rounds", "borb_max_keypoints",
]
# the pipelined sequence feeder (bound in stream.py)
STREAM_SYMBOLS = [
    "borb_stream_open", "borb_stream_close", "borb_stream_set_stereo", "borb_stream_set_camera",
    "borb_stream_input", "borb_stream_push", "borb_stream_pop", "borb_stream_release",
    "borb_stream_in_flight", "borb_stream_slots_free", "borb_stream_result", "borb_stream_stereo_result",
    "borb_stream_frame_grid", "borb_stream_pyramid", "borb_stream_link_bytes",
    "borb_stream_launches_per_push", "borb_stream_start", "borb_stream_join",
]
SYMBOLS += STREAM_SYMBOLS
# allocation-free variant of the per-keypoint step (no counterpart in the reference; see include/borb.h)
SYMBOLS += ["borb_undistort_grid_noalloc"]
# bag of words: Frame::ComputeBoW and ORBmatcher::SearchByBoW(KeyFrame*, Frame&)
SYMBOLS += ["borb_voc_load_text", "borb_voc_create", "borb_voc_destroy", "borb_voc_info", "borb_voc_transform",
            "borb_compute_bow", "borb_bow_result", "borb_search_by_bow"]
# the relocalisation matcher ORBmatcher::SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) and its logf taps
SYMBOLS += ["borb_search_by_projection_keyframe", "borb_debug_logf", "borb_debug_logf_device"]
# ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)
SYMBOLS += ["borb_voc_search_by_bow"]
# ORBmatcher::SearchForTriangulation
SYMBOLS += ["borb_voc_search_for_triangulation"]
# keyframes on the device and the matchers that take one: ORBmatcher::Fuse(pKF, vpMapPoints, th)
SYMBOLS += ["borb_load_frame", "borb_fuse_search"]
# the loop-closing matchers: Fuse(pKF, Scw, ...), SearchByProjection(pKF, Scw, ...), SearchBySim3
SYMBOLS += ["borb_project_search", "borb_search_by_projection_sim3", "borb_search_by_sim3"]


class BorbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"borb status {code}: {msg}")
        self.code = code


def load_library() -> C.CDLL:
    """dlopen liborb_b200.so (in-tree build).  Raises if it has not been built: no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise FileNotFoundError(
            f"{LIB_PATH} is missing: build it with `make -C {_HERE / 'csrc'}` "
            "(or __graft_entry__.build()); this package has no CPU fallback")
    lib = C.CDLL(str(LIB_PATH))
    lib.borb_last_error.restype = C.c_char_p
    lib.borb_version.restype = C.c_char_p
    lib.borb_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                C.POINTER(_vp)]
    lib.borb_destroy.argtypes = [_vp]
    lib.borb_levels.argtypes = [_vp]
    lib.borb_scale_factor.argtypes = [_vp]
    lib.borb_scale_factor.restype = C.c_float
    lib.borb_table.argtypes = [_vp, C.c_int, _vp]
    lib.borb_features_per_level.argtypes = [_vp, _vp]
    lib.borb_umax.argtypes = [_vp, _vp]
    lib.borb_extract_host.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t,
                                      C.c_uint]
    lib.borb_extract_device.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t]
    lib.borb_fetch.argtypes = [_vp, C.c_uint]
    lib.borb_sync.argtypes = [_vp]
    lib.borb_result.argtypes = [_vp, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_int)]
    lib.borb_pyramid.argtypes = [_vp, C.c_int, C.c_int, C.POINTER(_vp), C.POINTER(C.c_int),
                                 C.POINTER(C.c_int), C.POINTER(C.c_size_t)]
    lib.borb_level_size.argtypes = [_vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.borb_launches_per_call.argtypes = [_vp]
    lib.borb_last_device_ms.argtypes = [_vp]
    lib.borb_last_device_ms.restype = C.c_float
    lib.borb_debug_candidates.argtypes = [_vp, C.c_int, C.c_int, _vp, _vp, _vp, C.c_int,
                                          C.POINTER(C.c_int)]
    lib.borb_debug_blurred.argtypes = [_vp, C.c_int, C.c_int, _vp]
    lib.borb_debug_sincos.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.borb_debug_set_match_rounds.argtypes = [C.c_int]
    lib.borb_debug_set_match_rounds.restype = None
    lib.borb_debug_describe.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int, C.c_float, _vp]
    lib.borb_debug_fast_score.argtypes = [_vp, C.c_int, C.c_int]
    lib.borb_submit_host.argtypes = lib.borb_extract_host.argtypes
    lib.borb_wait.argtypes = [_vp]
    lib.borb_host_alloc.argtypes = [C.c_size_t]
    lib.borb_host_alloc.restype = _vp
    lib.borb_host_free.argtypes = [_vp]
    lib.borb_timer_start.argtypes = [_vp]
    lib.borb_timer_stop.argtypes = [_vp]
    lib.borb_timer_ms.argtypes = [_vp]
    lib.borb_timer_ms.restype = C.c_float
    lib.borb_stage_times.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, _vp]
    lib.borb_counts.argtypes = [_vp, _vp]
    lib.borb_set_stage_timing.argtypes = [_vp, C.c_int]
    lib.borb_last_stage_ms.argtypes = [_vp, _vp]
    lib.borb_set_camera.argtypes = [_vp, _vp, _vp, C.c_int]
    lib.borb_frame_grid.argtypes = [_vp, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _vp]
    lib.borb_last_prestep_ms.argtypes = [_vp]
    lib.borb_last_prestep_ms.restype = C.c_float
    lib.borb_set_input_format.argtypes = [_vp, C.c_int]
    lib.borb_set_rectify_map.argtypes = [_vp, _vp, _vp, C.c_int, C.c_int, C.c_size_t]
    lib.borb_stereo_match.argtypes = [_vp, _vp, C.c_float, C.c_float, C.c_uint]
    lib.borb_search_by_projection.argtypes = [_vp, C.c_int, _vp, _vp, C.c_float, C.c_float, C.c_float, C.c_uint, _vp,
                                              C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.borb_search_local_points.argtypes = [_vp, C.c_int, _vp, _vp, C.c_float, C.c_float, C.c_uint, _vp,
                                             C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.borb_search_for_initialization.argtypes = [_vp, C.c_int, _vp, _vp, C.c_int, C.c_float, C.c_uint, _vp, C.POINTER(C.c_int)]
    lib.borb_search_by_projection_keyframe.argtypes = [_vp, C.c_int, _vp, _vp, _vp, C.c_float, C.c_int, C.c_uint, _vp,
                                                       C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.borb_voc_search_by_bow.argtypes = [_vp, _vp, _vp, C.c_float, C.c_uint, _vp, C.POINTER(C.c_int)]
    lib.borb_voc_search_for_triangulation.argtypes = [_vp, _vp, _vp, _vp, C.c_uint, _vp, C.POINTER(C.c_int)]
    lib.borb_load_frame.argtypes = [_vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int]
    lib.borb_fuse_search.argtypes = [_vp, C.c_int, _vp, _vp, _vp, C.c_float, C.c_float, C.c_uint, _vp, C.POINTER(C.c_int)]
    lib.borb_project_search.argtypes = [_vp, C.c_int, _vp, _vp, C.c_float, C.c_int, _vp, C.POINTER(C.c_int)]
    lib.borb_search_by_projection_sim3.argtypes = [_vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.borb_search_by_sim3.argtypes = [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, C.c_float, _vp, _vp, C.POINTER(C.c_int)]
    lib.borb_debug_logf.argtypes = [_vp, _vp, C.c_int]
    lib.borb_debug_logf.restype = None
    lib.borb_debug_logf_device.argtypes = [_vp, _vp, C.c_int]
    lib.borb_match_rounds.argtypes = [_vp]
    lib.borb_max_keypoints.argtypes = [_vp]
    lib.borb_stereo_result.argtypes = [_vp, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_int),
                                       C.POINTER(C.c_int)]
    lib.borb_undistort_grid_noalloc.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp]
    lib.borb_voc_load_text.argtypes = [C.c_char_p, C.c_int, C.POINTER(_vp)]
    lib.borb_voc_create.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, _vp, _vp, C.c_int, C.POINTER(_vp)]
    lib.borb_voc_destroy.argtypes = [_vp]
    lib.borb_voc_info.argtypes = [_vp] + [C.POINTER(C.c_int)] * 6
    lib.borb_voc_transform.argtypes = [_vp, _vp, C.c_int, C.c_int, _vp, _vp, C.POINTER(C.c_int), _vp, _vp, _vp, C.POINTER(C.c_int)]
    lib.borb_compute_bow.argtypes = [_vp, C.c_int, _vp, C.c_int]
    lib.borb_bow_result.argtypes = [_vp, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_int), C.POINTER(_vp),
                                    C.POINTER(_vp), C.POINTER(_vp), C.POINTER(C.c_int)]
    lib.borb_search_by_bow.argtypes = [_vp, C.c_int, _vp, C.c_float, C.c_uint, _vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    _lib = lib
    return lib


def _check(lib, rc: int) -> None:
    if rc != 0:
        raise BorbError(rc, lib.borb_last_error().decode())


def debug_sincos(radians: float):
    lib = load_library()
    s, c = C.c_float(), C.c_float()
    lib.borb_debug_sincos(radians, C.byref(s), C.byref(c))
    return s.value, c.value


def debug_logf(x: np.ndarray, device: bool = False) -> np.ndarray:
    """csrc/orb_logf.cuh (glibc's logf restated): the host copy, or the device copy through one small kernel"""
    lib = load_library()
    x = np.ascontiguousarray(x, np.float32)
    y = np.empty_like(x)
    if device:
        _check(lib, lib.borb_debug_logf_device(x.ctypes.data, y.ctypes.data, len(x)))
    else:
        lib.borb_debug_logf(x.ctypes.data, y.ctypes.data, len(x))
    return y


class _InitFrame(C.Structure):  # == borb_init_frame
    _fields_ = [("n", C.c_int), ("octave", _vp), ("angle", _vp), ("desc", _vp)]


class _MapPoints(C.Structure):  # == borb_map_points
    _fields_ = [("n", C.c_int), ("proj", _vp), ("desc", _vp), ("flags", _vp), ("level", _vp)]


class _LastFrame(C.Structure):  # == borb_last_frame
    _fields_ = [("n", C.c_int), ("xyz", _vp), ("desc", _vp), ("flags", _vp), ("octave", _vp), ("angle", _vp),
                ("Tcw", _vp)]


class _Projection(C.Structure):  # == borb_projection
    _fields_ = [("T1", _vp), ("T2", _vp), ("Ow", _vp)]


def _fuse_points(pts: dict, keep: list) -> "_FusePoints":
    a = [np.ascontiguousarray(pts["xyz"], np.float32), np.ascontiguousarray(pts["desc"], np.uint8),
         np.ascontiguousarray(pts["flags"], np.uint8), np.ascontiguousarray(pts["dist"], np.float32)]
    nrm = np.ascontiguousarray(pts["normal"], np.float32) if pts.get("normal") is not None else None
    keep += a + [nrm]
    return _FusePoints(len(a[2]), a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, a[3].ctypes.data,
                       nrm.ctypes.data if nrm is not None else None)


def _projection(T1, T2, Ow, keep: list) -> "_Projection":
    a = [np.ascontiguousarray(x, np.float32) if x is not None else None for x in (T1, T2, Ow)]
    keep += a
    return _Projection(*[x.ctypes.data if x is not None else None for x in a])


class _FusePoints(C.Structure):  # == borb_fuse_points
    _fields_ = [("n", C.c_int), ("xyz", _vp), ("desc", _vp), ("flags", _vp), ("dist", _vp), ("normal", _vp)]


class _KeyFramePoints(C.Structure):  # == borb_keyframe_points
    _fields_ = [("n", C.c_int), ("xyz", _vp), ("desc", _vp), ("flags", _vp), ("dist", _vp), ("angle", _vp)]


class _KeyFrame(C.Structure):  # == borb_keyframe
    _fields_ = [("n", C.c_int), ("desc", _vp), ("angle", _vp), ("mp_flags", _vp), ("n_nodes", C.c_int), ("node_id", _vp),
                ("node_start", _vp), ("feat_idx", _vp)]


class _Triangulation(C.Structure):  # == borb_triangulation
    _fields_ = [("F12", _vp), ("Ow1", _vp), ("Tcw2", _vp), ("fx2", C.c_float), ("fy2", C.c_float), ("cx2", C.c_float),
                ("cy2", C.c_float), ("xy1", _vp), ("xy2", _vp), ("octave2", _vp), ("u_right1", _vp), ("u_right2", _vp),
                ("nlevels", C.c_int), ("scale_factors2", _vp), ("level_sigma2_2", _vp)]


class ORBVocabulary:
    """ORB_SLAM2::ORBVocabulary (include/ORBVocabulary.h = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB>) on the
    device: loadFromTextFile's format (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1338-1421) or the tree as arrays."""

    def __init__(self, path: str | None = None, device: int = -1, arrays=None):
        self._lib = load_library()
        self._h = _vp()
        if path is not None:
            _check(self._lib, self._lib.borb_voc_load_text(str(path).encode(), device, C.byref(self._h)))
        else:
            k, L, scoring, weighting, parent, leaf, desc, weight = arrays
            keep = [np.ascontiguousarray(parent, np.int32), np.ascontiguousarray(leaf, np.uint8),
                    np.ascontiguousarray(desc, np.uint8), np.ascontiguousarray(weight, np.float64)]
            _check(self._lib, self._lib.borb_voc_create(k, L, scoring, weighting, len(keep[0]), *[a.ctypes.data for a in keep],
                                                        device, C.byref(self._h)))
        v = [C.c_int() for _ in range(6)]
        _check(self._lib, self._lib.borb_voc_info(self._h, *[C.byref(x) for x in v]))
        self.k, self.L, self.n_nodes, self.n_words, self.scoring, self.weighting = [x.value for x in v]

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.borb_voc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def transform(self, desc: np.ndarray, levelsup: int = 4) -> dict:
        """transform(features, BowVector, FeatureVector, levelsup) for descriptors in host memory
        (KeyFrame::ComputeBoW).  -> dict word_id, word_val, node_id, node_start, feat_idx (see include/borb.h)."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(d)
        wid, wv = np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64)
        nid, ns, fi = np.zeros(max(n, 1), np.uint32), np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.uint32)
        nw, nn = C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_voc_transform(self._h, d.ctypes.data, n, levelsup, wid.ctypes.data, wv.ctypes.data,
                                                       C.byref(nw), nid.ctypes.data, ns.ctypes.data, fi.ctypes.data, C.byref(nn)))
        return {"word_id": wid[:nw.value].copy(), "word_val": wv[:nw.value].copy(), "node_id": nid[:nn.value].copy(),
                "node_start": ns[:nn.value + 1].copy(), "feat_idx": fi[:ns[nn.value]].copy()}


    def search_by_bow(self, kf1: dict, kf2: dict, nnratio: float = 0.75, check_orientation: bool = True):
        """ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (src/ORBmatcher.cc:634-777) on the device.  kf1 / kf2: dicts
        as ORBextractor.search_by_bow takes (desc, angle, mp_flags, fv).
        -> (match12 [n1] i32: pKF2 keypoint index or -1, nmatches)."""
        keep, recs = [], []
        for kf in (kf1, kf2):
            fv = kf["fv"]
            a = [np.ascontiguousarray(kf["desc"], np.uint8), np.ascontiguousarray(kf["angle"], np.float32),
                 np.ascontiguousarray(kf["mp_flags"], np.uint8), np.ascontiguousarray(fv["node_id"], np.uint32),
                 np.ascontiguousarray(fv["node_start"], np.int32), np.ascontiguousarray(fv["feat_idx"], np.uint32)]
            keep.append(a)
            recs.append(_KeyFrame(len(a[2]), a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data, len(a[3]),
                                  a[3].ctypes.data, a[4].ctypes.data, a[5].ctypes.data))
        out = np.full(max(recs[0].n, 1), -1, np.int32)
        nm = C.c_int()
        _check(self._lib, self._lib.borb_voc_search_by_bow(self._h, C.addressof(recs[0]), C.addressof(recs[1]), nnratio,
                                                           0 if check_orientation else 2, out.ctypes.data, C.byref(nm)))
        return out[:recs[0].n].copy(), nm.value


    def search_for_triangulation(self, kf1: dict, kf2: dict, geo: dict, only_stereo: bool = False,
                                 check_orientation: bool = True):
        """ORBmatcher::SearchForTriangulation(pKF1, pKF2, F12, vMatchedPairs, bOnlyStereo) (src/ORBmatcher.cc:779-966) on the
        device.  kf1 / kf2: dicts with desc, angle, mp_flags, fv (as search_by_bow) and xy [n,2] f32 (mvKeysUn positions),
        u_right [n] f32 or None; kf2 also octave [n] i32.  geo: F12 [3,3], Ow1 [3], Tcw2 [3,4], cam2 (fx, fy, cx, cy),
        scale2, sigma2_2.  -> (match12 [n1] i32: pKF2 keypoint or -1, nmatches); vMatchedPairs = the (i1, match12[i1]) >= 0."""
        keep, recs = [], []

        def arr(a, dt):
            a = np.ascontiguousarray(a, dt)
            keep.append(a)
            return a.ctypes.data
        for kf in (kf1, kf2):
            fv = kf["fv"]
            recs.append(_KeyFrame(len(kf["mp_flags"]), arr(kf["desc"], np.uint8), arr(kf["angle"], np.float32),
                                  arr(kf["mp_flags"], np.uint8), len(fv["node_id"]), arr(fv["node_id"], np.uint32),
                                  arr(fv["node_start"], np.int32), arr(fv["feat_idx"], np.uint32)))
        cam = [float(x) for x in geo["cam2"]]
        g = _Triangulation(arr(geo["F12"], np.float32), arr(geo["Ow1"], np.float32), arr(geo["Tcw2"], np.float32), cam[0], cam[1],
                           cam[2], cam[3], arr(kf1["xy"], np.float32), arr(kf2["xy"], np.float32), arr(kf2["octave"], np.int32),
                           arr(kf1["u_right"], np.float32) if kf1.get("u_right") is not None else None,
                           arr(kf2["u_right"], np.float32) if kf2.get("u_right") is not None else None,
                           len(geo["scale2"]), arr(geo["scale2"], np.float32), arr(geo["sigma2_2"], np.float32))
        out = np.full(max(recs[0].n, 1), -1, np.int32)
        nm = C.c_int()
        _check(self._lib, self._lib.borb_voc_search_for_triangulation(
            self._h, C.addressof(recs[0]), C.addressof(recs[1]), C.addressof(g),
            (0 if check_orientation else 2) | (8 if only_stereo else 0), out.ctypes.data, C.byref(nm)))
        return out[:recs[0].n].copy(), nm.value


class ORBextractor:
    """Same constructor arguments and accessor names as ORB_SLAM2::ORBextractor.

    `extractor(image)` returns (keypoints, descriptors): a structured array with cv::KeyPoint's
    fields and an [n, 32] uint8 array, in the reference's output order.  `max_batch` > 1 enables
    `extract_batch` (frames of one call go through the kernels together).
    """

    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures: int, scaleFactor: float, nlevels: int, iniThFAST: int,
                 minThFAST: int, device: int = -1, max_batch: int = 1, want_pyramid: bool = True):
        self._lib = load_library()
        self._h = _vp()
        _check(self._lib, self._lib.borb_create(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST,
                                                device, max_batch, C.byref(self._h)))
        self.nlevels = nlevels
        self.max_batch = max_batch
        self.want_pyramid = want_pyramid
        self.mvImagePyramid: list[np.ndarray] = []
        self._n = 0
        self._views = {}
        self._ch = 1

    # -- the step before the extractor: cvtColor (src/Tracking.cc:176-201) and cv::remap
    #    (Examples/Stereo/stereo_euroc.cc:136-137) on the device --
    FORMATS = {"gray": (0, 1), "bgr": (1, 3), "rgb": (2, 3), "bgra": (3, 4), "rgba": (4, 4)}

    def set_input_format(self, fmt: str) -> None:
        code, ch = self.FORMATS[fmt]
        _check(self._lib, self._lib.borb_set_input_format(self._h, code))
        self._ch = ch

    def set_rectify_map(self, map_x: np.ndarray | None, map_y: np.ndarray | None) -> None:
        """CV_32F maps of cv::initUndistortRectifyMap; their shape is the rectified image's."""
        if map_x is None and map_y is None:
            _check(self._lib, self._lib.borb_set_rectify_map(self._h, None, None, 0, 0, 0))
            return
        map_x = np.ascontiguousarray(map_x, np.float32)
        map_y = np.ascontiguousarray(map_y, np.float32)
        assert map_x.shape == map_y.shape and map_x.ndim == 2
        rows, cols = map_x.shape
        _check(self._lib, self._lib.borb_set_rectify_map(self._h, map_x.ctypes.data, map_y.ctypes.data, cols,
                                                         rows, cols))

    # -- the per-keypoint step after the extractor (src/Frame.cc:542-572, :574-609, :292-307) --
    def set_camera(self, K=None, dist=()) -> None:
        """K = (fx, fy, cx, cy), dist = (k1, k2, p1, p2[, k3]); None turns the step off."""
        if K is None:
            _check(self._lib, self._lib.borb_set_camera(self._h, None, None, 0))
            return
        k = np.asarray(K, np.float32)
        d = np.asarray(dist, np.float32)
        _check(self._lib, self._lib.borb_set_camera(self._h, k.ctypes.data, d.ctypes.data if len(d) else None, len(d)))

    def frame_grid(self, frame: int = 0):
        """(mvKeysUn positions [n,2], cell_start [3073], cell_idx [n_in_grid], bounds [4])."""
        xy, cs, ci = _vp(), _vp(), _vp()
        bounds = np.zeros(4, np.float32)
        _check(self._lib, self._lib.borb_frame_grid(self._h, frame, C.byref(xy), C.byref(cs), C.byref(ci),
                                                    bounds.ctypes.data))
        n = self.count(frame)
        start = np.ctypeslib.as_array(C.cast(cs, C.POINTER(C.c_int32)), shape=(64 * 48 + 1,)).copy()
        un = (np.ctypeslib.as_array(C.cast(xy, C.POINTER(C.c_float)), shape=(n, 2)).copy() if n
              else np.zeros((0, 2), np.float32))
        m = int(start[-1])
        idx = (np.ctypeslib.as_array(C.cast(ci, C.POINTER(C.c_int32)), shape=(m,)).copy() if m
               else np.zeros(0, np.int32))
        return un, start, idx, bounds

    def last_prestep_ms(self) -> float:
        return self._lib.borb_last_prestep_ms(self._h)

    def _check_image(self, shape, ndim_gray: int):
        if self._ch == 1:
            if len(shape) != ndim_gray:
                raise TypeError("image must be CV_8UC1 (2-D uint8)")  # reference: assert, :1337
        elif len(shape) != ndim_gray + 1 or shape[-1] != self._ch:
            raise TypeError(f"image must have {self._ch} interleaved channels")

    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.borb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- getters (include/ORBextractor.h:78-98) --
    def GetLevels(self) -> int:
        return self._lib.borb_levels(self._h)

    def GetScaleFactor(self) -> float:
        return self._lib.borb_scale_factor(self._h)

    def _table(self, which: int) -> np.ndarray:
        out = np.zeros(self.nlevels, np.float32)
        _check(self._lib, self._lib.borb_table(self._h, which, out.ctypes.data))
        return out

    def GetScaleFactors(self):
        return self._table(0)

    def GetInverseScaleFactors(self):
        return self._table(1)

    def GetScaleSigmaSquares(self):
        return self._table(2)

    def GetInverseScaleSigmaSquares(self):
        return self._table(3)

    def features_per_level(self) -> np.ndarray:
        out = np.zeros(self.nlevels, np.int32)
        _check(self._lib, self._lib.borb_features_per_level(self._h, out.ctypes.data))
        return out

    def umax(self) -> np.ndarray:
        out = np.zeros(16, np.int32)
        _check(self._lib, self._lib.borb_umax(self._h, out.ctypes.data))
        return out

    # -- operator() (src/ORBextractor.cc:1330-1397) --
    def __call__(self, image: np.ndarray, mask=None):
        """mask is ignored, as in the reference (include/ORBextractor.h:72)."""
        if image is None or image.size == 0:
            return None  # the reference returns without touching its outputs (:1333-1334)
        if image.dtype != np.uint8:
            raise TypeError("image must be CV_8UC1 (2-D uint8)")  # reference: assert, :1337
        self._check_image(image.shape, 2)
        if image.strides[-1] != 1 or (image.ndim == 3 and image.strides[1] != self._ch):
            image = np.ascontiguousarray(image)
        rows, cols = image.shape[:2]
        flags = BORB_WANT_PYRAMID if self.want_pyramid else 0
        _check(self._lib, self._lib.borb_extract_host(self._h, image.ctypes.data, 1, cols, rows,
                                                      image.strides[0], 0, flags))
        self._n = 1
        if self.want_pyramid:
            self.mvImagePyramid = [self.pyramid_level(0, l) for l in range(self.nlevels)]
        return self.result(0)

    def extract_batch(self, images: np.ndarray, want_pyramid: bool = False):
        """images: [n, rows, cols] (or [n, rows, cols, channels]) uint8, C-contiguous.  Returns a list
        of (kps, desc)."""
        assert images.dtype == np.uint8 and images.flags.c_contiguous
        self._check_image(images.shape, 3)
        n, rows, cols = images.shape[:3]
        _check(self._lib, self._lib.borb_extract_host(self._h, images.ctypes.data, n, cols, rows, cols * self._ch,
                                                      rows * cols * self._ch,
                                                      BORB_WANT_PYRAMID if want_pyramid else 0))
        self._n = n
        return [self.result(f) for f in range(n)]

    def extract_device(self, dev_ptr: int, n: int, cols: int, rows: int, step: int | None = None,
                       frame_stride: int | None = None) -> None:
        """Frames already in device memory (e.g. tensor.data_ptr()); asynchronous."""
        step = cols * self._ch if step is None else step
        frame_stride = rows * step if frame_stride is None else frame_stride
        _check(self._lib, self._lib.borb_extract_device(self._h, dev_ptr, n, cols, rows, step, frame_stride))
        self._n = n

    def fetch(self, want_pyramid: bool = False):
        _check(self._lib, self._lib.borb_fetch(self._h, BORB_WANT_PYRAMID if want_pyramid else 0))
        return [self.result(f) for f in range(self._n)]

    def submit_batch(self, images: np.ndarray, want_pyramid: bool = False) -> None:
        """Asynchronous half of extract_batch; pair with wait()."""
        assert images.dtype == np.uint8 and images.flags.c_contiguous
        self._check_image(images.shape, 3)
        n, rows, cols = images.shape[:3]
        _check(self._lib, self._lib.borb_submit_host(self._h, images.ctypes.data, n, cols, rows, cols * self._ch,
                                                     rows * cols * self._ch,
                                                     BORB_WANT_PYRAMID if want_pyramid else 0))
        self._n = n

    def wait(self) -> None:
        _check(self._lib, self._lib.borb_wait(self._h))

    def timer_start(self) -> None:
        _check(self._lib, self._lib.borb_timer_start(self._h))

    def timer_stop(self) -> None:
        _check(self._lib, self._lib.borb_timer_stop(self._h))

    def timer_ms(self) -> float:
        return self._lib.borb_timer_ms(self._h)

    def stage_times(self, dev_ptr: int, n: int, cols: int, rows: int) -> dict:
        ms = (C.c_float * 5)()
        _check(self._lib, self._lib.borb_stage_times(self._h, dev_ptr, n, cols, rows, cols, rows * cols, ms))
        return dict(zip(("pyramid", "fast", "quadtree", "blur", "describe"), [float(v) for v in ms]))

    STAGES = ("pyramid", "fast", "quadtree", "blur", "describe")

    def set_stage_timing(self, on: bool) -> None:
        _check(self._lib, self._lib.borb_set_stage_timing(self._h, 1 if on else 0))

    def last_stage_ms(self) -> dict:
        ms = (C.c_float * 5)()
        _check(self._lib, self._lib.borb_last_stage_ms(self._h, ms))
        return dict(zip(self.STAGES, [float(v) for v in ms]))

    def counts(self) -> np.ndarray:
        out = np.zeros(self._n, np.int32)
        _check(self._lib, self._lib.borb_counts(self._h, out.ctypes.data))
        return out

    def count(self, frame: int = 0) -> int:
        n = C.c_int()
        _check(self._lib, self._lib.borb_result(self._h, frame, None, None, C.byref(n)))
        return n.value

    def sync(self) -> None:
        _check(self._lib, self._lib.borb_sync(self._h))

    def last_device_ms(self) -> float:
        return self._lib.borb_last_device_ms(self._h)

    def launches_per_call(self) -> int:
        return self._lib.borb_launches_per_call(self._h)

    def result(self, frame: int = 0, copy: bool = True):
        kp, de, n = _vp(), _vp(), C.c_int()
        _check(self._lib, self._lib.borb_result(self._h, frame, C.byref(kp), C.byref(de), C.byref(n)))
        cnt = n.value
        if cnt == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        # numpy views over the library's pinned result slots, made once per (slot address, capacity): building
        # them per call through np.ctypeslib costs more than the call's kernels
        K = self._lib.borb_max_keypoints(self._h)
        key = (kp.value, de.value, K)
        v = self._views.get(key)
        if v is None:
            if len(self._views) > 4096:
                self._views.clear()
            kb = (C.c_uint8 * (K * 28)).from_address(kp.value)
            db = (C.c_uint8 * (K * 32)).from_address(de.value)
            v = self._views[key] = (np.frombuffer(kb, dtype=KP_DTYPE), np.frombuffer(db, dtype=np.uint8).reshape(K, 32))
        kps, desc = v[0][:cnt], v[1][:cnt]
        return (kps.copy(), desc.copy()) if copy else (kps, desc)

    def stereo_match(self, right: "ORBextractor", mbf: float, mb: float, keep_on_device: bool = False) -> None:
        """Frame::ComputeStereoMatches (src/Frame.cc:619-834) of this (left) extractor's last call
        against `right`'s, on the device; see include/borb.h."""
        _check(self._lib, self._lib.borb_stereo_match(self._h, right._h, mbf, mb, 1 if keep_on_device else 0))

    def stereo_result(self, frame: int = 0):
        """(mvuRight, mvDepth, matched count) of frame `frame` of the last stereo_match."""
        u, z, n, m = _vp(), _vp(), C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_stereo_result(self._h, frame, C.byref(u), C.byref(z), C.byref(n),
                                                       C.byref(m)))
        cnt = n.value
        if cnt == 0:
            return np.zeros(0, np.float32), np.zeros(0, np.float32), 0
        ur = np.ctypeslib.as_array(C.cast(u, C.POINTER(C.c_float)), shape=(cnt,)).copy()
        dz = np.ctypeslib.as_array(C.cast(z, C.POINTER(C.c_float)), shape=(cnt,)).copy()
        return ur, dz, m.value

    def search_by_projection(self, last: dict, Tcw_cur, mbf: float, mb: float, th: float, frame: int = 0,
                             mono: bool = False, check_orientation: bool = True, use_stereo: bool = False):
        """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (src/ORBmatcher.cc:1503-1668)
        with frame `frame` of this extractor's last call as the CurrentFrame, on the device (needs
        set_camera).  last: dict with mp_xyz [n,3] f32, mp_desc [n,32] u8, mp_flags [n] u8, octave [n] i32,
        angle [n] f32, Tcw [3,4] f32 (see include/borb.h).  -> (match [n_cur] i32, nmatches)."""
        keep = [np.ascontiguousarray(last["mp_xyz"], np.float32), np.ascontiguousarray(last["mp_desc"], np.uint8),
                np.ascontiguousarray(last["mp_flags"], np.uint8), np.ascontiguousarray(last["octave"], np.int32),
                np.ascontiguousarray(last["angle"], np.float32), np.ascontiguousarray(last["Tcw"], np.float32)]
        lf = _LastFrame(len(keep[2]), *[a.ctypes.data for a in keep])
        tc = np.ascontiguousarray(Tcw_cur, np.float32)
        out = np.full(max(self._lib.borb_max_keypoints(self._h), 1), -1, np.int32)
        n, nm = C.c_int(), C.c_int()
        flags = (1 if mono else 0) | (0 if check_orientation else 2) | (4 if use_stereo else 0)
        _check(self._lib, self._lib.borb_search_by_projection(self._h, frame, C.addressof(lf), tc.ctypes.data, mbf, mb, th,
                                                              flags, out.ctypes.data, C.byref(n), C.byref(nm)))
        return out[:n.value].copy(), nm.value

    def search_local_points(self, mps: dict, occupied=None, th: float = 1.0, nnratio: float = 0.8, frame: int = 0,
                            use_stereo: bool = False):
        """ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:59-154) with frame `frame` of this
        extractor's last call as F, on the device.  mps: dict with proj [n,4] f32, desc [n,32] u8, flags [n] u8,
        level [n] i32 (see include/borb.h); occupied [n_cur] u8 or None.  -> (match [n_cur] i32, nmatches)."""
        keep = [np.ascontiguousarray(mps["proj"], np.float32), np.ascontiguousarray(mps["desc"], np.uint8),
                np.ascontiguousarray(mps["flags"], np.uint8), np.ascontiguousarray(mps["level"], np.int32)]
        mp = _MapPoints(len(keep[2]), *[a.ctypes.data for a in keep])
        occ = np.ascontiguousarray(occupied, np.uint8) if occupied is not None else None
        out = np.full(max(self._lib.borb_max_keypoints(self._h), 1), -1, np.int32)
        n, nm = C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_search_local_points(self._h, frame, C.addressof(mp), occ.ctypes.data if occ is not None else None,
                                                             th, nnratio, 4 if use_stereo else 0, out.ctypes.data,
                                                             C.byref(n), C.byref(nm)))
        return out[:n.value].copy(), nm.value

    def load_frame(self, kps: np.ndarray, desc: np.ndarray, cols: int, rows: int, u_right=None) -> None:
        """Put a frame extracted earlier (a KeyFrame's mvKeys + mDescriptors, optionally its mvuRight) back into the handle as
        if an extract call had produced it: mvKeysUn, mGrid and the bounds are rebuilt on the device (set_camera first)."""
        k = np.ascontiguousarray(kps, KP_DTYPE)
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        ur = np.ascontiguousarray(u_right, np.float32) if u_right is not None else None
        _check(self._lib, self._lib.borb_load_frame(self._h, k.ctypes.data, d.ctypes.data, ur.ctypes.data if ur is not None else None,
                                                    len(k), cols, rows))
        self._n = 1

    def fuse_search(self, mps: dict, Tcw, Ow, bf: float, th: float = 3.0, frame: int = 0, use_stereo: bool = False):
        """The search of ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:968-1131) with frame `frame` of this handle
        (a keyframe put there by load_frame) as pKF.  mps: flags [n] u8 (bit 0 map point, bit 1 bad, bit 2 already in the
        keyframe), xyz [n,3], desc [n,32], dist [n,2] (mfMinDistance, mfMaxDistance), normal [n,3].
        -> (best [n] i32: keypoint of pKF per map point or -1, nFused)."""
        keep = [np.ascontiguousarray(mps["xyz"], np.float32), np.ascontiguousarray(mps["desc"], np.uint8),
                np.ascontiguousarray(mps["flags"], np.uint8), np.ascontiguousarray(mps["dist"], np.float32),
                np.ascontiguousarray(mps["normal"], np.float32)]
        rec = _FusePoints(len(keep[2]), *[a.ctypes.data for a in keep])
        T, O = np.ascontiguousarray(Tcw, np.float32), np.ascontiguousarray(Ow, np.float32)
        out = np.full(max(rec.n, 1), -1, np.int32)
        nf = C.c_int()
        _check(self._lib, self._lib.borb_fuse_search(self._h, frame, C.addressof(rec), T.ctypes.data, O.ctypes.data, bf, th,
                                                     4 if use_stereo else 0, out.ctypes.data, C.byref(nf)))
        return out[:rec.n].copy(), nf.value

    def project_search(self, pts: dict, T1, T2=None, Ow=None, th: float = 4.0, max_dist: int = 50, frame: int = 0):
        """The search of ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:1133-1260) or one direction
        of SearchBySim3 (:1262-1501) with this handle's frame (load_frame) as the keyframe.  pts: flags, xyz, desc, dist,
        normal (None = no viewing-angle test); T1 / T2 / Ow as include/borb.h: borb_projection.
        -> (best [n] i32: keypoint per point or -1, number found)."""
        keep = []
        rec, pr = _fuse_points(pts, keep), _projection(T1, T2, Ow, keep)
        out = np.full(max(rec.n, 1), -1, np.int32)
        nf = C.c_int()
        _check(self._lib, self._lib.borb_project_search(self._h, frame, C.addressof(rec), C.addressof(pr), th, max_dist,
                                                        out.ctypes.data, C.byref(nf)))
        return out[:rec.n].copy(), nf.value

    def search_by_projection_sim3(self, pts: dict, T, Ow, occupied=None, th: int = 10, frame: int = 0):
        """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:353-485), this handle's frame as
        pKF.  T = [Rcw | tcw], Ow as the reference derives them from Scw.  -> (match [n_kf] i32: point put into vpMatched[idx]
        or -1, nmatches)."""
        keep = []
        rec, pr = _fuse_points(pts, keep), _projection(T, None, Ow, keep)
        occ = np.ascontiguousarray(occupied, np.uint8) if occupied is not None else None
        out = np.full(max(self._lib.borb_max_keypoints(self._h), 1), -1, np.int32)
        n, nm = C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_search_by_projection_sim3(self._h, frame, C.addressof(rec), C.addressof(pr),
                                                                   occ.ctypes.data if occ is not None else None, int(th),
                                                                   out.ctypes.data, C.byref(n), C.byref(nm)))
        return out[:n.value].copy(), nm.value

    def search_by_sim3(self, other: "ORBextractor", pts1: dict, pts2: dict, T1w, T2w, T12, T21, th: float = 7.5, pre12=None):
        """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1262-1501): self holds pKF1, other pKF2 (load_frame).
        -> (vpMatches12 as pKF2 keypoint indices [n1] i32, nFound)."""
        keep = []
        a, b = _fuse_points(dict(pts1, normal=None), keep), _fuse_points(dict(pts2, normal=None), keep)
        mats = [np.ascontiguousarray(x, np.float32) for x in (T1w, T2w, T12, T21)]
        pre = np.ascontiguousarray(pre12, np.int32) if pre12 is not None else None
        out = np.full(max(a.n, 1), -1, np.int32)
        nf = C.c_int()
        _check(self._lib, self._lib.borb_search_by_sim3(self._h, other._h, C.addressof(a), C.addressof(b), *[m.ctypes.data for m in mats],
                                                        th, pre.ctypes.data if pre is not None else None, out.ctypes.data, C.byref(nf)))
        return out[:a.n].copy(), nf.value

    def search_by_projection_keyframe(self, kf: dict, Tcw_cur, occupied=None, th: float = 10.0, orb_dist: int = 100,
                                      frame: int = 0, check_orientation: bool = True):
        """The relocalisation matcher ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)
        (src/ORBmatcher.cc:1670-1798) with frame `frame` of this extractor's last call as CurrentFrame, on the device.
        kf: dict with flags [n] u8 (bit 0 map point, bit 1 bad, bit 2 already found), xyz [n,3] f32, desc [n,32] u8,
        dist [n,2] f32 (mfMinDistance, mfMaxDistance), angle [n] f32; occupied [n_cur] u8 or None.
        -> (match [n_cur] i32: keyframe keypoint whose map point the slot receives or -1, nmatches)."""
        keep = [np.ascontiguousarray(kf["xyz"], np.float32), np.ascontiguousarray(kf["desc"], np.uint8),
                np.ascontiguousarray(kf["flags"], np.uint8), np.ascontiguousarray(kf["dist"], np.float32),
                np.ascontiguousarray(kf["angle"], np.float32)]
        rec = _KeyFramePoints(len(keep[2]), *[a.ctypes.data for a in keep])
        T = np.ascontiguousarray(Tcw_cur, np.float32)
        occ = np.ascontiguousarray(occupied, np.uint8) if occupied is not None else None
        out = np.full(max(self._lib.borb_max_keypoints(self._h), 1), -1, np.int32)
        n, nm = C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_search_by_projection_keyframe(
            self._h, frame, C.addressof(rec), T.ctypes.data, occ.ctypes.data if occ is not None else None, th, orb_dist,
            0 if check_orientation else 2, out.ctypes.data, C.byref(n), C.byref(nm)))
        return out[:n.value].copy(), nm.value

    def search_for_initialization(self, f1: dict, prev, window: int = 100, nnratio: float = 0.9, frame: int = 0,
                                  check_orientation: bool = True):
        """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:487-632) with frame `frame` of this extractor's last
        call as F2, on the device.  f1: dict with kps (octave, angle) and desc of the initial frame; prev [n1,2] f32 =
        vbPrevMatched.  -> (vnMatches12 [n1] i32, updated vbPrevMatched, nmatches)."""
        keep = [np.ascontiguousarray(f1["kps"]["octave"], np.int32), np.ascontiguousarray(f1["kps"]["angle"], np.float32),
                np.ascontiguousarray(f1["desc"], np.uint8)]
        fr = _InitFrame(len(keep[0]), *[a.ctypes.data for a in keep])
        pv = np.array(prev, np.float32, copy=True, order="C")
        out = np.full(max(fr.n, 1), -1, np.int32)
        nm = C.c_int()
        _check(self._lib, self._lib.borb_search_for_initialization(self._h, frame, C.addressof(fr), pv.ctypes.data, window,
                                                                   nnratio, 0 if check_orientation else 2, out.ctypes.data,
                                                                   C.byref(nm)))
        return out[:fr.n].copy(), pv, nm.value

    def compute_bow(self, voc: ORBVocabulary, frame: int = 0, levelsup: int = 4) -> dict:
        """Frame::ComputeBoW (src/Frame.cc:532-539) on frame `frame` of the last call: mBowVec as (word_id, word_val),
        mFeatVec as CSR (node_id, node_start, feat_idx).  The device copies stay for search_by_bow."""
        _check(self._lib, self._lib.borb_compute_bow(self._h, frame, voc._h, levelsup))
        p = [_vp() for _ in range(5)]
        nw, nn = C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_bow_result(self._h, frame, C.byref(p[0]), C.byref(p[1]), C.byref(nw), C.byref(p[2]),
                                                    C.byref(p[3]), C.byref(p[4]), C.byref(nn)))

        def view(ptr, n, dt):
            return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_ubyte)), (n * np.dtype(dt).itemsize,)).view(dt).copy() if n else np.zeros(0, dt)
        ns = view(p[3], nn.value + 1, np.int32)
        return {"word_id": view(p[0], nw.value, np.uint32), "word_val": view(p[1], nw.value, np.float64),
                "node_id": view(p[2], nn.value, np.uint32), "node_start": ns, "feat_idx": view(p[4], int(ns[-1]), np.uint32)}

    def search_by_bow(self, kf: dict, nnratio: float = 0.7, frame: int = 0, check_orientation: bool = True):
        """ORBmatcher::SearchByBoW(pKF, F, vpMapPointMatches) (src/ORBmatcher.cc:195-351) with frame `frame` of the last
        call (after compute_bow) as F.  kf: dict desc [n,32] u8, angle [n] f32 (mvKeysUn), mp_flags [n] u8 (bit 0 map
        point, bit 1 bad), fv = the keyframe's feature vector (dict as compute_bow / ORBVocabulary.transform return).
        -> (match [n_cur] i32: keyframe keypoint index or -1, nmatches)."""
        fv = kf["fv"]
        keep = [np.ascontiguousarray(kf["desc"], np.uint8), np.ascontiguousarray(kf["angle"], np.float32),
                np.ascontiguousarray(kf["mp_flags"], np.uint8), np.ascontiguousarray(fv["node_id"], np.uint32),
                np.ascontiguousarray(fv["node_start"], np.int32), np.ascontiguousarray(fv["feat_idx"], np.uint32)]
        rec = _KeyFrame(len(keep[2]), keep[0].ctypes.data, keep[1].ctypes.data, keep[2].ctypes.data, len(keep[3]),
                        keep[3].ctypes.data, keep[4].ctypes.data, keep[5].ctypes.data)
        out = np.full(max(self._lib.borb_max_keypoints(self._h), 1), -1, np.int32)
        n, nm = C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_search_by_bow(self._h, frame, C.addressof(rec), nnratio, 0 if check_orientation else 2,
                                                       out.ctypes.data, C.byref(n), C.byref(nm)))
        return out[:n.value].copy(), nm.value

    def match_rounds(self) -> int:
        return self._lib.borb_match_rounds(self._h) & 0xff

    def match_rescans(self) -> int:
        return self._lib.borb_match_rounds(self._h) >> 8

    def level_size(self, level: int):
        c, r = C.c_int(), C.c_int()
        _check(self._lib, self._lib.borb_level_size(self._h, level, C.byref(c), C.byref(r)))
        return c.value, r.value

    def pyramid_level(self, frame: int, level: int, border: int = 0) -> np.ndarray:
        """Level image as a view with `border` (<= 19) pixels of its reflect-101 frame."""
        data, c, r, st = _vp(), C.c_int(), C.c_int(), C.c_size_t()
        _check(self._lib, self._lib.borb_pyramid(self._h, frame, level, C.byref(data), C.byref(c),
                                                 C.byref(r), C.byref(st)))
        step = st.value
        base = data.value - border * step - border
        H, W = r.value + 2 * border, c.value + 2 * border
        buf = (C.c_uint8 * (H * step)).from_address(base)
        return np.lib.stride_tricks.as_strided(np.frombuffer(buf, np.uint8), shape=(H, W), strides=(step, 1)).copy()

    # -- stage taps for the parity tests --
    def debug_candidates(self, frame: int, level: int):
        cap = 1 << 20
        xs = np.zeros(cap, np.int32); ys = np.zeros(cap, np.int32); sc = np.zeros(cap, np.int32)
        n = C.c_int()
        _check(self._lib, self._lib.borb_debug_candidates(self._h, frame, level, xs.ctypes.data, ys.ctypes.data,
                                                          sc.ctypes.data, cap, C.byref(n)))
        m = min(n.value, cap)
        return xs[:m].copy(), ys[:m].copy(), sc[:m].copy()

    def debug_describe(self, img: np.ndarray, x: int, y: int, angle_deg: float) -> np.ndarray:
        """computeOrbDescriptor of pixel (x, y) of `img` at `angle_deg` by the device function of the extract kernels"""
        a = np.ascontiguousarray(img, np.uint8)
        out = np.zeros(32, np.uint8)
        _check(self._lib, self._lib.borb_debug_describe(self._h, a.ctypes.data, a.shape[1], a.shape[0], a.strides[0], x, y,
                                                        float(angle_deg), out.ctypes.data))
        return out

    def debug_blurred(self, frame: int, level: int) -> np.ndarray:
        c, r = self.level_size(level)
        out = np.zeros((r, c), np.uint8)
        _check(self._lib, self._lib.borb_debug_blurred(self._h, frame, level, out.ctypes.data))
        return out


def pinned_empty(shape) -> np.ndarray:
    """uint8 array in page-locked host memory (borb_host_alloc); freed when garbage collected."""
    lib = load_library()
    n = int(np.prod(shape))
    p = lib.borb_host_alloc(n)
    if not p:
        raise BorbError(-2, lib.borb_last_error().decode())
    buf = (C.c_uint8 * n).from_address(p)
    arr = np.frombuffer(buf, np.uint8).reshape(shape)

    class _Owner:
        def __del__(self, p=p, lib=lib):
            lib.borb_host_free(p)

    _PINNED[arr.ctypes.data] = _Owner()
    return arr


_PINNED: dict = {}
