"""ctypes bindings for the CPU checker -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Loads oracle/liborb_oracle.so (our restatement, orb_oracle.cpp) and, when present,
oracle/_ref/libreforb{,_mono}.so (the reference's own src/ORBextractor.cc compiled in place).
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
CAND_DTYPE = np.dtype([("x", "<i4"), ("y", "<i4"), ("score", "<i4")])
assert KP_DTYPE.itemsize == 28

_u8p = C.POINTER(C.c_uint8)
_vp = C.c_void_p


def build(ref: str = "/root/reference") -> None:
    """Compile the oracle (and oracle/_ref when the reference tree is present)."""
    subprocess.run(["make", "-s", "-C", str(HERE), f"REF={ref}", "all"], check=True)


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_u8p)


class _Base:
    """Common driver for the three libraries (same call shapes, different symbol prefix)."""

    prefix = ""

    def __init__(self, lib: C.CDLL, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.lib = lib
        p = self.prefix
        f = getattr(lib, p + "create")
        f.restype = _vp
        f.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        self.h = _vp(f(nfeatures, scale_factor, nlevels, ini_th, min_th))
        self.nfeatures, self.nlevels = nfeatures, nlevels
        ex = getattr(lib, p + "extract")
        ex.restype = C.c_int
        ex.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_size_t, _vp, _vp, C.c_int]
        self._extract = ex
        getattr(lib, p + "destroy").argtypes = [_vp]
        for name in ("table",):
            getattr(lib, p + name).argtypes = [_vp, C.c_int, _vp]
        getattr(lib, p + "quotas").argtypes = [_vp, _vp]
        getattr(lib, p + "umax").argtypes = [_vp, _vp]
        getattr(lib, p + "pyramid_copy").argtypes = [_vp, C.c_int, C.c_int, _vp]

    def close(self):
        if self.h:
            getattr(self.lib, self.prefix + "destroy")(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def extract(self, img: np.ndarray | None):
        """-> (keypoints structured array, descriptors [n,32] u8) or an int status < 0."""
        cap = self.nfeatures + 64 * self.nlevels + 4096
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        if img is None or img.size == 0:
            n = self._extract(self.h, None, 0, 0, 0, kps.ctypes.data, desc.ctypes.data, cap)
            return n
        assert img.dtype == np.uint8 and img.ndim == 2 and img.strides[1] == 1
        n = self._extract(self.h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0],
                          kps.ctypes.data, desc.ctypes.data, cap)
        if n < 0:
            return n
        assert n <= cap
        return kps[:n].copy(), desc[:n].copy()

    def table(self, which: int) -> np.ndarray:
        out = np.zeros(self.nlevels, np.float32)
        getattr(self.lib, self.prefix + "table")(self.h, which, out.ctypes.data)
        return out

    def quotas(self) -> np.ndarray:
        out = np.zeros(self.nlevels, np.int32)
        getattr(self.lib, self.prefix + "quotas")(self.h, out.ctypes.data)
        return out

    def umax(self) -> np.ndarray:
        out = np.zeros(16, np.int32)
        getattr(self.lib, self.prefix + "umax")(self.h, out.ctypes.data)
        return out


class Oracle(_Base):
    """Our restatement (liborb_oracle.so, canonical quadtree tie-break)."""

    prefix = "orbo_"
    _lib = None

    @classmethod
    def load(cls) -> C.CDLL:
        if cls._lib is None:
            path = HERE / "liborb_oracle.so"
            if not path.exists():
                build()
            lib = C.CDLL(str(path))
            lib.orbo_level_size.argtypes = [_vp, C.c_int, _vp, _vp]
            lib.orbo_blurred_copy.argtypes = [_vp, C.c_int, _vp]
            lib.orbo_candidates.argtypes = [_vp, C.c_int, _vp, C.c_int]
            lib.orbo_level_keypoints.argtypes = [_vp, C.c_int, _vp, C.c_int]
            lib.orbo_distribute.argtypes = [_vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                            C.c_int, _vp, C.c_int]
            lib.orbo_describe.argtypes = [_vp, C.c_size_t, C.c_int, C.c_int, C.c_float, _vp]
            lib.orbo_ic_angle.argtypes = [_vp, C.c_size_t, C.c_int, C.c_int]
            lib.orbo_ic_angle.restype = C.c_float
            lib.orbo_pattern.restype = _vp
            cls._lib = lib
        return cls._lib

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        super().__init__(self.load(), nfeatures, scale_factor, nlevels, ini_th, min_th)

    def level_size(self, level: int):
        c, r = C.c_int(), C.c_int()
        assert self.lib.orbo_level_size(self.h, level, C.byref(c), C.byref(r)) == 0
        return c.value, r.value

    def pyramid(self, level: int, border: int = 19) -> np.ndarray:
        c, r = self.level_size(level)
        out = np.zeros((r + 2 * border, c + 2 * border), np.uint8)
        assert self.lib.orbo_pyramid_copy(self.h, level, border, out.ctypes.data) == 0
        return out

    def blurred(self, level: int):
        c, r = self.level_size(level)
        out = np.zeros((r, c), np.uint8)
        if self.lib.orbo_blurred_copy(self.h, level, out.ctypes.data) != 0:
            return None
        return out

    def candidates(self, level: int) -> np.ndarray:
        n = self.lib.orbo_candidates(self.h, level, None, 0)
        out = np.zeros(max(n, 1), CAND_DTYPE)
        self.lib.orbo_candidates(self.h, level, out.ctypes.data, n)
        return out[:n]

    def level_keypoints(self, level: int) -> np.ndarray:
        n = self.lib.orbo_level_keypoints(self.h, level, None, 0)
        out = np.zeros(max(n, 1), KP_DTYPE)
        self.lib.orbo_level_keypoints(self.h, level, out.ctypes.data, n)
        return out[:n]


def distribute(xs, ys, resp, min_x, max_x, min_y, max_y, n_target):
    """Stand-alone canonical quadtree distribution -> indices of survivors in output order."""
    lib = Oracle.load()
    xs = np.ascontiguousarray(xs, np.int32)
    ys = np.ascontiguousarray(ys, np.int32)
    resp = np.ascontiguousarray(resp, np.int32)
    out = np.zeros(len(xs) + 8, np.int32)
    n = lib.orbo_distribute(xs.ctypes.data, ys.ctypes.data, resp.ctypes.data, len(xs), min_x, max_x,
                            min_y, max_y, n_target, out.ctypes.data, len(out))
    if n < 0:
        return n
    return out[:n].copy()


class _SoLevel(C.Structure):
    _fields_ = [("data", C.c_void_p), ("cols", C.c_int), ("rows", C.c_int), ("step", C.c_size_t)]


def stereo_match(kl, dl, kr, dr, pyr_l, pyr_r, scale, inv_scale, mbf: float, mb: float):
    """Frame::ComputeStereoMatches restated (oracle/stereo_oracle.cpp).  pyr_l / pyr_r: lists of
    2-D uint8 level images (un-padded).  -> (uRight[n], depth[n], sad[n], kept)."""
    lib = Oracle.load()
    lib.so_stereo_match.argtypes = [_vp, _vp, C.c_int, _vp, _vp, C.c_int, _vp, _vp, C.c_int, _vp, _vp,
                                    C.c_float, C.c_float, _vp, _vp, _vp]
    kl = np.ascontiguousarray(kl); kr = np.ascontiguousarray(kr)
    dl = np.ascontiguousarray(dl); dr = np.ascontiguousarray(dr)
    keep = []

    def levels(pyr):
        arr = (_SoLevel * len(pyr))()
        for i, im in enumerate(pyr):
            im = np.ascontiguousarray(im)
            keep.append(im)
            arr[i] = _SoLevel(im.ctypes.data, im.shape[1], im.shape[0], im.strides[0])
        return arr

    pl, pr = levels(pyr_l), levels(pyr_r)
    scale = np.ascontiguousarray(scale, np.float32); inv_scale = np.ascontiguousarray(inv_scale, np.float32)
    n = len(kl)
    u = np.zeros(max(n, 1), np.float32); z = np.zeros(max(n, 1), np.float32); sad = np.zeros(max(n, 1), np.int32)
    kept = lib.so_stereo_match(kl.ctypes.data, dl.ctypes.data, n, kr.ctypes.data, dr.ctypes.data, len(kr),
                               C.addressof(pl), C.addressof(pr), len(pyr_l), scale.ctypes.data, inv_scale.ctypes.data,
                               mbf, mb, u.ctypes.data, z.ctypes.data, sad.ctypes.data)
    return u[:n], z[:n], sad[:n], kept


class _MatchFrame(C.Structure):
    _fields_ = [("n", C.c_int), ("kps", _vp), ("un_xy", _vp), ("desc", _vp), ("u_right", _vp),
                ("grid_count", _vp), ("grid_idx", _vp), ("Tcw", _vp)]


def _match_call(fn, cur: dict, last: dict, cam, bounds, scale, th: float, mono: bool, check_ori: bool):
    """Common marshalling for mo_search_by_projection / refmatch_search_by_projection.
    cur / last: dicts with kps (KP_DTYPE), un_xy [n,2] f32, desc [n,32] u8, Tcw [3,4] f32; cur also
    u_right [n] f32 (or None), cell_start [3073] i32, cell_idx; last also mp_flags [n] u8 (bit 0 map
    point, bit 1 Observations() > 0, bit 2 outlier), mp_xyz [n,3] f32, mp_desc [n,32] u8.
    -> (match [n_cur] i32: index of the last-frame map point or -1, nmatches)."""
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data

    def frame(d, with_grid):
        f = _MatchFrame()
        f.n = len(d["kps"])
        f.kps = arr(d["kps"], KP_DTYPE)
        f.un_xy = arr(d["un_xy"], np.float32)
        f.desc = arr(d["desc"], np.uint8)
        f.u_right = arr(d["u_right"], np.float32) if d.get("u_right") is not None else None
        if with_grid:
            f.grid_count = arr(np.diff(d["cell_start"]), np.int32)
            f.grid_idx = arr(d["cell_idx"], np.int32)
        f.Tcw = arr(d["Tcw"], np.float32)
        return f

    fc, fl = frame(cur, True), frame(last, False)
    scale = np.ascontiguousarray(scale, np.float32)
    out = np.full(max(fc.n, 1), -1, np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, C.c_float, C.c_int, C.c_int, _vp]
    nm = fn(C.addressof(fc), C.addressof(fl), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data,
            len(scale), arr(last["mp_flags"], np.uint8), arr(last["mp_xyz"], np.float32), arr(last["mp_desc"], np.uint8),
            float(th), int(mono), int(check_ori), out.ctypes.data)
    return out[:fc.n], int(nm)


def search_by_projection(cur, last, cam, bounds, scale, th, mono=False, check_ori=True):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) restated (oracle/match_oracle.cpp).
    cam = (fx, fy, cx, cy, mbf, mb)."""
    return _match_call(Oracle.load().mo_search_by_projection, cur, last, cam, bounds, scale, th, mono, check_ori)


def _local_call(fn, cur: dict, bounds, scale, mps: dict, occupied, th: float, nnratio: float):
    """Common marshalling for mo_search_local_points / refmatch_search_local_points.  mps: proj [n,4] f32
    (mTrackProjX, mTrackProjY, mTrackProjXR, mTrackViewCos), desc [n,32] u8, flags [n] u8 (bit 0 in view and not
    bad, bit 1 Observations() > 0), level [n] i32.  occupied [n_cur] u8 (2 = holds a point with observations).
    -> (match [n_cur] i32: map point put there by this call or -1, nmatches)."""
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data

    f = _MatchFrame()
    f.n = len(cur["kps"])
    f.kps = arr(cur["kps"], KP_DTYPE)
    f.un_xy = arr(cur["un_xy"], np.float32)
    f.desc = arr(cur["desc"], np.uint8)
    f.u_right = arr(cur["u_right"], np.float32) if cur.get("u_right") is not None else None
    f.grid_count = arr(np.diff(cur["cell_start"]), np.int32)
    f.grid_idx = arr(cur["cell_idx"], np.int32)
    f.Tcw = None
    scale = np.ascontiguousarray(scale, np.float32)
    out = np.full(max(f.n, 1), -1, np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_vp, _vp, _vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_float, C.c_float, _vp]
    nm = fn(C.addressof(f), arr(bounds, np.float32), scale.ctypes.data, len(scale), len(mps["flags"]), arr(mps["proj"], np.float32),
            arr(mps["desc"], np.uint8), arr(mps["flags"], np.uint8), arr(mps["level"], np.int32), arr(occupied, np.uint8),
            float(th), float(nnratio), out.ctypes.data)
    return out[:f.n], int(nm)


def search_local_points(cur, bounds, scale, mps, occupied, th=1.0, nnratio=0.8):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:59-154) restated (oracle/match_oracle.cpp)."""
    return _local_call(Oracle.load().mo_search_local_points, cur, bounds, scale, mps, occupied, th, nnratio)


def _init_call(fn, f1: dict, f2: dict, bounds, prev, window: int, nnratio: float, check_ori: bool):
    """Common marshalling for mo_search_for_initialization / refmatch_search_for_initialization.
    f1: kps, un_xy, desc; f2: kps, un_xy, desc, cell_start, cell_idx; prev [n1,2] f32 = vbPrevMatched.
    -> (vnMatches12 [n1] i32, updated vbPrevMatched [n1,2], nmatches)."""
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data

    def frame(d, with_grid):
        f = _MatchFrame()
        f.n = len(d["kps"])
        f.kps = arr(d["kps"], KP_DTYPE)
        f.un_xy = arr(d["un_xy"], np.float32)
        f.desc = arr(d["desc"], np.uint8)
        f.u_right = None
        if with_grid:
            f.grid_count = arr(np.diff(d["cell_start"]), np.int32)
            f.grid_idx = arr(d["cell_idx"], np.int32)
        f.Tcw = None
        return f

    a, b = frame(f1, False), frame(f2, True)
    pv = np.array(prev, np.float32, copy=True, order="C")
    out = np.full(max(a.n, 1), -1, np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_vp, _vp, _vp, _vp, C.c_int, C.c_float, C.c_int, _vp]
    nm = fn(C.addressof(a), C.addressof(b), arr(bounds, np.float32), pv.ctypes.data, int(window), float(nnratio),
            int(check_ori), out.ctypes.data)
    return out[:a.n], pv, int(nm)


def search_for_initialization(f1, f2, bounds, prev, window=100, nnratio=0.9, check_ori=True):
    """ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:487-632) restated (oracle/match_oracle.cpp)."""
    return _init_call(Oracle.load().mo_search_for_initialization, f1, f2, bounds, prev, window, nnratio, check_ori)


def _reloc_call(fn, cur: dict, cam, bounds, scale, log_scale: float, kf: dict, occupied, th: float, orb_dist: int,
                check_ori: bool):
    """Common marshalling for mo_search_by_projection_kf / refmatch_search_by_projection_kf.  cur as _match_call's (with
    Tcw and the grid); kf: flags [n] u8 (bit 0 map point, bit 1 bad, bit 2 already found), xyz [n,3], desc [n,32],
    dist [n,2] f32 (mfMinDistance, mfMaxDistance), angle [n] f32; occupied [n_cur] u8 or None.
    -> (match [n_cur] i32: keyframe keypoint whose map point this call put there or -1, nmatches)."""
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data

    f = _MatchFrame()
    f.n = len(cur["kps"])
    f.kps = arr(cur["kps"], KP_DTYPE)
    f.un_xy = arr(cur["un_xy"], np.float32)
    f.desc = arr(cur["desc"], np.uint8)
    f.u_right = None
    f.grid_count = arr(np.diff(cur["cell_start"]), np.int32)
    f.grid_idx = arr(cur["cell_idx"], np.int32)
    f.Tcw = arr(cur["Tcw"], np.float32)
    scale = np.ascontiguousarray(scale, np.float32)
    out = np.full(max(f.n, 1), -1, np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_vp, _vp, _vp, _vp, C.c_int, C.c_float, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.c_float, C.c_int,
                   C.c_int, _vp]
    nm = fn(C.addressof(f), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data, len(scale), float(log_scale),
            len(kf["flags"]), arr(kf["flags"], np.uint8), arr(kf["xyz"], np.float32), arr(kf["desc"], np.uint8),
            arr(kf["dist"], np.float32), arr(kf["angle"], np.float32),
            arr(occupied, np.uint8) if occupied is not None else None, float(th), int(orb_dist), int(check_ori),
            out.ctypes.data)
    return out[:f.n], int(nm)


def search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occupied, th=10.0, orb_dist=100, check_ori=True):
    """The relocalisation matcher (src/ORBmatcher.cc:1670-1798) restated (oracle/match_oracle.cpp)."""
    return _reloc_call(Oracle.load().mo_search_by_projection_kf, cur, cam, bounds, scale, log_scale, kf, occupied, th,
                       orb_dist, check_ori)


def _fuse_call(fn, kf: dict, cam, bounds, scale, inv_sigma2, log_scale: float, mps: dict, th: float):
    """Common marshalling for mo_fuse / refmatch_fuse.  kf as _match_call's current frame (kps, un_xy, desc, u_right, grid,
    Tcw); mps: flags [n] u8 (bit 0 map point, bit 1 bad, bit 2 already in the keyframe), xyz [n,3], desc [n,32], dist [n,2]
    (mfMinDistance, mfMaxDistance), normal [n,3].  -> (best [n] i32: the keyframe keypoint chosen per map point or -1, nFused)."""
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data

    f = _MatchFrame()
    f.n = len(kf["kps"])
    f.kps = arr(kf["kps"], KP_DTYPE)
    f.un_xy = arr(kf["un_xy"], np.float32)
    f.desc = arr(kf["desc"], np.uint8)
    f.u_right = arr(kf["u_right"], np.float32) if kf.get("u_right") is not None else None
    f.grid_count = arr(np.diff(kf["cell_start"]), np.int32)
    f.grid_idx = arr(kf["cell_idx"], np.int32)
    f.Tcw = arr(kf["Tcw"], np.float32)
    scale = np.ascontiguousarray(scale, np.float32)
    n = len(mps["flags"])
    out = np.full(max(n, 1), -1, np.int32)
    fn.restype = C.c_int
    fn.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int, C.c_float, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_float, _vp]
    nf = fn(C.addressof(f), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data, arr(inv_sigma2, np.float32),
            len(scale), float(log_scale), n, arr(mps["flags"], np.uint8), arr(mps["xyz"], np.float32), arr(mps["desc"], np.uint8),
            arr(mps["dist"], np.float32), arr(mps["normal"], np.float32), float(th), out.ctypes.data)
    return out[:n], int(nf)


def _kf_frame(kf: dict, keep: list):
    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data
    f = _MatchFrame()
    f.n = len(kf["kps"])
    f.kps = arr(kf["kps"], KP_DTYPE)
    f.un_xy = arr(kf["un_xy"], np.float32)
    f.desc = arr(kf["desc"], np.uint8)
    f.u_right = arr(kf["u_right"], np.float32) if kf.get("u_right") is not None else None
    f.grid_count = arr(np.diff(kf["cell_start"]), np.int32)
    f.grid_idx = arr(kf["cell_idx"], np.int32)
    f.Tcw = arr(kf["Tcw"], np.float32)
    keep.append(f)
    return f, arr


def sim3_decompose(Scw) -> tuple:
    """Scw [4,4] -> (T [3,4] = [Rcw | tcw], Ow [3]) as src/ORBmatcher.cc:360-365 / :1140-1146 compute them"""
    S = np.ascontiguousarray(Scw, np.float32)
    T, Ow = np.zeros((3, 4), np.float32), np.zeros(3, np.float32)
    Oracle.load().mo_sim3_decompose(_ptr(S), _ptr(T), _ptr(Ow))
    return T, Ow


def sim3_relative(s12: float, R12, t12) -> tuple:
    """-> (T12 = [s12 R12 | t12], T21 = [sR21 | t21]) as src/ORBmatcher.cc:1279-1281"""
    R, t = np.ascontiguousarray(R12, np.float32), np.ascontiguousarray(t12, np.float32)
    T12, T21 = np.zeros((3, 4), np.float32), np.zeros((3, 4), np.float32)
    fn = Oracle.load().mo_sim3_relative
    fn.argtypes = [C.c_float, _vp, _vp, _vp, _vp]
    fn(float(s12), _ptr(R), _ptr(t), _ptr(T12), _ptr(T21))
    return T12, T21


def project_search(kf, cam, bounds, scale, log_scale, T1, T2, Ow, pts, th, max_dist):
    """mo_project_search (oracle/match_oracle.cpp): the choice per point of Fuse(pKF, Scw, ...) / one direction of SearchBySim3.
    pts: flags, xyz, desc, dist, normal (or None)"""
    keep = []
    f, arr = _kf_frame(kf, keep)
    scale = np.ascontiguousarray(scale, np.float32)
    n = len(pts["flags"])
    out = np.full(max(n, 1), -1, np.int32)
    fn = Oracle.load().mo_project_search
    fn.restype = C.c_int
    fn.argtypes = [_vp, _vp, _vp, _vp, C.c_int, C.c_float, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_float, C.c_int, _vp]
    nf = fn(C.addressof(f), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data, len(scale), float(log_scale),
            arr(T1, np.float32), arr(T2, np.float32) if T2 is not None else None, arr(Ow, np.float32) if Ow is not None else None,
            n, arr(pts["flags"], np.uint8), arr(pts["xyz"], np.float32), arr(pts["desc"], np.uint8), arr(pts["dist"], np.float32),
            arr(pts["normal"], np.float32) if pts.get("normal") is not None else None, float(th), int(max_dist), out.ctypes.data)
    return out[:n], int(nf)


def fuse_sim3(kf, cam, bounds, scale, log_scale, Scw, pts, th=4.0):
    """ORBmatcher::Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (src/ORBmatcher.cc:1133-1260): the choice per point, restated"""
    T, Ow = sim3_decompose(Scw)
    return project_search(kf, cam, bounds, scale, log_scale, T, None, Ow, pts, th, 50)


def search_by_projection_kf_sim3(kf, cam, bounds, scale, log_scale, Scw, pts, occupied, th=10):
    """ORBmatcher::SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (src/ORBmatcher.cc:353-485), restated"""
    keep = []
    f, arr = _kf_frame(kf, keep)
    T, Ow = sim3_decompose(Scw)
    scale = np.ascontiguousarray(scale, np.float32)
    out = np.full(max(f.n, 1), -1, np.int32)
    fn = Oracle.load().mo_search_by_projection_kf_sim3
    fn.restype = C.c_int
    fn.argtypes = [_vp, _vp, _vp, _vp, C.c_int, C.c_float, _vp, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp]
    nm = fn(C.addressof(f), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data, len(scale), float(log_scale),
            arr(T, np.float32), arr(Ow, np.float32), len(pts["flags"]), arr(pts["flags"], np.uint8), arr(pts["xyz"], np.float32),
            arr(pts["desc"], np.uint8), arr(pts["dist"], np.float32), arr(pts["normal"], np.float32),
            arr(occupied, np.uint8) if occupied is not None else None, int(th), out.ctypes.data)
    return out[:f.n], int(nm)


def _sim3_args(k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12, th, keep, with_sigma):
    f1, arr = _kf_frame(k1, keep)
    f2, _ = _kf_frame(k2, keep)
    scale = np.ascontiguousarray(scale, np.float32)
    keep.append(scale)
    a = [C.addressof(f1), C.addressof(f2), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data]
    if with_sigma:
        a.append(arr(inv_sigma2, np.float32))
    a += [len(scale), float(log_scale)]
    for p in (p1, p2):
        a += [arr(p["flags"], np.uint8), arr(p["xyz"], np.float32), arr(p["desc"], np.uint8), arr(p["dist"], np.float32)]
    a += [arr(pre12, np.int32), float(s12), arr(R12, np.float32), arr(t12, np.float32), float(th)]
    return a, f1.n


def search_by_sim3(k1, k2, cam, bounds, scale, log_scale, p1, p2, pre12, s12, R12, t12, th=7.5):
    """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1262-1501), restated -> (vpMatches12 as pKF2 keypoint indices, nFound)"""
    keep = []
    a, n1 = _sim3_args(k1, k2, cam, bounds, scale, None, log_scale, p1, p2, pre12, s12, R12, t12, th, keep, False)
    out = np.full(max(n1, 1), -1, np.int32)
    fn = Oracle.load().mo_search_by_sim3
    fn.restype = C.c_int
    fn.argtypes = [_vp] * 5 + [C.c_int, C.c_float] + [_vp] * 9 + [C.c_float, _vp, _vp, C.c_float, _vp]
    nf = fn(*a, out.ctypes.data)
    return out[:n1], int(nf)


def fuse(kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th=3.0):
    """ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:968-1131): the choice per map point, restated"""
    return _fuse_call(Oracle.load().mo_fuse, kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th)


class RefMatcher:
    """The reference's own src/ORBmatcher.cc + src/Frame.cc (oracle/_ref/librefmatch.so)."""

    @classmethod
    def _load(cls):
        if cls._lib is None:
            cls._lib = C.CDLL(str(HERE / "_ref" / "librefmatch.so"))
        return cls._lib

    @classmethod
    def fuse_sim3(cls, kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, th=4.0):
        keep = []
        f, arr = _kf_frame(kf, keep)
        scale = np.ascontiguousarray(scale, np.float32)
        n = len(pts["flags"])
        out = np.full(max(n, 1), -1, np.int32)
        fn = cls._load().refmatch_fuse_sim3
        fn.restype = C.c_int
        fn.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int, C.c_float, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, C.c_float, _vp]
        nf = fn(C.addressof(f), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data, arr(inv_sigma2, np.float32),
                len(scale), float(log_scale), arr(Scw, np.float32), n, arr(pts["flags"], np.uint8), arr(pts["xyz"], np.float32),
                arr(pts["desc"], np.uint8), arr(pts["dist"], np.float32), arr(pts["normal"], np.float32), float(th), out.ctypes.data)
        return out[:n], int(nf)

    @classmethod
    def search_by_projection_kf_sim3(cls, kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occupied, th=10):
        keep = []
        f, arr = _kf_frame(kf, keep)
        scale = np.ascontiguousarray(scale, np.float32)
        out = np.full(max(f.n, 1), -1, np.int32)
        fn = cls._load().refmatch_search_by_projection_kf_sim3
        fn.restype = C.c_int
        fn.argtypes = [_vp, _vp, _vp, _vp, _vp, C.c_int, C.c_float, _vp, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, C.c_int, _vp]
        nm = fn(C.addressof(f), arr(cam, np.float32), arr(bounds, np.float32), scale.ctypes.data, arr(inv_sigma2, np.float32),
                len(scale), float(log_scale), arr(Scw, np.float32), len(pts["flags"]), arr(pts["flags"], np.uint8),
                arr(pts["xyz"], np.float32), arr(pts["desc"], np.uint8), arr(pts["dist"], np.float32), arr(pts["normal"], np.float32),
                arr(occupied, np.uint8) if occupied is not None else None, int(th), out.ctypes.data)
        return out[:f.n], int(nm)

    @classmethod
    def search_by_sim3(cls, k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12, th=7.5):
        keep = []
        a, n1 = _sim3_args(k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12, th, keep, True)
        out = np.full(max(n1, 1), -1, np.int32)
        fn = cls._load().refmatch_search_by_sim3
        fn.restype = C.c_int
        fn.argtypes = [_vp] * 6 + [C.c_int, C.c_float] + [_vp] * 9 + [C.c_float, _vp, _vp, C.c_float, _vp]
        nf = fn(*a, out.ctypes.data)
        return out[:n1], int(nf)

    @classmethod
    def fuse(cls, kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th=3.0):
        if cls._lib is None:
            cls._lib = C.CDLL(str(HERE / "_ref" / "librefmatch.so"))
        return _fuse_call(cls._lib.refmatch_fuse, kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th)

    @classmethod
    def search_by_projection_kf(cls, cur, cam, bounds, scale, log_scale, kf, occupied, th=10.0, orb_dist=100, check_ori=True):
        if cls._lib is None:
            cls._lib = C.CDLL(str(HERE / "_ref" / "librefmatch.so"))
        return _reloc_call(cls._lib.refmatch_search_by_projection_kf, cur, cam, bounds, scale, log_scale, kf, occupied, th,
                           orb_dist, check_ori)

    _lib = None

    @classmethod
    def available(cls) -> bool:
        return (HERE / "_ref" / "librefmatch.so").exists()

    @classmethod
    def search_by_projection(cls, cur, last, cam, bounds, scale, th, mono=False, check_ori=True):
        if cls._lib is None:
            cls._lib = C.CDLL(str(HERE / "_ref" / "librefmatch.so"))
        return _match_call(cls._lib.refmatch_search_by_projection, cur, last, cam, bounds, scale, th, mono, check_ori)

    @classmethod
    def search_local_points(cls, cur, bounds, scale, mps, occupied, th=1.0, nnratio=0.8):
        if cls._lib is None:
            cls._lib = C.CDLL(str(HERE / "_ref" / "librefmatch.so"))
        return _local_call(cls._lib.refmatch_search_local_points, cur, bounds, scale, mps, occupied, th, nnratio)

    @classmethod
    def search_for_initialization(cls, f1, f2, bounds, prev, window=100, nnratio=0.9, check_ori=True):
        if cls._lib is None:
            cls._lib = C.CDLL(str(HERE / "_ref" / "librefmatch.so"))
        return _init_call(cls._lib.refmatch_search_for_initialization, f1, f2, bounds, prev, window, nnratio, check_ori)


def _cam(K, dist) -> np.ndarray:
    d = list(dist) + [0.0] * (5 - len(dist))
    return np.asarray(list(K) + d, np.float32)


def undistort_points(xy: np.ndarray, K, dist) -> np.ndarray:
    """cv::undistortPoints(xy, ., mK, mDistCoef, Mat(), mK) restated (oracle/frame_oracle.c)."""
    lib = Oracle.load()
    lib.fo_undistort_points.argtypes = [_vp, C.c_int, _vp, _vp]
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    out = np.empty_like(xy)
    cam = _cam(K, dist)
    lib.fo_undistort_points(xy.ctypes.data, len(xy), cam.ctypes.data, out.ctypes.data)
    return out


def frame_post(kps: np.ndarray, cols: int, rows: int, K, dist):
    """Frame::UndistortKeyPoints + ComputeImageBounds + AssignFeaturesToGrid restated.
    -> (xy_un [n,2] f32, cell_start [3073] i32, cell_idx [n_in_grid] i32, bounds [4] f32)."""
    lib = Oracle.load()
    lib.fo_undistort_keypoints.argtypes = [_vp, C.c_int, _vp, _vp]
    lib.fo_image_bounds.argtypes = [C.c_int, C.c_int, _vp, _vp]
    lib.fo_assign_grid.argtypes = [_vp, C.c_int, _vp, _vp, _vp]
    cam = _cam(K, dist)
    n = len(kps)
    xy = np.ascontiguousarray(np.stack([kps["x"], kps["y"]], 1).astype(np.float32)) if n else np.zeros((0, 2), np.float32)
    un = np.zeros((max(n, 1), 2), np.float32)
    lib.fo_undistort_keypoints(xy.ctypes.data, n, cam.ctypes.data, un.ctypes.data)
    bounds = np.zeros(4, np.float32)
    lib.fo_image_bounds(cols, rows, cam.ctypes.data, bounds.ctypes.data)
    start = np.zeros(64 * 48 + 1, np.int32)
    idx = np.zeros(max(n, 1), np.int32)
    m = lib.fo_assign_grid(un.ctypes.data, n, bounds.ctypes.data, start.ctypes.data, idx.ctypes.data)
    return un[:n], start, idx[:m], bounds


def cvt_gray(img: np.ndarray, rgb_order: bool) -> np.ndarray:
    """cv::cvtColor(..., CV_{RGB,BGR,RGBA,BGRA}2GRAY) restated (oracle/cvprim_pre.c)."""
    lib = Oracle.load()
    lib.cvp_cvt_gray.argtypes = [_vp, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, _vp, C.c_size_t]
    img = np.ascontiguousarray(img)
    rows, cols = img.shape[:2]
    ch = 1 if img.ndim == 2 else img.shape[2]
    out = np.empty((rows, cols), np.uint8)
    lib.cvp_cvt_gray(img.ctypes.data, img.strides[0], cols, rows, ch, int(rgb_order), out.ctypes.data, cols)
    return out


def remap_linear(img: np.ndarray, map_x: np.ndarray, map_y: np.ndarray) -> np.ndarray:
    """cv::remap(img, ., map_x, map_y, INTER_LINEAR) restated (oracle/cvprim_pre.c)."""
    lib = Oracle.load()
    lib.cvp_remap_linear.argtypes = [_vp, C.c_size_t, C.c_int, C.c_int, C.c_int, _vp, _vp, C.c_size_t, _vp,
                                     C.c_size_t, C.c_int, C.c_int]
    img = np.ascontiguousarray(img)
    map_x = np.ascontiguousarray(map_x, np.float32); map_y = np.ascontiguousarray(map_y, np.float32)
    assert map_x.shape == map_y.shape
    rows, cols = img.shape[:2]
    ch = 1 if img.ndim == 2 else img.shape[2]
    dr, dc = map_x.shape
    out = np.empty((dr, dc) if img.ndim == 2 else (dr, dc, ch), np.uint8)
    lib.cvp_remap_linear(img.ctypes.data, img.strides[0], cols, rows, ch, map_x.ctypes.data, map_y.ctypes.data,
                         dc, out.ctypes.data, dc * ch, dc, dr)
    return out


def describe(img: np.ndarray, x: int, y: int, angle_deg: float) -> np.ndarray:
    out = np.zeros(32, np.uint8)
    Oracle.load().orbo_describe(img.ctypes.data, img.strides[0], x, y, angle_deg, out.ctypes.data)
    return out


def ic_angle(img: np.ndarray, x: int, y: int) -> float:
    return float(Oracle.load().orbo_ic_angle(img.ctypes.data, img.strides[0], x, y))


def pattern() -> np.ndarray:
    p = Oracle.load().orbo_pattern()
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_int8)), shape=(512, 2)).copy()


class Reference(_Base):
    """The reference's own ORBextractor.cc (oracle/_ref).  mono=True -> monotone-allocator build,
    whose heap-address tie-break coincides with the canonical rule."""

    prefix = "reforb_"
    _libs: dict = {}

    @classmethod
    def available(cls, mono: bool = True) -> bool:
        return (HERE / "_ref" / ("libreforb_mono.so" if mono else "libreforb.so")).exists()

    @classmethod
    def load(cls, mono: bool) -> C.CDLL:
        if mono not in cls._libs:
            path = HERE / "_ref" / ("libreforb_mono.so" if mono else "libreforb.so")
            lib = C.CDLL(str(path))
            lib.reforb_pyramid_info.argtypes = [_vp, C.c_int, _vp, _vp]
            lib.reforb_distribute.argtypes = [_vp, _vp, _vp, _vp, C.c_int, C.c_int, C.c_int, C.c_int,
                                              C.c_int, C.c_int, _vp, C.c_int]
            cls._libs[mono] = lib
        return cls._libs[mono]

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7, mono=True):
        super().__init__(self.load(mono), nfeatures, scale_factor, nlevels, ini_th, min_th)

    def level_size(self, level: int):
        c, r = C.c_int(), C.c_int()
        assert self.lib.reforb_pyramid_info(self.h, level, C.byref(c), C.byref(r)) == 0
        return c.value, r.value

    def pyramid(self, level: int, border: int = 19) -> np.ndarray:
        c, r = self.level_size(level)
        out = np.zeros((r + 2 * border, c + 2 * border), np.uint8)
        assert self.lib.reforb_pyramid_copy(self.h, level, border, out.ctypes.data) == 0
        return out

    def distribute(self, xs, ys, resp, min_x, max_x, min_y, max_y, n_target) -> np.ndarray:
        """The reference's own DistributeOctTree on these candidates -> survivor indices, in order."""
        xs = np.ascontiguousarray(xs, np.int32)
        ys = np.ascontiguousarray(ys, np.int32)
        resp = np.ascontiguousarray(resp, np.int32)
        out = np.zeros(len(xs) + 8, np.int32)
        n = self.lib.reforb_distribute(self.h, xs.ctypes.data, ys.ctypes.data, resp.ctypes.data, len(xs),
                                       min_x, max_x, min_y, max_y, n_target, out.ctypes.data, len(out))
        return out[:n].copy()


class _RefFrameOut(C.Structure):
    _fields_ = [("n_left", C.c_int), ("n_right", C.c_int), ("bounds", C.c_float * 4)]


class RefFrame:
    """The reference's own src/Frame.cc (oracle/_ref/libreframe.so, built by oracle/Makefile):
    its stereo constructor run on the reference's own extractor objects."""

    _lib = None

    @classmethod
    def available(cls) -> bool:
        return (HERE / "_ref" / "libreframe.so").exists()

    @classmethod
    def load(cls) -> C.CDLL:
        if cls._lib is None:
            lib = C.CDLL(str(HERE / "_ref" / "libreframe.so"))
            lib.refframe_stereo.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int,
                                            _vp, _vp, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int] + [_vp] * 10
            cls._lib = lib
        return cls._lib

    @staticmethod
    def bind(lib: C.CDLL) -> C.CDLL:
        """argtypes of the refframe_stereo entry point (oracle/frame_driver.cpp) on any library that exports it --
        oracle/_ref/libreframe.so, or tests/shim/libshimframe.so (the same reference Frame.cc over the drop-in class)."""
        lib.refframe_stereo.argtypes = [_vp, _vp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int,
                                        _vp, _vp, C.c_int, C.c_float, C.c_float, C.c_int, C.c_int] + [_vp] * 10
        return lib

    @classmethod
    def stereo(cls, left: np.ndarray, right: np.ndarray, K, dist, bf: float, mb: float, nfeatures=2000,
               scale=1.2, nlevels=8, ini_th=20, min_th=7, canonical=True, lib=None) -> dict:
        """Frame(imLeft, imRight, ...) of the reference -> its mvKeys / mDescriptors (both eyes),
        mvKeysUn positions, mvuRight, mvDepth, mGrid (as CSR) and image bounds.  `mb` is the value the
        uninitialised member holds when ComputeStereoMatches reads it (see oracle/frame_driver.cpp);
        canonical=True runs the two extraction threads with the monotone node allocator, so the
        keypoints are the canonical ones (= oracle = CUDA path)."""
        lib = lib if lib is not None else cls.load()
        left = np.ascontiguousarray(left); right = np.ascontiguousarray(right)
        rows, cols = left.shape
        cap = nfeatures + 64 * nlevels + 4096
        kl = np.zeros(cap, KP_DTYPE); kr = np.zeros(cap, KP_DTYPE)
        dl = np.zeros((cap, 32), np.uint8); dr = np.zeros((cap, 32), np.uint8)
        un = np.zeros((cap, 2), np.float32); u = np.zeros(cap, np.float32); z = np.zeros(cap, np.float32)
        gc = np.zeros(64 * 48, np.int32); gi = np.zeros(cap, np.int32)
        k4 = np.asarray(K, np.float32); d = np.asarray(dist, np.float32)
        out = _RefFrameOut()
        rc = lib.refframe_stereo(left.ctypes.data, right.ctypes.data, cols, rows, nfeatures, scale, nlevels, ini_th,
                                 min_th, k4.ctypes.data, d.ctypes.data, len(d), bf, mb, int(canonical), cap, kl.ctypes.data,
                                 dl.ctypes.data, kr.ctypes.data, dr.ctypes.data, un.ctypes.data, u.ctypes.data,
                                 z.ctypes.data, gc.ctypes.data, gi.ctypes.data, C.addressof(out))
        assert rc == 0
        nl, nr = out.n_left, out.n_right
        start = np.concatenate([[0], np.cumsum(gc)]).astype(np.int32)
        return {"kps_l": kl[:nl].copy(), "desc_l": dl[:nl].copy(), "kps_r": kr[:nr].copy(), "desc_r": dr[:nr].copy(),
                "un_xy": un[:nl].copy(), "u_right": u[:nl].copy(), "depth": z[:nl].copy(), "cell_start": start,
                "cell_idx": gi[:start[-1]].copy(), "bounds": np.array(list(out.bounds), np.float32)}


# ---- bag of words: TemplatedVocabulary::transform / Frame::ComputeBoW / ORBmatcher::SearchByBoW(KeyFrame*, Frame&) ----

def _bow_out(n: int):
    return (np.zeros(max(n, 1), np.uint32), np.zeros(max(n, 1), np.float64), np.zeros(max(n, 1), np.uint32),
            np.zeros(n + 1, np.int32), np.zeros(max(n, 1), np.uint32), C.c_int(0), C.c_int(0))


def _bow_pack(out):
    wid, wv, nid, ns, fi, nw, nn = out
    return {"word_id": wid[:nw.value].copy(), "word_val": wv[:nw.value].copy(), "node_id": nid[:nn.value].copy(),
            "node_start": ns[:nn.value + 1].copy(), "feat_idx": fi[:ns[nn.value]].copy()}


class BowVocabulary:
    """Our restatement (oracle/bow_oracle.cpp in liborb_oracle.so) of DBoW2's vocabulary: loadFromTextFile,
    transform, L1 score, and ORBmatcher::SearchByBoW(KeyFrame*, Frame&) on its feature vectors."""

    def __init__(self, path: str):
        lib = Oracle.load()
        lib.bo_load_text.restype = _vp
        lib.bo_load_text.argtypes = [C.c_char_p]
        lib.bo_free.argtypes = [_vp]
        lib.bo_info.argtypes = [_vp] + [_vp] * 6
        lib.bo_arrays.argtypes = [_vp] * 5
        lib.bo_descend.argtypes = [_vp, _vp, C.c_int, C.c_int, _vp, _vp, _vp]
        lib.bo_transform.argtypes = [_vp, _vp, C.c_int, C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _vp]
        lib.bo_score_l1.restype = C.c_double
        lib.bo_score_l1.argtypes = [_vp, _vp, C.c_int, _vp, _vp, C.c_int]
        lib.bo_search_by_bow.restype = C.c_int
        lib.bo_search_by_bow.argtypes = ([C.c_int] + [_vp] * 6 + [C.c_int] + [C.c_int] + [_vp] * 5 + [C.c_int] +
                                         [C.c_float, C.c_int, _vp])
        self.lib = lib
        self.h = lib.bo_load_text(str(path).encode())
        if not self.h:
            raise ValueError(f"not a vocabulary text file: {path}")
        v = [C.c_int() for _ in range(6)]
        lib.bo_info(self.h, *[C.byref(x) for x in v])
        self.k, self.L, self.n_nodes, self.n_words, self.scoring, self.weighting = [x.value for x in v]

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.bo_free(self.h)
            self.h = None

    def arrays(self):
        """-> parent [N-1] i32, leaf [N-1] u8, desc [N-1, 32] u8, weight [N-1] f64 for nodes 1..N-1"""
        n = self.n_nodes - 1
        parent, leaf = np.zeros(n, np.int32), np.zeros(n, np.uint8)
        desc, weight = np.zeros((n, 32), np.uint8), np.zeros(n, np.float64)
        self.lib.bo_arrays(self.h, _ptr(parent), _ptr(leaf), _ptr(desc), _ptr(weight))
        return parent, leaf, desc, weight

    def descend(self, desc: np.ndarray, levelsup: int = 4):
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(d)
        word, node, w = np.zeros(n, np.int32), np.zeros(n, np.int32), np.zeros(n, np.float64)
        self.lib.bo_descend(self.h, _ptr(d), n, levelsup, _ptr(word), _ptr(node), _ptr(w))
        return word, node, w

    def transform(self, desc: np.ndarray, levelsup: int = 4) -> dict:
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        out = _bow_out(len(d))
        self.lib.bo_transform(self.h, _ptr(d), len(d), levelsup, _ptr(out[0]), _ptr(out[1]), C.byref(out[5]),
                              _ptr(out[2]), _ptr(out[3]), _ptr(out[4]), C.byref(out[6]))
        return _bow_pack(out)

    def score(self, a: dict, b: dict) -> float:
        return float(self.lib.bo_score_l1(_ptr(a["word_id"]), _ptr(a["word_val"]), len(a["word_id"]),
                                          _ptr(b["word_id"]), _ptr(b["word_val"]), len(b["word_id"])))

    def search_by_bow(self, kf_kps, kf_desc, kf_mp, f_kps, f_desc, nnratio=0.7, check_ori=True, kf_fv=None, f_fv=None):
        """-> (match [n_f] i32: keyframe keypoint index or -1, nmatches)"""
        kd = np.ascontiguousarray(kf_desc, np.uint8).reshape(-1, 32)
        fd = np.ascontiguousarray(f_desc, np.uint8).reshape(-1, 32)
        kfv = kf_fv or self.transform(kd, 4)
        ffv = f_fv or self.transform(fd, 4)
        ka = np.ascontiguousarray(kf_kps["angle"], np.float32)
        fa = np.ascontiguousarray(f_kps["angle"], np.float32)
        mp = np.ascontiguousarray(kf_mp, np.uint8)
        out = np.full(max(len(fd), 1), -1, np.int32)
        nm = self.lib.bo_search_by_bow(len(kd), _ptr(ka), _ptr(kd), _ptr(mp), _ptr(kfv["node_id"]), _ptr(kfv["node_start"]),
                                       _ptr(kfv["feat_idx"]), len(kfv["node_id"]), len(fd), _ptr(fa), _ptr(fd),
                                       _ptr(ffv["node_id"]), _ptr(ffv["node_start"]), _ptr(ffv["feat_idx"]),
                                       len(ffv["node_id"]), nnratio, int(check_ori), _ptr(out))
        return out[:len(fd)], int(nm)


def _bow_kf_restated(self, kps1, desc1, mp1, kps2, desc2, mp2, nnratio=0.75, check_ori=True, fv1=None, fv2=None):
    """ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) restated (oracle/bow_oracle.cpp)
    -> (match12 [n1] i32: pKF2 keypoint index or -1, nmatches)"""
    d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
    f1 = fv1 or self.transform(d1, 4)
    f2 = fv2 or self.transform(d2, 4)
    a1 = np.ascontiguousarray(kps1["angle"], np.float32)
    a2 = np.ascontiguousarray(kps2["angle"], np.float32)
    m1, m2 = np.ascontiguousarray(mp1, np.uint8), np.ascontiguousarray(mp2, np.uint8)
    out = np.full(max(len(d1), 1), -1, np.int32)
    fn = self.lib.bo_search_by_bow_kf
    fn.restype = C.c_int
    fn.argtypes = ([C.c_int] + [_vp] * 6 + [C.c_int]) * 2 + [C.c_float, C.c_int, _vp]
    nm = fn(len(d1), _ptr(a1), _ptr(d1), _ptr(m1), _ptr(f1["node_id"]), _ptr(f1["node_start"]), _ptr(f1["feat_idx"]),
            len(f1["node_id"]), len(d2), _ptr(a2), _ptr(d2), _ptr(m2), _ptr(f2["node_id"]), _ptr(f2["node_start"]),
            _ptr(f2["feat_idx"]), len(f2["node_id"]), nnratio, int(check_ori), _ptr(out))
    return out[:len(d1)], int(nm)


BowVocabulary.search_by_bow_kf = _bow_kf_restated


def camera_centre(Tcw) -> np.ndarray:
    """Ow = -Rcw.t() * tcw as cv::MatExpr evaluates it (cv::gemm GEMM_1_T: double accumulation, one rounding)"""
    T = np.asarray(Tcw, np.float32)
    out = np.zeros(3, np.float32)
    for y in range(3):
        s = 0.0
        for k in range(3):
            s += float(T[k, y]) * float(T[k, 3])
        out[y] = np.float32(s * -1.0)
    return out


def _bow_triangulation_restated(self, kf1: dict, kf2: dict, geo: dict, only_stereo=False, check_ori=True, fv1=None, fv2=None):
    """ORBmatcher::SearchForTriangulation restated (oracle/bow_oracle.cpp); arguments as RefBow.search_for_triangulation"""
    keep = []

    def arr(a, dt):
        a = np.ascontiguousarray(a, dt)
        keep.append(a)
        return a.ctypes.data
    d1 = np.ascontiguousarray(kf1["desc"], np.uint8).reshape(-1, 32)
    d2 = np.ascontiguousarray(kf2["desc"], np.uint8).reshape(-1, 32)
    f1 = fv1 or self.transform(d1, 4)
    f2 = fv2 or self.transform(d2, 4)
    fn = self.lib.bo_search_for_triangulation
    fn.restype = C.c_int
    fn.argtypes = ([C.c_int] + [_vp] * 8 + [C.c_int]) * 2 + [_vp] * 4 + [C.c_int, C.c_int, _vp]
    n1 = len(d1)
    out = np.full(max(n1, 1), -1, np.int32)

    def side(k, d, f, pose):
        return [len(d), arr(k["kps"], KP_DTYPE), arr(d, np.uint8), arr(k["mp"], np.uint8),
                arr(k["u_right"], np.float32) if k.get("u_right") is not None else None, pose,
                arr(f["node_id"], np.uint32), arr(f["node_start"], np.int32), arr(f["feat_idx"], np.uint32), len(f["node_id"])]
    nm = fn(*side(kf1, d1, f1, arr(camera_centre(kf1["Tcw"]), np.float32)), *side(kf2, d2, f2, arr(kf2["Tcw"], np.float32)),
            arr(geo["cam2"], np.float32), arr(geo["scale2"], np.float32), arr(geo["sigma2_2"], np.float32),
            arr(geo["F12"], np.float32), int(only_stereo), int(check_ori), _ptr(out))
    return out[:n1], int(nm)


BowVocabulary.search_for_triangulation = _bow_triangulation_restated


class RefBow:
    """The reference's own DBoW2 + Frame::ComputeBoW + ORBmatcher::SearchByBoW (oracle/_ref/librefbow.so)."""

    _lib = None

    @classmethod
    def available(cls) -> bool:
        return (HERE / "_ref" / "librefbow.so").exists()

    def __init__(self, path: str):
        if RefBow._lib is None:
            lib = C.CDLL(str(HERE / "_ref" / "librefbow.so"))
            lib.refbow_load.restype = _vp
            lib.refbow_load.argtypes = [C.c_char_p]
            lib.refbow_free.argtypes = [_vp]
            lib.refbow_info.argtypes = [_vp] * 6
            lib.refbow_compute_bow.argtypes = [_vp, _vp, C.c_int, C.c_int] + [_vp] * 7
            lib.refbow_score.restype = C.c_double
            lib.refbow_score.argtypes = [_vp, _vp, _vp, C.c_int, _vp, _vp, C.c_int]
            lib.refbow_search_by_bow.restype = C.c_int
            lib.refbow_search_by_bow.argtypes = [_vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, _vp, C.c_float, C.c_int, _vp]
            RefBow._lib = lib
        self.lib = RefBow._lib
        self.h = self.lib.refbow_load(str(path).encode())
        if not self.h:
            raise ValueError(f"the reference refused {path}")
        v = [C.c_int() for _ in range(5)]
        self.lib.refbow_info(self.h, *[C.byref(x) for x in v])
        self.k, self.L, self.n_words, self.scoring, self.weighting = [x.value for x in v]

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.refbow_free(self.h)
            self.h = None

    def transform(self, desc: np.ndarray, levelsup: int = 4) -> dict:
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        out = _bow_out(len(d))
        self.lib.refbow_compute_bow(self.h, _ptr(d), len(d), levelsup, _ptr(out[0]), _ptr(out[1]), C.byref(out[5]),
                                    _ptr(out[2]), _ptr(out[3]), _ptr(out[4]), C.byref(out[6]))
        return _bow_pack(out)

    def score(self, a: dict, b: dict) -> float:
        return float(self.lib.refbow_score(self.h, _ptr(a["word_id"]), _ptr(a["word_val"]), len(a["word_id"]),
                                           _ptr(b["word_id"]), _ptr(b["word_val"]), len(b["word_id"])))

    def search_by_bow(self, kf_kps, kf_desc, kf_mp, f_kps, f_desc, nnratio=0.7, check_ori=True):
        kd = np.ascontiguousarray(kf_desc, np.uint8).reshape(-1, 32)
        fd = np.ascontiguousarray(f_desc, np.uint8).reshape(-1, 32)
        kk, fk = np.ascontiguousarray(kf_kps), np.ascontiguousarray(f_kps)
        mp = np.ascontiguousarray(kf_mp, np.uint8)
        out = np.full(max(len(fd), 1), -1, np.int32)
        nm = self.lib.refbow_search_by_bow(self.h, len(kd), _ptr(kk), _ptr(kd), _ptr(mp), len(fd), _ptr(fk), _ptr(fd),
                                           nnratio, int(check_ori), _ptr(out))
        return out[:len(fd)], int(nm)

    def search_by_bow_kf(self, kps1, desc1, mp1, kps2, desc2, mp2, nnratio=0.75, check_ori=True):
        """ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) -> (match12 [n1] i32: pKF2 keypoint index or -1, nmatches)"""
        d1 = np.ascontiguousarray(desc1, np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(desc2, np.uint8).reshape(-1, 32)
        k1, k2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
        m1, m2 = np.ascontiguousarray(mp1, np.uint8), np.ascontiguousarray(mp2, np.uint8)
        out = np.full(max(len(d1), 1), -1, np.int32)
        fn = self.lib.refbow_search_by_bow_kf
        fn.restype = C.c_int
        fn.argtypes = [_vp, C.c_int, _vp, _vp, _vp, C.c_int, _vp, _vp, _vp, C.c_float, C.c_int, _vp]
        nm = fn(self.h, len(d1), _ptr(k1), _ptr(d1), _ptr(m1), len(d2), _ptr(k2), _ptr(d2), _ptr(m2), nnratio, int(check_ori),
                _ptr(out))
        return out[:len(d1)], int(nm)

    def search_for_triangulation(self, kf1: dict, kf2: dict, geo: dict, only_stereo=False, check_ori=True):
        """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:779-966).  kf: kps (KP_DTYPE, mvKeysUn), desc, mp [n] u8,
        u_right [n] f32 or None, Tcw [3,4]; geo: cam2 (fx, fy, cx, cy), scale2, sigma2_2, F12 [3,3].
        -> (match12 [n1] i32, nmatches)"""
        keep = []

        def arr(a, dt):
            a = np.ascontiguousarray(a, dt)
            keep.append(a)
            return a.ctypes.data
        fn = self.lib.refbow_search_for_triangulation
        fn.restype = C.c_int
        fn.argtypes = [_vp] + [C.c_int, _vp, _vp, _vp, _vp, _vp] * 2 + [_vp, C.c_int, _vp, _vp, _vp, C.c_int, C.c_int, _vp]
        n1, n2 = len(kf1["kps"]), len(kf2["kps"])
        out = np.full(max(n1, 1), -1, np.int32)
        side = lambda k: [len(k["kps"]), arr(k["kps"], KP_DTYPE), arr(np.asarray(k["desc"]).reshape(-1, 32), np.uint8),
                          arr(k["mp"], np.uint8), arr(k["u_right"], np.float32) if k.get("u_right") is not None else None,
                          arr(k["Tcw"], np.float32)]
        nm = fn(self.h, *side(kf1), *side(kf2), arr(geo["cam2"], np.float32), len(geo["scale2"]), arr(geo["scale2"], np.float32),
                arr(geo["sigma2_2"], np.float32), arr(geo["F12"], np.float32), int(only_stereo), int(check_ori), _ptr(out))
        return out[:n1], int(nm)
