"""Scenario builder for the frame-to-frame matcher row (ORBmatcher::SearchByProjection(Frame, Frame),
src/ORBmatcher.cc:1503-1668): two consecutive synthetic stereo frames, the last frame's map points
un-projected from its stereo depths (Frame::UnprojectStereo, src/Frame.cc:836-850), and a current pose
that predicts the (3, 1) px shift of the synthetic sequence.  TEST INFRASTRUCTURE (uses the oracle)."""
from __future__ import annotations

import numpy as np

from orb_slam2_test_b200 import synth

KITTI_K = (718.856, 718.856, 607.1928, 185.2157)
KITTI_BF = 386.1448
TUM1_K = (517.306408, 516.469215, 318.643040, 255.313989)
TUM1_D = (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)


def _rot(rx: float, ry: float, rz: float) -> np.ndarray:
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def pose(rx=0.0, ry=0.0, rz=0.0, t=(0.0, 0.0, 0.0)) -> np.ndarray:
    T = np.zeros((3, 4))
    T[:, :3] = _rot(rx, ry, rz)
    T[:, 3] = t
    return T.astype(np.float32)


def compose(A: np.ndarray, B: np.ndarray) -> np.ndarray:
    """Tcw of A after B (A * B as 4x4)."""
    A4 = np.vstack([A.astype(np.float64), [0, 0, 0, 1]])
    B4 = np.vstack([B.astype(np.float64), [0, 0, 0, 1]])
    return (A4 @ B4)[:3].astype(np.float32)


def oracle_frame(left, right, K, dist, bf, mb, nf=2000):
    """One stereo (or mono, right=None) frame through the CPU oracle: the fields a Frame holds."""
    from oracle import pyoracle
    rows, cols = left.shape
    ol = pyoracle.Oracle(nf, 1.2, 8, 20, 7)
    kl, dl = ol.extract(left)
    un, start, idx, bounds = pyoracle.frame_post(kl, cols, rows, K, dist)
    scale = ol.table(0)
    f = {"kps": kl, "desc": dl, "un_xy": un, "cell_start": start, "cell_idx": idx, "bounds": bounds, "scale": scale,
         "u_right": None, "depth": None}
    if right is not None:
        orr = pyoracle.Oracle(nf, 1.2, 8, 20, 7)
        kr, dr = orr.extract(right)
        pl = [ol.pyramid(l, 0) for l in range(8)]
        pr = [orr.pyramid(l, 0) for l in range(8)]
        u, z, _, _ = pyoracle.stereo_match(kl, dl, kr, dr, pl, pr, scale, ol.table(1), bf, mb)
        f["u_right"], f["depth"] = u, z
    return f


def map_points(last: dict, Tcw_last: np.ndarray, K, rng: np.random.Generator, obs_rate=0.7, outlier_rate=0.05,
               flip_bits=6, synth_depth=None, copies=1):
    """Map points for the last frame: keypoint i with depth z > 0 (or a made-up depth when the frame
    is monocular) is un-projected with the last pose; its descriptor is the keypoint's with a few bits
    flipped.  copies > 1 repeats every entry (crowding: several queries want the same keypoint)."""
    fx, fy, cx, cy = K
    n = len(last["kps"])
    z = last["depth"] if last["depth"] is not None else np.full(n, -1.0, np.float32)
    z = z.astype(np.float64).copy()
    if synth_depth is not None:
        lo, hi = synth_depth
        fill = rng.uniform(lo, hi, n)
        z = np.where(z > 0, z, np.where(rng.random(n) < 0.8, fill, -1.0))
    has = z > 0
    u, v = last["un_xy"][:, 0].astype(np.float64), last["un_xy"][:, 1].astype(np.float64)
    Xc = np.stack([(u - cx) * z / fx, (v - cy) * z / fy, z], 1)
    R, t = Tcw_last[:, :3].astype(np.float64), Tcw_last[:, 3].astype(np.float64)
    Xw = (Xc - t) @ R  # R^T (Xc - t)
    flags = (has.astype(np.uint8) | ((rng.random(n) < obs_rate).astype(np.uint8) << 1)
             | ((rng.random(n) < outlier_rate).astype(np.uint8) << 2))
    desc = last["desc"].copy()
    for i in range(n):
        for b in rng.integers(0, 256, rng.integers(0, flip_bits + 1)):
            desc[i, b >> 3] ^= np.uint8(1 << (b & 7))
    out = dict(last)
    out.update({"mp_flags": flags, "mp_xyz": np.where(has[:, None], Xw, 0).astype(np.float32), "mp_desc": desc,
                "Tcw": Tcw_last})
    if copies > 1:
        for k in ("kps", "desc", "un_xy", "mp_flags", "mp_xyz", "mp_desc"):
            out[k] = np.concatenate([out[k]] * copies)
        # the copies differ a little so that they do not all tie
        jitter = rng.normal(0, 0.02, out["mp_xyz"].shape).astype(np.float32)
        out["mp_xyz"] = (out["mp_xyz"] + jitter * (out["mp_xyz"][:, 2:3] > 0)).astype(np.float32)
        out["mp_flags"] = out["mp_flags"].copy()
        out["mp_flags"][n:] = (out["mp_flags"][n:] & 5) | ((rng.random(len(out["mp_flags"]) - n) < obs_rate).astype(np.uint8) << 1)
    return out


def make_case(seed=0, shape=synth.KITTI_SHAPE, K=KITTI_K, dist=(0.0, 0.0, 0.0, 0.0), bf=KITTI_BF, stereo=True,
              dz=0.0, nf=2000, obs_rate=0.7, outlier_rate=0.05, copies=1, frame=0, mb=None):
    """-> (cur, last, cam, bounds, scale): dicts as oracle.pyoracle.search_by_projection takes them."""
    rng = np.random.default_rng(1000 + seed)
    mb = bf / K[0] if mb is None else mb
    imgs = {(f, e): synth.make_frame(shape, seed=seed, frame=frame + f, eye=e) for f in (0, 1) for e in ((0, 1) if stereo else (0,))}
    last = oracle_frame(imgs[(0, 0)], imgs.get((0, 1)), K, dist, bf, mb, nf)
    cur = oracle_frame(imgs[(1, 0)], imgs.get((1, 1)), K, dist, bf, mb, nf)
    T_last = pose(0.01, -0.02, 0.005, (0.3, -0.1, 0.2))
    # frame i+1 is frame i shifted by (3, 1) px: a small yaw / pitch of the camera predicts that shift
    rel = pose(rx=1.0 / K[1], ry=-3.0 / K[0], t=(0.0, 0.0, -dz))
    T_cur = compose(rel, T_last)
    last = map_points(last, T_last, K, rng, obs_rate, outlier_rate, synth_depth=None if stereo else (4.0, 60.0), copies=copies)
    cur["Tcw"] = T_cur
    cam = np.array([K[0], K[1], K[2], K[3], bf, mb], np.float32)
    return cur, last, cam, cur["bounds"], cur["scale"]


def make_local_case(seed=0, copies=2, obs_rate=0.9, **kw):
    """Scenario for ORBmatcher::SearchByProjection(F, vpMapPoints, th) (src/ORBmatcher.cc:59-154), the matcher of
    Tracking::SearchLocalPoints: the current frame after the motion-model matcher has run (its matches occupy
    slots), and a local map = the last frame's map points not matched yet plus jittered copies, projected with
    the current pose the way Frame::isInFrustum leaves them (mTrackProjX/Y/XR, mTrackViewCos, mnTrackScaleLevel).
    -> (cur, bounds, scale, mps, occupied)"""
    from oracle import pyoracle
    cur, last, cam, bounds, scale = make_case(seed=seed, **kw)
    rng = np.random.default_rng(5000 + seed)
    m, _ = pyoracle.search_by_projection(cur, last, cam, bounds, scale, 7.0 if kw.get("stereo", True) else 15.0,
                                         mono=not kw.get("stereo", True))
    occupied = np.zeros(len(m), np.uint8)
    held = m >= 0
    occupied[held] = np.where(last["mp_flags"][m[held]] & 2, 2, 1)
    n = len(last["kps"])
    matched = np.zeros(n, bool)
    matched[m[held]] = True
    xyz = np.concatenate([last["mp_xyz"]] + [last["mp_xyz"] + rng.normal(0, 0.03, last["mp_xyz"].shape).astype(np.float32)
                                             for _ in range(copies - 1)]).astype(np.float64)
    has = np.concatenate([(last["mp_flags"] & 1).astype(bool) & ~matched] + [(last["mp_flags"] & 1).astype(bool)] * (copies - 1))
    desc = np.concatenate([last["mp_desc"]] * copies)
    octv = np.concatenate([last["kps"]["octave"]] * copies)
    T = cur["Tcw"].astype(np.float64)
    Xc = xyz @ T[:, :3].T + T[:, 3]
    z = np.where(np.abs(Xc[:, 2]) > 1e-6, Xc[:, 2], 1.0)
    u = cam[0] * Xc[:, 0] / z + cam[2]
    v = cam[1] * Xc[:, 1] / z + cam[3]
    inside = (Xc[:, 2] > 0) & (u > bounds[0]) & (u < bounds[1]) & (v > bounds[2]) & (v < bounds[3])
    N = len(xyz)
    proj = np.stack([u, v, u - cam[4] / z, rng.uniform(0.99, 1.0, N)], 1).astype(np.float32)
    proj[~np.isfinite(proj)] = 0
    level = np.clip(octv + rng.integers(-1, 2, N), 0, 7).astype(np.int32)
    flags = ((has & inside).astype(np.uint8) | ((rng.random(N) < obs_rate).astype(np.uint8) << 1))
    mps = {"proj": proj, "desc": desc, "flags": flags, "level": level}
    return cur, bounds, scale, mps, occupied


def make_init_case(seed=0, shape=synth.TUM_SHAPE, K=TUM1_K, dist=TUM1_D, nf=2000, step=1, frame=0, copies=1):
    """Scenario for ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:487-632): the monocular-initialisation
    extractor (2 x nFeatures, src/Tracking.cc:124-125) on an initial frame and on a later frame of the same
    synthetic sequence; vbPrevMatched starts as the initial frame's undistorted keypoints (src/Tracking.cc:587-589).
    -> (f1, f2, bounds, prev)"""
    f1 = oracle_frame(synth.make_frame(shape, seed=seed, frame=frame), None, K, dist, 0.0, 0.0, nf)
    f2 = oracle_frame(synth.make_frame(shape, seed=seed, frame=frame + step), None, K, dist, 0.0, 0.0, nf)
    if copies > 1:  # several initial keypoints want the same current keypoint, at different distances: take-overs
        rng = np.random.default_rng(7000 + seed)
        n = len(f1["kps"])
        desc = [f1["desc"]]
        for _ in range(copies - 1):
            d = f1["desc"].copy()
            for i in range(n):
                for b in rng.integers(0, 256, rng.integers(0, 9)):
                    d[i, b >> 3] ^= np.uint8(1 << (b & 7))
            desc.append(d)
        f1 = dict(f1)
        f1["kps"] = np.concatenate([f1["kps"]] * copies)
        f1["un_xy"] = np.concatenate([f1["un_xy"]] * copies)
        f1["desc"] = np.concatenate(desc)
    return f1, f2, f2["bounds"], f1["un_xy"].copy()


def make_reloc_case(seed=0, copies=1, found_rate=0.2, bad_rate=0.05, occupied_rate=0.15, dist_jitter=0.25, **kw):
    """Scenario for the relocalisation matcher ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist)
    (src/ORBmatcher.cc:1670-1798): the keyframe is the previous frame of make_case with its map points (copies > 1 crowds the
    queries), some of them bad, some already found by the PnP step; the scale-invariance distances are what
    MapPoint::UpdateNormalAndDepth (src/MapPoint.cc) would give for the keyframe's own observation, jittered so that a share
    of the points fails the distance test and the predicted level moves; some slots of the current frame are occupied.
    -> (cur, cam, bounds, scale, log_scale, kf, occupied); kf: flags, xyz, desc, dist [n,2] = (mfMinDistance, mfMaxDistance), angle."""
    cur, last, cam, bounds, scale = make_case(seed=seed, copies=copies, **kw)
    rng = np.random.default_rng(9000 + seed)
    n = len(last["mp_flags"])
    T = last["Tcw"].astype(np.float64)
    Ow = -T[:, :3].T @ T[:, 3]
    d = np.linalg.norm(last["mp_xyz"].astype(np.float64) - Ow, axis=1)
    sc = np.asarray(scale, np.float64)
    max_d = d * sc[last["kps"]["octave"]] * rng.uniform(1 - dist_jitter, 1 + dist_jitter, n)
    min_d = max_d / sc[-1]
    flags = ((last["mp_flags"] & 1) | ((rng.random(n) < bad_rate).astype(np.uint8) << 1)
             | ((rng.random(n) < found_rate).astype(np.uint8) << 2)).astype(np.uint8)
    kf = {"flags": flags, "xyz": last["mp_xyz"], "desc": last["mp_desc"],
          "dist": np.stack([min_d, max_d], 1).astype(np.float32), "angle": last["kps"]["angle"].astype(np.float32)}
    occupied = (rng.random(len(cur["kps"])) < occupied_rate).astype(np.uint8)
    # Frame: mfLogScaleFactor = log(mfScaleFactor) on floats = the host libm's logf
    import ctypes
    import ctypes.util
    libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.logf.restype = ctypes.c_float
    libm.logf.argtypes = [ctypes.c_float]
    log_scale = libm.logf(float(np.float32(1.2)))
    return cur, cam, bounds, scale, float(log_scale), kf, occupied


def host_logf(x: float) -> float:
    import ctypes
    import ctypes.util
    libm = ctypes.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    libm.logf.restype = ctypes.c_float
    libm.logf.argtypes = [ctypes.c_float]
    return float(libm.logf(float(np.float32(x))))


def make_fuse_case(seed=0, copies=1, bad_rate=0.05, in_kf_rate=0.1, dist_jitter=0.25, normal_tilt=0.6, **kw):
    """Scenario for ORBmatcher::Fuse(pKF, vpMapPoints, th) (src/ORBmatcher.cc:968-1131): the keyframe is the current frame of
    make_case (stereo keypoints where the scenario is stereo), the map points are the previous frame's (a neighbour
    keyframe's), with scale-invariance distances as in make_reloc_case and viewing normals = the direction they were seen
    from, tilted at random so that some fail the 60 degree test.
    -> (kf, cam, bounds, scale, inv_sigma2, log_scale, mps)"""
    cur, last, cam, bounds, scale = make_case(seed=seed, copies=copies, **kw)
    rng = np.random.default_rng(11000 + seed)
    n = len(last["mp_flags"])
    T = last["Tcw"].astype(np.float64)
    Ow = -T[:, :3].T @ T[:, 3]
    PO = last["mp_xyz"].astype(np.float64) - Ow
    d = np.linalg.norm(PO, axis=1)
    sc = np.asarray(scale, np.float64)
    max_d = d * sc[last["kps"]["octave"]] * rng.uniform(1 - dist_jitter, 1 + dist_jitter, n)
    nrm = PO / np.maximum(d, 1e-9)[:, None] + rng.normal(0, normal_tilt, (n, 3))
    nrm /= np.maximum(np.linalg.norm(nrm, axis=1), 1e-9)[:, None]
    flags = ((last["mp_flags"] & 1) | ((rng.random(n) < bad_rate).astype(np.uint8) << 1)
             | ((rng.random(n) < in_kf_rate).astype(np.uint8) << 2)).astype(np.uint8)
    mps = {"flags": flags, "xyz": last["mp_xyz"], "desc": last["mp_desc"],
           "dist": np.stack([max_d / sc[-1], max_d], 1).astype(np.float32), "normal": nrm.astype(np.float32)}
    sc32 = np.asarray(scale, np.float32)
    sigma2 = (sc32 * sc32).astype(np.float32)            # mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i]
    inv_sigma2 = (np.float32(1.0) / sigma2).astype(np.float32)   # mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i]
    return cur, cam, bounds, scale, inv_sigma2, host_logf(1.2), mps


def sim3_of(Tcw: np.ndarray, s: float) -> np.ndarray:
    """the 4x4 similarity [s R | s t; 0 0 0 1] of a pose (what Converter::toCvMat(g2o::Sim3) hands LoopClosing's matchers)"""
    S = np.eye(4, dtype=np.float32)
    S[:3, :3] = (np.float32(s) * Tcw[:, :3].astype(np.float32)).astype(np.float32)
    S[:3, 3] = (np.float32(s) * Tcw[:, 3].astype(np.float32)).astype(np.float32)
    return S


def make_sim3_points_case(seed=0, s=1.03, found_rate=0.1, occupied_rate=0.2, **kw):
    """Scenario for the two Scw matchers of LoopClosing -- Fuse(pKF, Scw, vpPoints, th, vpReplacePoint)
    (src/ORBmatcher.cc:1133-1260) and SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:353-485): make_fuse_case's
    keyframe and points, the keyframe's pose given as a similarity of scale s; bit 2 of the point flags = already found
    (in pKF->GetMapPoints() / in vpMatched), and enough occupied vpMatched slots to hold those.
    -> (kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occupied)"""
    kf, cam, bounds, scale, inv_sigma2, log_scale, mps = make_fuse_case(seed=seed, in_kf_rate=found_rate, **kw)
    rng = np.random.default_rng(12000 + seed)
    Scw = sim3_of(kf["Tcw"], s)
    occupied = (rng.random(len(kf["kps"])) < occupied_rate).astype(np.uint8)
    pts = dict(mps)
    pts["flags"] = (mps["flags"] | 1).astype(np.uint8)  # vpPoints holds no NULL entries
    found = np.nonzero(pts["flags"] & 4)[0]
    pts["flags"][found[int(occupied.sum()):]] &= np.uint8(3)  # vpMatched can only hold as many points as it has set slots
    return kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occupied


def make_sim3_pair_case(seed=0, s12=1.02, pre_rate=0.1, bad_rate=0.05, dist_jitter=0.25, **kw):
    """Scenario for ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1262-1501): keyframe 1 = the previous frame of make_case,
    keyframe 2 = the current one; each keypoint with a depth carries a map point (un-projected with its own keyframe's
    pose, as map_points does for the previous frame); the relative similarity is the true relative pose scaled by s12;
    pre_rate of pKF1's points are matched already (vpMatches12 on entry).
    -> (k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12)"""
    K = kw.get("K", KITTI_K)
    cur, last, cam, bounds, scale = make_case(seed=seed, **kw)
    rng = np.random.default_rng(13000 + seed)
    k1 = {k: last[k] for k in ("kps", "desc", "un_xy", "cell_start", "cell_idx", "u_right", "Tcw")}
    cur_mp = map_points(cur, cur["Tcw"], K, rng, synth_depth=None if kw.get("stereo", True) else (4.0, 60.0))
    k2 = {k: cur[k] for k in ("kps", "desc", "un_xy", "cell_start", "cell_idx", "u_right", "Tcw")}
    sc = np.asarray(scale, np.float64)

    def pts(fr, mp):
        n = len(mp["mp_flags"])
        T = fr["Tcw"].astype(np.float64)
        Ow = -T[:, :3].T @ T[:, 3]
        d = np.linalg.norm(mp["mp_xyz"].astype(np.float64) - Ow, axis=1)
        max_d = d * sc[fr["kps"]["octave"]] * rng.uniform(1 - dist_jitter, 1 + dist_jitter, n)
        flags = ((mp["mp_flags"] & 1) | ((rng.random(n) < bad_rate).astype(np.uint8) << 1)).astype(np.uint8)
        return {"flags": flags, "xyz": mp["mp_xyz"], "desc": mp["mp_desc"], "dist": np.stack([max_d / sc[-1], max_d], 1).astype(np.float32)}
    p1, p2 = pts(k1, last), pts(k2, cur_mp)
    T1 = np.vstack([k1["Tcw"].astype(np.float64), [0, 0, 0, 1]])
    T2 = np.vstack([k2["Tcw"].astype(np.float64), [0, 0, 0, 1]])
    T12 = T1 @ np.linalg.inv(T2)  # camera 2 -> camera 1
    R12 = T12[:3, :3].astype(np.float32)
    t12 = (T12[:3, 3] * s12).astype(np.float32)
    n1, n2 = len(p1["flags"]), len(p2["flags"])
    pre12 = np.full(n1, -1, np.int32)
    have2 = np.nonzero(p2["flags"] & 1)[0]
    pick = np.nonzero((rng.random(n1) < pre_rate) & ((p1["flags"] & 1) != 0))[0]
    pre12[pick] = rng.choice(have2, len(pick), replace=False) if len(pick) <= len(have2) else -1
    sc32 = np.asarray(scale, np.float32)
    inv_sigma2 = (np.float32(1.0) / (sc32 * sc32)).astype(np.float32)
    return k1, k2, cam, bounds, scale, inv_sigma2, host_logf(1.2), p1, p2, pre12, float(np.float32(s12)), R12, t12
