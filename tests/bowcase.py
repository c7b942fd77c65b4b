"""Scenario builder for the bag-of-words rows (Frame::ComputeBoW src/Frame.cc:532-539 -> DBoW2 transform;
ORBmatcher::SearchByBoW(KeyFrame*, Frame&) src/ORBmatcher.cc:195-351): synthetic vocabularies in the reference's text
format (tests/golden/make_vocabulary.py) and ORB descriptors of consecutive synthetic frames.  TEST INFRASTRUCTURE."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

from orb_slam2_test_b200 import synth

sys.path.insert(0, str(Path(__file__).resolve().parent / "golden"))
import make_vocabulary  # noqa: E402

# (name, k, L, seed, ragged, scoring, weighting): the reference's ORBvoc.txt is k=10 L=6 L1 TF-IDF
VOCABULARIES = [
    ("k10L4", 10, 4, 1, False, 0, 0),
    ("k9L5ragged", 9, 5, 2, True, 0, 0),
    ("k4L6", 4, 6, 3, False, 0, 0),
    ("k10L3", 10, 3, 4, False, 0, 0),
]
_cache: dict = {}


def vocabulary(tmpdir, name="k10L4", scoring=None, weighting=None) -> str:
    """Path of the named vocabulary file (written once per session into tmpdir)."""
    spec = next(v for v in VOCABULARIES if v[0] == name)
    _, k, L, seed, ragged, sc, we = spec
    sc = sc if scoring is None else scoring
    we = we if weighting is None else weighting
    path = Path(tmpdir) / f"voc_{name}_{sc}_{we}.txt"
    if str(path) not in _cache:
        make_vocabulary.write_vocabulary(str(path), k, L, seed, sc, we, ragged)
        _cache[str(path)] = True
    return str(path)


def tree(name="k10L4"):
    _, k, L, seed, ragged, _, _ = next(v for v in VOCABULARIES if v[0] == name)
    return make_vocabulary.make_tree(k, L, seed, ragged)


def frame_features(shape=synth.TUM_SHAPE, seed=0, frame=0, nf=1000):
    """(keypoints, descriptors) of a synthetic frame through the CPU extractor oracle."""
    from oracle import pyoracle
    key = ("ff", shape, seed, frame, nf)
    if key not in _cache:
        _cache[key] = pyoracle.Oracle(nf, 1.2, 8, 20, 7).extract(synth.make_frame(shape, seed=seed, frame=frame))
    return _cache[key]


def word_like(name: str, n: int, seed: int, noise_bits=24) -> np.ndarray:
    """n descriptors near the vocabulary's words (what a trained vocabulary sees)."""
    _, leaf, desc, _ = tree(name)
    return make_vocabulary.words_for(desc, leaf, n, seed, noise_bits)


def keyframe_flags(n: int, seed: int, has=0.7, bad=0.05) -> np.ndarray:
    """per keyframe keypoint: bit 0 = has a map point, bit 1 = that map point isBad()"""
    rng = np.random.default_rng(seed)
    f = (rng.random(n) < has).astype(np.uint8)
    f |= ((rng.random(n) < bad) & (f > 0)).astype(np.uint8) << 1
    return f


def same_bow(a: dict, b: dict) -> bool:
    return (np.array_equal(a["word_id"], b["word_id"]) and
            np.array_equal(a["word_val"].view(np.uint64), b["word_val"].view(np.uint64)) and
            np.array_equal(a["node_id"], b["node_id"]) and np.array_equal(a["node_start"], b["node_start"]) and
            np.array_equal(a["feat_idx"], b["feat_idx"]))


def triangulation_case(shape=synth.TUM_SHAPE, seed=3, t=0, nf=1000, baseline=(0.08, 0.0, 0.02), stereo_rate=0.0, has=0.5,
                       K=(517.306408, 516.469215, 318.643040, 255.313989)):
    """Scenario for ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:779-966): keyframe 1 = frame t, keyframe 2 =
    frame t + 1 of the synthetic sequence (shifted by (3, 1) px: a small yaw / pitch of the camera), plus a translation
    `baseline` that gives the pair an epipolar geometry; F12 as LocalMapping::ComputeF12 builds it (computed in double
    here: the matcher takes it as an argument).  `has` of the keypoints of either side already carry a map point (left
    out by the matcher), `stereo_rate` of them have mvuRight >= 0.
    -> (kf1, kf2, geo) as oracle.pyoracle.RefBow.search_for_triangulation takes them"""
    import matchcase
    k1, d1 = frame_features(shape, seed, t, nf)
    k2, d2 = frame_features(shape, seed, t + 1, nf)
    rng = np.random.default_rng(8000 + seed + 17 * t)
    T1 = matchcase.pose(0.01, -0.02, 0.005, (0.3, -0.1, 0.2))
    rel = matchcase.pose(rx=1.0 / K[1], ry=-3.0 / K[0], t=baseline)
    T2 = matchcase.compose(rel, T1)
    Km = np.array([[K[0], 0, K[2]], [0, K[1], K[3]], [0, 0, 1]], np.float64)
    R1, t1 = T1[:, :3].astype(np.float64), T1[:, 3].astype(np.float64)
    R2, t2 = T2[:, :3].astype(np.float64), T2[:, 3].astype(np.float64)
    R12 = R1 @ R2.T
    t12 = -R1 @ R2.T @ t2 + t1
    tx = np.array([[0, -t12[2], t12[1]], [t12[2], 0, -t12[0]], [-t12[1], t12[0], 0]])
    F12 = np.linalg.inv(Km.T) @ tx @ R12 @ np.linalg.inv(Km)
    F12 = (F12 / np.abs(F12).max()).astype(np.float32)

    def side(k, d, T, s):
        n = len(k)
        ur = np.where(rng.random(n) < stereo_rate, k["x"] - rng.uniform(2, 40, n).astype(np.float32), -1.0).astype(np.float32)
        return {"kps": k, "desc": d, "mp": (rng.random(n) < has).astype(np.uint8), "u_right": ur if stereo_rate > 0 else None,
                "Tcw": T}
    sc = np.float32(1.2) ** np.arange(8, dtype=np.float32)
    scale = np.ones(8, np.float32)
    for i in range(1, 8):
        scale[i] = np.float32(np.float64(scale[i - 1]) * 1.2)  # mvScaleFactor[i] = mvScaleFactor[i-1] * scaleFactor (double member)
    geo = {"cam2": np.array(K, np.float32), "scale2": scale, "sigma2_2": (scale * scale).astype(np.float32), "F12": F12}
    return side(k1, d1, T1, 0), side(k2, d2, T2, 1), geo
