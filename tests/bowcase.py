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
