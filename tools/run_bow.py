"""One pass over the bag-of-words kernels for ncu launch lists: B1 k_bow_descend, B2 k_bow_assemble (Frame::ComputeBoW),
B3 k_bow_search, B4 k_bow_finish (ORBmatcher::SearchByBoW(KeyFrame*, Frame&)) on a KITTI-shape frame and a synthetic
k = 10, L = 5 vocabulary (tests/golden/make_vocabulary.py).
usage: python tools/run_bow.py"""
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests" / "golden"))
import make_vocabulary  # noqa: E402
from orb_slam2_test_b200 import ORBextractor, ORBVocabulary, synth  # noqa: E402

with tempfile.TemporaryDirectory() as td:
    path = str(Path(td) / "voc.txt")
    n = make_vocabulary.write_vocabulary(path, 10, 5, seed=5)
    voc = ORBVocabulary(path)
    ext = ORBextractor(2000, 1.2, 8, 20, 7)
    kk, kd = ext(synth.make_frame(synth.KITTI_SHAPE, seed=81, frame=0))
    kk, kd = kk.copy(), kd.copy()
    ext(synth.make_frame(synth.KITTI_SHAPE, seed=81, frame=1))
    kf = {"desc": kd, "angle": kk["angle"], "mp_flags": (np.random.default_rng(81).random(len(kk)) < 0.7).astype(np.uint8),
          "fv": voc.transform(kd, 4)}
    for levelsup in (4, 3):  # 4: the reference's constant (10 nodes with this L = 5 tree); 3: 100 nodes, ORBvoc's granularity
        kf["fv"] = voc.transform(kd, levelsup)
        ts = []
        for _ in range(20):
            t0 = time.perf_counter()
            bow = ext.compute_bow(voc, levelsup=levelsup)
            t1 = time.perf_counter()
            m, nm = ext.search_by_bow(kf, 0.7)
            ts.append(((t1 - t0) * 1e6, (time.perf_counter() - t1) * 1e6))
        print(f"vocabulary: {n} nodes, levelsup {levelsup}; ComputeBoW p50 {np.median([t[0] for t in ts]):.1f} us ({len(bow['word_id'])} words, "
              f"{len(bow['node_id'])} nodes); SearchByBoW p50 {np.median([t[1] for t in ts]):.1f} us ({nm} matches, {ext.match_rescans()} rescans)")
