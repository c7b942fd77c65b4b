#!/usr/bin/env python3
"""Generate tests/golden/*: outputs of the REFERENCE's own ORBextractor.cc on fixed frames.

Run in the build container (needs /root/reference; `make -C oracle` compiles the reference's
src/ORBextractor.cc in place into oracle/_ref/libreforb_mono.so -- monotone allocator, so that its
heap-address tie-break equals the canonical rule):

    python tests/golden/make_golden.py

Writes
  frames.npz    the input frames themselves (small: 320x240 and 416x128, so the bytes do not
                depend on any generator being reproducible on another machine)
  expected.npz  per case: keypoints (cv::KeyPoint records) and descriptors from the reference
  golden.json   per case: constructor parameters, SHA-256 of input / keypoints / descriptors, and
                the same hashes for the full-size BASELINE configs on synth frames (inputs hashed
                too: if a machine's numpy renders different synth bytes those cases are skipped)
"""
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle.pyoracle import Reference  # noqa: E402
from orb_slam2_test_b200 import synth  # noqa: E402

HERE = Path(__file__).resolve().parent


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def small_frames():
    out = {}
    out["qvga_a"] = synth.make_frame((240, 320), seed=21, frame=0)
    out["qvga_b"] = synth.make_frame((240, 320), seed=21, frame=5)
    out["wide"] = synth.make_frame((128, 416), seed=22, frame=0)  # nIni = 4 at level 0
    rng = np.random.default_rng(5)
    out["noise"] = rng.integers(0, 256, (200, 260), dtype=np.uint8)  # dense candidates, many ties
    chk = ((np.add.outer(np.arange(240) // 9, np.arange(320) // 11) % 2) * 120 + 60).astype(np.uint8)
    out["checker"] = chk  # exact response ties everywhere
    return out


SMALL_CASES = [  # (case name, frame, nfeatures, scale, levels, iniTh, minTh)
    ("qvga_a_500", "qvga_a", 500, 1.2, 8, 20, 7),
    ("qvga_b_1000", "qvga_b", 1000, 1.2, 8, 20, 7),
    ("qvga_a_kitti04", "qvga_a", 2000, 1.2, 8, 12, 7),   # KITTI04-12.yaml thresholds
    ("wide_300", "wide", 300, 1.2, 4, 20, 7),
    ("noise_400", "noise", 400, 1.2, 5, 20, 7),
    ("checker_200", "checker", 200, 1.2, 6, 20, 7),
    ("qvga_a_s15", "qvga_a", 600, 1.5, 4, 20, 7),
]

FULL_CASES = [  # BASELINE.json configs on synth frames
    ("T1k", synth.TUM_SHAPE, 1000, 3, 0), ("T1k", synth.TUM_SHAPE, 1000, 3, 1),
    ("T2k", synth.TUM_SHAPE, 2000, 3, 0),
    ("K2k", synth.KITTI_SHAPE, 2000, 3, 0), ("K2k", synth.KITTI_SHAPE, 2000, 3, 1),
    ("K2k", synth.KITTI_SHAPE, 2000, 7, 0),
]


def main():
    assert Reference.available(mono=True), "build oracle/_ref first (make -C oracle)"
    frames = small_frames()
    np.savez_compressed(HERE / "frames.npz", **frames)
    expected, meta = {}, {"small": {}, "full": []}
    for name, fr, nf, sf, nl, it, mt in SMALL_CASES:
        ref = Reference(nf, sf, nl, it, mt, mono=True)
        k, d = ref.extract(frames[fr])
        expected[name + "_kp"] = k.view(np.uint8).reshape(-1, 28)
        expected[name + "_desc"] = d
        meta["small"][name] = {"frame": fr, "nfeatures": nf, "scaleFactor": sf, "nlevels": nl, "iniThFAST": it,
                               "minThFAST": mt, "n": int(len(k)), "input_sha256": sha(frames[fr]),
                               "kp_sha256": sha(k.view(np.uint8)), "desc_sha256": sha(d)}
        print(name, len(k))
    np.savez_compressed(HERE / "expected.npz", **expected)
    for cfg, shape, nf, seed, frame in FULL_CASES:
        img = synth.make_frame(shape, seed=seed, frame=frame)
        k, d = Reference(nf, 1.2, 8, 20, 7, mono=True).extract(img)
        meta["full"].append({"config": cfg, "rows": shape[0], "cols": shape[1], "nfeatures": nf, "seed": seed,
                             "frame": frame, "n": int(len(k)), "input_sha256": sha(img),
                             "kp_sha256": sha(k.view(np.uint8)), "desc_sha256": sha(d)})
        print(cfg, seed, frame, len(k))
    meta["generator"] = "reference src/ORBextractor.cc compiled in place (oracle/_ref/libreforb_mono.so)"
    (HERE / "golden.json").write_text(json.dumps(meta, indent=1) + "\n")


if __name__ == "__main__":
    main()
