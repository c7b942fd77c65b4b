#!/usr/bin/env python3
"""Generate tests/golden/frame_*.npz: outputs of the REFERENCE's own src/Frame.cc (stereo
constructor: extraction on both eyes, UndistortKeyPoints, ComputeStereoMatches, ComputeImageBounds,
AssignFeaturesToGrid) compiled in place into oracle/_ref/libreframe.so, and of the cv2 4.13.0 wheel
for the step before the extractor (cvtColor, remap).

Run in the build container (needs /root/reference and cv2):

    python tests/golden/make_golden_frame.py

Writes
  frame_inputs.npz    the small stereo pairs, the colour frame, the rectification maps
  frame_expected.npz  per stereo case: mvuRight, mvDepth, mvKeysUn positions, mGrid (CSR), bounds,
                      number of keypoints, SHA-256 of the reference's keypoints + descriptors;
                      gray (cv2.cvtColor) and rectified (cv2.remap) images
The extraction threads inside the reference Frame use the monotone node allocator, i.e. the
canonical quadtree tie-break.
"""
import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
from oracle.pyoracle import RefFrame  # noqa: E402
from orb_slam2_test_b200 import synth  # noqa: E402
from test_oracle_frame import CAMS  # noqa: E402

HERE = Path(__file__).resolve().parent

# name -> (shape, seed, K, dist, bf, nfeatures); mb = bf / fx
CASES = {
    "pair_nodist": ((240, 416), 61, (320.0, 320.0, 208.0, 100.0), (0.0, 0.0, 0.0, 0.0), 170.0, 600),
    "pair_tum1": ((240, 320), 62, tuple(v / 2 for v in CAMS["TUM1"][0]), CAMS["TUM1"][1], 20.0, 500),
}


def sha(*arrays) -> str:
    h = hashlib.sha256()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def main():
    import cv2
    cv2.setNumThreads(1)
    inputs, expected = {}, {}
    for name, (shape, seed, K, dist, bf, nf) in CASES.items():
        left = synth.make_frame(shape, seed=seed, eye=0)
        right = synth.make_frame(shape, seed=seed, eye=1)
        r = RefFrame.stereo(left, right, K, dist, bf, bf / K[0], nfeatures=nf, canonical=True)
        inputs[f"{name}_left"], inputs[f"{name}_right"] = left, right
        for k in ("u_right", "depth", "un_xy", "cell_start", "cell_idx", "bounds"):
            expected[f"{name}_{k}"] = r[k]
        expected[f"{name}_n"] = np.array([len(r["kps_l"]), len(r["kps_r"])], np.int32)
        expected[f"{name}_sha"] = np.frombuffer(
            sha(r["kps_l"], r["desc_l"], r["kps_r"], r["desc_r"]).encode(), np.uint8)
        print(name, "keypoints", len(r["kps_l"]), len(r["kps_r"]), "matched", int((r["depth"] > 0).sum()),
              "in grid", int(r["cell_start"][-1]))
    colour = synth.make_color_frame((120, 160), seed=63, channels=3)
    inputs["colour_rgb"] = colour
    expected["gray_from_rgb"] = cv2.cvtColor(colour, cv2.COLOR_RGB2GRAY)
    expected["gray_from_bgr"] = cv2.cvtColor(colour, cv2.COLOR_BGR2GRAY)
    raw = synth.make_frame((120, 188), seed=64)
    mx, my = synth.rectify_maps(*synth.EUROC[0], size=(752, 480))
    mx = np.ascontiguousarray(mx[::4, ::4] / 4)  # a quarter-size version of the EuRoC map
    my = np.ascontiguousarray(my[::4, ::4] / 4)
    inputs["raw_gray"], inputs["map_x"], inputs["map_y"] = raw, mx, my
    expected["rectified"] = cv2.remap(raw, mx, my, cv2.INTER_LINEAR)
    np.savez_compressed(HERE / "frame_inputs.npz", **inputs)
    np.savez_compressed(HERE / "frame_expected.npz", **expected)
    print("written", (HERE / "frame_inputs.npz").stat().st_size, (HERE / "frame_expected.npz").stat().st_size)


if __name__ == "__main__":
    main()
