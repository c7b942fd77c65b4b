"""C++ call latency of ORBextractor::SearchByProjection through the shim (tests/shim/frame_like `match` mode with
FRAME_LIKE_TIME=1): the same scenario as tests/test_shim.py::test_frame_to_frame_matcher_through_the_shim."""
import os, struct, subprocess, sys, tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import matchcase  # noqa: E402
from orb_slam2_test_b200 import synth  # noqa: E402

tmp = Path(tempfile.mkdtemp())
bf = np.float32(386.1448); mb = bf / np.float32(718.856)
cur, last, cam, bounds, scale = matchcase.make_case(seed=33, bf=float(bf), mb=float(mb), dz=0.0)
shape = synth.KITTI_SHAPE
(tmp / "l.raw").write_bytes(synth.make_frame(shape, seed=33, frame=1, eye=0).tobytes())
(tmp / "r.raw").write_bytes(synth.make_frame(shape, seed=33, frame=1, eye=1).tobytes())
n = len(last["kps"])
T4 = lambda T: np.vstack([T, [0, 0, 0, 1]]).astype("<f4")
blob = struct.pack("<i4ffffi", n, *matchcase.KITTI_K, float(bf), float(mb), 7.0, 0)
blob += T4(cur["Tcw"]).tobytes() + T4(last["Tcw"]).tobytes()
blob += last["mp_xyz"].astype("<f4").tobytes() + last["mp_desc"].tobytes() + last["mp_flags"].tobytes()
blob += last["kps"]["octave"].astype("<i4").tobytes() + last["kps"]["angle"].astype("<f4").tobytes()
(tmp / "last.bin").write_bytes(blob)
r = subprocess.run([str(ROOT / "tests/shim/frame_like"), "match", str(tmp / "l.raw"), str(tmp / "r.raw"), str(shape[1]), str(shape[0]),
                    "2000", str(tmp / "last.bin"), str(tmp / "m.bin")], capture_output=True, text=True,
                   env={**os.environ, "FRAME_LIKE_TIME": "1"})
print(r.stdout, r.stderr[-500:])
