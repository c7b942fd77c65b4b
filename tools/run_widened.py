"""One pass over every kernel of the widened rows for ncu launch lists:
K0 colour->gray and rectification, the extraction chain, K8 undistort+grid, S0/S1/S2 stereo match,
M1a/M1b frame-to-frame matcher.
usage: python tools/run_widened.py [pairs]   (default 1 stereo pair per call; e.g. 64 for batches)"""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
from orb_slam2_test_b200 import ORBextractor, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
BF, FX = 386.1448, 718.856
# EuRoC-style stereo: raw gray frames rectified on the device, then matched on the device
mxl, myl = synth.rectify_maps(*synth.EUROC[0], size=(752, 480))
mxr, myr = synth.rectify_maps(*synth.EUROC[1], size=(752, 480))
L = ORBextractor(1200, 1.2, 8, 20, 7, max_batch=n, want_pyramid=False)
R = ORBextractor(1200, 1.2, 8, 20, 7, max_batch=n, want_pyramid=False)
L.set_rectify_map(mxl, myl)
R.set_rectify_map(mxr, myr)
L.set_camera((435.2046959714599, 435.2046959714599, 367.4517211914062, 252.2008514404297), (0, 0, 0, 0))
li = synth.make_sequence(synth.EUROC_SHAPE, seed=40, start=0, count=n, eye=0)
ri = synth.make_sequence(synth.EUROC_SHAPE, seed=40, start=0, count=n, eye=1)
for _ in range(3):
    L.extract_batch(li)
    R.extract_batch(ri)
    L.stereo_match(R, 47.90639384423901, 47.90639384423901 / 435.2046959714599)
print("euroc pairs:", n, "matched", [L.stereo_result(f)[2] for f in range(min(n, 4))])
# TUM colour frames with the TUM1 camera
C = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=n, want_pyramid=False)
C.set_input_format("rgb")
C.set_camera((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314))
ci = np.stack([synth.make_color_frame(synth.TUM_SHAPE, seed=21, frame=f) for f in range(n)])
for _ in range(3):
    C.extract_batch(ci)
print("tum colour frames:", n, "in grid", int(C.frame_grid(0)[1][-1]), "of", C.count(0))
# KITTI stereo pair (no rectification): the K2k match
KL = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=n, want_pyramid=False)
KR = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=n, want_pyramid=False)
kl = synth.make_sequence(synth.KITTI_SHAPE, seed=31, start=0, count=n, eye=0)
kr = synth.make_sequence(synth.KITTI_SHAPE, seed=31, start=0, count=n, eye=1)
for _ in range(3):
    KL.extract_batch(kl)
    KR.extract_batch(kr)
    KL.stereo_match(KR, BF, BF / FX)
print("kitti pairs:", n, "matched", [KL.stereo_result(f)[2] for f in range(min(n, 4))])
# the frame-to-frame matcher (M1a / M1b) on a KITTI stereo frame pair: last frame's stereo depths as map points
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / "tests"))
import matchcase  # noqa: E402

K = matchcase.KITTI_K
ML = ORBextractor(2000, 1.2, 8, 20, 7)
MR = ORBextractor(2000, 1.2, 8, 20, 7)
ML.set_camera(K, (0.0, 0.0, 0.0, 0.0))
frames = {}
for f in (0, 1):
    ML(synth.make_frame(synth.KITTI_SHAPE, seed=71, frame=f, eye=0))
    MR(synth.make_frame(synth.KITTI_SHAPE, seed=71, frame=f, eye=1))
    ML.stereo_match(MR, BF, BF / FX)
    kps, desc = ML.result(0)
    u, z, _ = ML.stereo_result(0)
    frames[f] = {"kps": kps, "desc": desc, "un_xy": ML.frame_grid(0)[0], "depth": z}
T_last = matchcase.pose(0.01, -0.02, 0.005, (0.3, -0.1, 0.2))
T_cur = matchcase.compose(matchcase.pose(rx=1.0 / K[1], ry=-3.0 / K[0]), T_last)
last = matchcase.map_points(frames[0], T_last, K, np.random.default_rng(71))
last["octave"], last["angle"] = last["kps"]["octave"], last["kps"]["angle"]
for th in (7.0, 15.0, 7.0):
    m, nm = ML.search_by_projection(last, T_cur, BF, BF / FX, th, use_stereo=True)
print("matcher: nmatches", nm, "rounds", ML.match_rounds(), "rescans", ML.match_rescans())
# the local-map matcher (MODE 2) and the monocular-initialisation matcher (MODE 3 + M3) on the same kind of frames
occ = np.where(m >= 0, 2, 0).astype(np.uint8)
rng = np.random.default_rng(72)
Xc = last["mp_xyz"].astype(np.float64) @ T_cur[:, :3].T.astype(np.float64) + T_cur[:, 3]
z = np.where(np.abs(Xc[:, 2]) > 1e-6, Xc[:, 2], 1.0)
proj = np.stack([K[0] * Xc[:, 0] / z + K[2], K[1] * Xc[:, 1] / z + K[3], K[0] * Xc[:, 0] / z + K[2] - BF / z,
                 rng.uniform(0.99, 1.0, len(z))], 1).astype(np.float32)
mps = {"proj": proj, "desc": last["mp_desc"], "flags": ((last["mp_flags"] & 1) | 2).astype(np.uint8),
       "level": last["kps"]["octave"].astype(np.int32)}
for th in (1.0, 3.0, 1.0):
    lm, lnm = ML.search_local_points(mps, occ, th, 0.8, use_stereo=True)
print("local-map matcher: nmatches", lnm, "rounds", ML.match_rounds())
INI = ORBextractor(2000, 1.2, 8, 20, 7)
INI.set_camera(matchcase.TUM1_K, matchcase.TUM1_D)
k0, d0 = INI(synth.make_frame(synth.TUM_SHAPE, seed=73, frame=0))
f1 = {"kps": k0, "desc": d0}
prev = INI.frame_grid(0)[0].copy()
INI(synth.make_frame(synth.TUM_SHAPE, seed=73, frame=1))
for _ in range(3):
    im, ipv, inm = INI.search_for_initialization(f1, prev, 100, 0.9)
print("initialisation matcher: nmatches", inm, "rounds", INI.match_rounds())
