#!/usr/bin/env python3
"""Generate tests/golden/match_expected.npz: outputs of the REFERENCE's own src/ORBmatcher.cc
(ORBmatcher::SearchByProjection(Frame&, const Frame&, th, bMono) and SearchByProjection(Frame&, const
vector<MapPoint*>&, th), compiled in place into
oracle/_ref/librefmatch.so) on the scenarios of tests/matchcase.py, which are a pure function of
their arguments (synthetic frames -> oracle extractor -> map points), so only the outputs are stored.

Run in the build container (needs /root/reference):   python tests/golden/make_golden_match.py
"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import matchcase  # noqa: E402
from oracle.pyoracle import RefMatcher  # noqa: E402
from orb_slam2_test_b200 import synth  # noqa: E402

HERE = Path(__file__).resolve().parent

# name -> (make_case kwargs, mono flag, thresholds)
CASES = {
    "kitti_still": (dict(seed=11), False, (7.0, 15.0)),
    "kitti_forward": (dict(seed=12, dz=1.0), False, (7.0, 14.0)),
    "kitti_backward": (dict(seed=13, dz=-1.0), False, (7.0, 14.0)),
    "kitti_crowded": (dict(seed=14, copies=3, obs_rate=0.9), False, (15.0,)),
    "tum_mono": (dict(seed=15, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0, stereo=False,
                      nf=1000), True, (15.0, 30.0)),
}


# name -> (make_local_case kwargs, thresholds): ORBmatcher::SearchByProjection(F, vpMapPoints, th), nnratio 0.8
LOCAL_CASES = {
    "local_kitti": (dict(seed=16, copies=2, obs_rate=0.9), (1.0, 3.0)),
    "local_kitti_crowded": (dict(seed=17, copies=4, obs_rate=0.5), (1.0, 5.0)),
    "local_tum_mono": (dict(seed=18, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0, stereo=False,
                            nf=1000), (1.0, 5.0)),
}


# name -> (make_init_case kwargs, windows): ORBmatcher::SearchForInitialization, nnratio 0.9, orientation check on
INIT_CASES = {
    "init_tum": (dict(seed=19), (100, 30)),
    "init_tum_takeovers": (dict(seed=20, copies=3), (100,)),
    "init_kitti": (dict(seed=21, shape=synth.KITTI_SHAPE, K=matchcase.KITTI_K, dist=(0.0, 0.0, 0.0, 0.0)), (100,)),
}


# name -> (make_reloc_case kwargs, (th, ORBdist) settings): the relocalisation matcher
# ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), orientation check on and off
RELOC_CASES = {
    "reloc_kitti": (dict(seed=22), ((10.0, 100), (3.0, 64))),
    "reloc_kitti_ahead": (dict(seed=23, dz=2.0), ((10.0, 100),)),
    "reloc_kitti_crowded": (dict(seed=24, copies=3, found_rate=0.1), ((10.0, 100), (3.0, 64))),
    "reloc_tum_mono": (dict(seed=25, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0, stereo=False,
                            nf=1000, dist_jitter=0.4), ((10.0, 100),)),
}


# name -> make_fuse_case kwargs: ORBmatcher::Fuse(pKF, vpMapPoints, th), th 3 and 6
FUSE_CASES = {
    "fuse_kitti": dict(seed=26),
    "fuse_kitti_crowded": dict(seed=27, copies=3),
    "fuse_tum_mono": dict(seed=28, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0, stereo=False, nf=1000,
                          dist_jitter=0.4),
}
# name -> make_sim3_points_case kwargs: Fuse(pKF, Scw, ...) and SearchByProjection(pKF, Scw, ...), th 4 and 10
SIM3_POINT_CASES = {
    "sim3pts_kitti": dict(seed=29),
    "sim3pts_tum_mono": dict(seed=30, s=0.97, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0, stereo=False,
                             nf=1000, dist_jitter=0.4),
}
# name -> make_sim3_pair_case kwargs: SearchBySim3, th 7.5
SIM3_PAIR_CASES = {
    "sim3pair_kitti": dict(seed=31),
    "sim3pair_kitti_prematched": dict(seed=32, s12=0.96, dz=0.3, pre_rate=0.3),
}


def main():
    out = {}
    for name, kw in FUSE_CASES.items():
        kf, cam, bounds, scale, inv_sigma2, log_scale, mps = matchcase.make_fuse_case(**kw)
        for th in (3.0, 6.0):
            b, nf = RefMatcher.fuse(kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th)
            out[f"{name}_th{int(th)}_best"] = b
            out[f"{name}_th{int(th)}_nm"] = np.array([nf, len(mps["flags"])], np.int32)
            print(name, th, "nFused", nf)
    for name, kw in SIM3_POINT_CASES.items():
        kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occ = matchcase.make_sim3_points_case(**kw)
        for th in (4, 10):
            b, nf = RefMatcher.fuse_sim3(kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, float(th))
            m, nm = RefMatcher.search_by_projection_kf_sim3(kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occ, th)
            out[f"{name}_th{th}_best"] = b
            out[f"{name}_th{th}_match"] = m
            out[f"{name}_th{th}_nm"] = np.array([nf, nm, len(pts["flags"])], np.int32)
            print(name, th, "nFused", nf, "nmatches", nm)
    for name, kw in SIM3_PAIR_CASES.items():
        k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12 = matchcase.make_sim3_pair_case(**kw)
        m, nf = RefMatcher.search_by_sim3(k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12, 7.5)
        out[f"{name}_match"] = m
        out[f"{name}_nm"] = np.array([nf, len(p1["flags"])], np.int32)
        print(name, "nFound", nf)
    for name, (kw, settings) in RELOC_CASES.items():
        cur, cam, bounds, scale, log_scale, kf, occ = matchcase.make_reloc_case(**kw)
        for th, od in settings:
            for ori in (1, 0):
                m, nm = RefMatcher.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occ, th, od, bool(ori))
                key = f"{name}_th{int(th)}_d{od}_ori{ori}"
                out[key + "_match"] = m
                out[key + "_nm"] = np.array([nm, len(kf["flags"]), int((m >= 0).sum())], np.int32)
                print(key, "nmatches", nm, "slots", int((m >= 0).sum()))
    for name, (kw, wins) in INIT_CASES.items():
        f1, f2, bounds, prev = matchcase.make_init_case(**kw)
        for win in wins:
            m, pv, nm = RefMatcher.search_for_initialization(f1, f2, bounds, prev, win, 0.9, True)
            out[f"{name}_w{win}_match"] = m
            out[f"{name}_w{win}_prev"] = pv
            out[f"{name}_w{win}_nm"] = np.array([nm, len(f1["kps"])], np.int32)
            print(name, win, "nmatches", nm)
    for name, (kw, ths) in LOCAL_CASES.items():
        cur, bounds, scale, mps, occ = matchcase.make_local_case(**kw)
        for th in ths:
            m, nm = RefMatcher.search_local_points(cur, bounds, scale, mps, occ, th, 0.8)
            out[f"{name}_th{int(th)}_match"] = m
            out[f"{name}_th{int(th)}_nm"] = np.array([nm, len(mps["flags"]), int((m >= 0).sum())], np.int32)
            print(name, th, "nmatches", nm, "slots", int((m >= 0).sum()))
    for name, (kw, mono, ths) in CASES.items():
        cur, last, cam, bounds, scale = matchcase.make_case(**kw)
        for th in ths:
            for ori in (1, 0):
                m, nm = RefMatcher.search_by_projection(cur, last, cam, bounds, scale, th, mono=mono, check_ori=bool(ori))
                out[f"{name}_th{int(th)}_ori{ori}_match"] = m
                out[f"{name}_th{int(th)}_ori{ori}_nm"] = np.array([nm, len(last["kps"]), int((m >= 0).sum())], np.int32)
                print(name, th, ori, "nmatches", nm, "slots", int((m >= 0).sum()))
    np.savez_compressed(HERE / "match_expected.npz", **out)


if __name__ == "__main__":
    main()
