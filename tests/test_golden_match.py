"""Golden vectors of the matcher row: outputs of the reference's own ORBmatcher::SearchByProjection
(Frame, Frame) (tests/golden/make_golden_match.py, oracle/_ref/librefmatch.so) on the scenarios of
tests/matchcase.py.  The CPU restatement (not gpu) and the CUDA path (gpu) must both reproduce them."""
import sys
from pathlib import Path

import numpy as np
import pytest

import matchcase

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE / "golden"))
from make_golden_match import (CASES, FUSE_CASES, INIT_CASES, LOCAL_CASES, RELOC_CASES, SIM3_PAIR_CASES,  # noqa: E402
                               SIM3_POINT_CASES)


def _golden():
    return np.load(HERE / "golden" / "match_expected.npz")


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_reproduces_the_reference_matcher(name, oracle_built):
    g = _golden()
    kw, mono, ths = CASES[name]
    cur, last, cam, bounds, scale = matchcase.make_case(**kw)
    for th in ths:
        for ori in (1, 0):
            m, nm = oracle_built.search_by_projection(cur, last, cam, bounds, scale, th, mono=mono, check_ori=bool(ori))
            key = f"{name}_th{int(th)}_ori{ori}"
            assert len(last["kps"]) == g[key + "_nm"][1], "the scenario is not the one the golden file was made from"
            assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_reproduces_the_reference_matcher(name, oracle_built):
    from orb_slam2_test_b200 import ORBextractor, synth
    g = _golden()
    kw, mono, ths = CASES[name]
    cur, last, cam, bounds, scale = matchcase.make_case(**kw)
    stereo = kw.get("stereo", True)
    shape, seed, nf = kw.get("shape", synth.KITTI_SHAPE), kw["seed"], kw.get("nf", 2000)
    ext = ORBextractor(nf, 1.2, 8, 20, 7)
    ext.set_camera(kw.get("K", matchcase.KITTI_K), kw.get("dist", (0.0, 0.0, 0.0, 0.0)))
    ext(synth.make_frame(shape, seed=seed, frame=1, eye=0))
    if stereo:
        right = ORBextractor(nf, 1.2, 8, 20, 7)
        right(synth.make_frame(shape, seed=seed, frame=1, eye=1))
        ext.stereo_match(right, float(cam[4]), float(cam[5]))
    lf = dict(last)
    lf["octave"], lf["angle"] = last["kps"]["octave"], last["kps"]["angle"]
    for th in ths:
        for ori in (1, 0):
            m, nm = ext.search_by_projection(lf, cur["Tcw"], float(cam[4]), float(cam[5]), th, mono=mono,
                                             check_orientation=bool(ori), use_stereo=stereo)
            key = f"{name}_th{int(th)}_ori{ori}"
            assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key


@pytest.mark.parametrize("name", sorted(LOCAL_CASES))
def test_oracle_reproduces_the_reference_local_matcher(name, oracle_built):
    g = _golden()
    kw, ths = LOCAL_CASES[name]
    cur, bounds, scale, mps, occ = matchcase.make_local_case(**kw)
    for th in ths:
        m, nm = oracle_built.search_local_points(cur, bounds, scale, mps, occ, th, 0.8)
        key = f"{name}_th{int(th)}"
        assert len(mps["flags"]) == g[key + "_nm"][1], "the scenario is not the one the golden file was made from"
        assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(LOCAL_CASES))
def test_cuda_reproduces_the_reference_local_matcher(name, oracle_built):
    from orb_slam2_test_b200 import ORBextractor, synth
    g = _golden()
    kw, ths = LOCAL_CASES[name]
    cur, bounds, scale, mps, occ = matchcase.make_local_case(**kw)
    stereo = kw.get("stereo", True)
    shape, seed, nf = kw.get("shape", synth.KITTI_SHAPE), kw["seed"], kw.get("nf", 2000)
    K, bf = kw.get("K", matchcase.KITTI_K), kw.get("bf", matchcase.KITTI_BF)
    ext = ORBextractor(nf, 1.2, 8, 20, 7)
    ext.set_camera(K, kw.get("dist", (0.0, 0.0, 0.0, 0.0)))
    ext(synth.make_frame(shape, seed=seed, frame=1, eye=0))
    if stereo:
        right = ORBextractor(nf, 1.2, 8, 20, 7)
        right(synth.make_frame(shape, seed=seed, frame=1, eye=1))
        ext.stereo_match(right, bf, bf / K[0])
    for th in ths:
        m, nm = ext.search_local_points(mps, occ, th, 0.8, use_stereo=stereo)
        key = f"{name}_th{int(th)}"
        assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key


@pytest.mark.parametrize("name", sorted(INIT_CASES))
def test_oracle_reproduces_the_reference_initialisation_matcher(name, oracle_built):
    g = _golden()
    kw, wins = INIT_CASES[name]
    f1, f2, bounds, prev = matchcase.make_init_case(**kw)
    for win in wins:
        m, pv, nm = oracle_built.search_for_initialization(f1, f2, bounds, prev, win, 0.9, True)
        key = f"{name}_w{win}"
        assert len(f1["kps"]) == g[key + "_nm"][1], "the scenario is not the one the golden file was made from"
        assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key
        assert np.array_equal(pv.view(np.uint32), g[key + "_prev"].view(np.uint32)), key


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(INIT_CASES))
def test_cuda_reproduces_the_reference_initialisation_matcher(name, oracle_built):
    from orb_slam2_test_b200 import ORBextractor, synth
    g = _golden()
    kw, wins = INIT_CASES[name]
    f1, f2, bounds, prev = matchcase.make_init_case(**kw)
    ext = ORBextractor(kw.get("nf", 2000), 1.2, 8, 20, 7)
    ext.set_camera(kw.get("K", matchcase.TUM1_K), kw.get("dist", matchcase.TUM1_D))
    ext(synth.make_frame(kw.get("shape", synth.TUM_SHAPE), seed=kw["seed"], frame=kw.get("step", 1)))
    for win in wins:
        m, pv, nm = ext.search_for_initialization(f1, prev, win, 0.9)
        key = f"{name}_w{win}"
        assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key
        assert np.array_equal(pv.view(np.uint32), g[key + "_prev"].view(np.uint32)), key


@pytest.mark.parametrize("name", sorted(RELOC_CASES))
def test_oracle_reproduces_the_reference_relocalisation_matcher(name, oracle_built):
    g = _golden()
    kw, settings = RELOC_CASES[name]
    cur, cam, bounds, scale, log_scale, kf, occ = matchcase.make_reloc_case(**kw)
    for th, od in settings:
        for ori in (1, 0):
            m, nm = oracle_built.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occ, th, od, bool(ori))
            key = f"{name}_th{int(th)}_d{od}_ori{ori}"
            assert len(kf["flags"]) == g[key + "_nm"][1], "the scenario is not the one the golden file was made from"
            assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(RELOC_CASES))
def test_cuda_reproduces_the_reference_relocalisation_matcher(name, oracle_built):
    from orb_slam2_test_b200 import ORBextractor, synth
    g = _golden()
    kw, settings = RELOC_CASES[name]
    cur, cam, bounds, scale, log_scale, kf, occ = matchcase.make_reloc_case(**kw)
    shape, seed, nf = kw.get("shape", synth.KITTI_SHAPE), kw["seed"], kw.get("nf", 2000)
    ext = ORBextractor(nf, 1.2, 8, 20, 7)
    ext.set_camera(kw.get("K", matchcase.KITTI_K), kw.get("dist", (0.0, 0.0, 0.0, 0.0)))
    ext(synth.make_frame(shape, seed=seed, frame=1, eye=0))
    for th, od in settings:
        for ori in (1, 0):
            m, nm = ext.search_by_projection_keyframe(kf, cur["Tcw"], occ, th, od, check_orientation=bool(ori))
            key = f"{name}_th{int(th)}_d{od}_ori{ori}"
            assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][0], key


def _loaded_keyframe(kf, kw):
    from orb_slam2_test_b200 import ORBextractor, synth
    shape = kw.get("shape", synth.KITTI_SHAPE)
    h = ORBextractor(kw.get("nf", 2000), 1.2, 8, 20, 7)
    h.set_camera(kw.get("K", matchcase.KITTI_K), kw.get("dist", (0.0, 0.0, 0.0, 0.0)))
    h.load_frame(kf["kps"], kf["desc"], shape[1], shape[0], kf["u_right"] if kw.get("stereo", True) else None)
    return h


@pytest.mark.parametrize("name", sorted(FUSE_CASES))
@pytest.mark.parametrize("cuda", [False, pytest.param(True, marks=pytest.mark.gpu)])
def test_the_reference_fuse_search_is_reproduced(name, cuda, oracle_built):
    g = _golden()
    kw = FUSE_CASES[name]
    kf, cam, bounds, scale, inv_sigma2, log_scale, mps = matchcase.make_fuse_case(**kw)
    h = _loaded_keyframe(kf, kw) if cuda else None
    for th in (3.0, 6.0):
        key = f"{name}_th{int(th)}"
        assert len(mps["flags"]) == g[key + "_nm"][1], "the scenario is not the one the golden file was made from"
        if cuda:
            b, nf = h.fuse_search(mps, kf["Tcw"], oracle_built.camera_centre(kf["Tcw"]), float(cam[4]), th,
                                  use_stereo=kw.get("stereo", True))
        else:
            b, nf = oracle_built.fuse(kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th)
        assert np.array_equal(b, g[key + "_best"]) and nf == g[key + "_nm"][0], key


@pytest.mark.parametrize("name", sorted(SIM3_POINT_CASES))
@pytest.mark.parametrize("cuda", [False, pytest.param(True, marks=pytest.mark.gpu)])
def test_the_reference_sim3_point_matchers_are_reproduced(name, cuda, oracle_built):
    g = _golden()
    kw = SIM3_POINT_CASES[name]
    kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occ = matchcase.make_sim3_points_case(**kw)
    h = _loaded_keyframe(dict(kf, u_right=None), dict(kw, stereo=False)) if cuda else None
    T, Ow = oracle_built.sim3_decompose(Scw)
    for th in (4, 10):
        key = f"{name}_th{th}"
        assert len(pts["flags"]) == g[key + "_nm"][2], "the scenario is not the one the golden file was made from"
        if cuda:
            b, nf = h.project_search(pts, T, None, Ow, float(th), 50)
            m, nm = h.search_by_projection_sim3(pts, T, Ow, occ, th)
        else:
            b, nf = oracle_built.fuse_sim3(kf, cam, bounds, scale, log_scale, Scw, pts, float(th))
            m, nm = oracle_built.search_by_projection_kf_sim3(kf, cam, bounds, scale, log_scale, Scw, pts, occ, th)
        assert np.array_equal(b, g[key + "_best"]) and nf == g[key + "_nm"][0], key
        assert np.array_equal(m, g[key + "_match"]) and nm == g[key + "_nm"][1], key


@pytest.mark.parametrize("name", sorted(SIM3_PAIR_CASES))
@pytest.mark.parametrize("cuda", [False, pytest.param(True, marks=pytest.mark.gpu)])
def test_the_reference_search_by_sim3_is_reproduced(name, cuda, oracle_built):
    g = _golden()
    kw = SIM3_PAIR_CASES[name]
    k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12 = matchcase.make_sim3_pair_case(**kw)
    assert len(p1["flags"]) == g[name + "_nm"][1], "the scenario is not the one the golden file was made from"
    if cuda:
        h1 = _loaded_keyframe(dict(k1, u_right=None), dict(kw, stereo=False))
        h2 = _loaded_keyframe(dict(k2, u_right=None), dict(kw, stereo=False))
        T12, T21 = oracle_built.sim3_relative(s12, R12, t12)
        m, nf = h1.search_by_sim3(h2, p1, p2, k1["Tcw"], k2["Tcw"], T12, T21, 7.5, pre12)
    else:
        m, nf = oracle_built.search_by_sim3(k1, k2, cam, bounds, scale, log_scale, p1, p2, pre12, s12, R12, t12, 7.5)
    assert np.array_equal(m, g[name + "_match"]) and nf == g[name + "_nm"][0], name
