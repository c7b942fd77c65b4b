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
from make_golden_match import CASES, INIT_CASES, LOCAL_CASES, RELOC_CASES  # noqa: E402


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
