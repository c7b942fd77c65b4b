"""The restatement of ORBmatcher::SearchByProjection(Frame, Frame) (oracle/match_oracle.cpp) against the
reference's OWN src/ORBmatcher.cc + src/Frame.cc compiled in place (oracle/_ref/librefmatch.so, driven by
oracle/matcher_driver.cpp): CurrentFrame.mvpMapPoints and nmatches identical.  Needs /root/reference at
build time; tests/test_golden_match.py carries the same outputs to machines without it."""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

import matchcase


@pytest.fixture(scope="module")
def ref(oracle_built):
    if not oracle_built.RefMatcher.available():
        pytest.skip("oracle/_ref/librefmatch.so not built (no reference tree)")
    return oracle_built.RefMatcher


def _compare(py, ref, tag, ths, mono=False, **kw):
    cur, last, cam, bounds, scale = matchcase.make_case(**kw)
    total = 0
    for th in ths:
        for ori in (True, False):
            want, wnm = ref.search_by_projection(cur, last, cam, bounds, scale, th, mono=mono, check_ori=ori)
            got, nm = py.search_by_projection(cur, last, cam, bounds, scale, th, mono=mono, check_ori=ori)
            assert np.array_equal(got, want), f"{tag} th={th} ori={ori}: {(got != want).sum()} slots differ"
            assert nm == wnm, f"{tag} th={th} ori={ori}: nmatches {nm} vs {wnm}"
            total += wnm
    return total


@pytest.mark.parametrize("seed,dz", [(1, 0.0), (2, 1.0), (3, -1.0), (7, 0.3), (8, -0.6)])
def test_kitti_stereo(seed, dz, oracle_built, ref):
    """still / forward / backward (tlc.z against mb) and the two motions just either side of mb = 0.537 m"""
    assert _compare(oracle_built, ref, f"kitti {seed} dz {dz}", (7.0, 14.0, 30.0), seed=seed, dz=dz) > 1500


@pytest.mark.parametrize("obs_rate", [0.0, 0.5, 1.0])
def test_crowded_queries(obs_rate, oracle_built, ref):
    """three copies of every map point: the Observations() > 0 skip and the overwrite of points without"""
    _compare(oracle_built, ref, f"crowded {obs_rate}", (15.0, 40.0), seed=4, copies=3, obs_rate=obs_rate)


def test_monocular_distorted(oracle_built, ref):
    assert _compare(oracle_built, ref, "tum mono", (15.0, 30.0), mono=True, seed=5, shape=synth.TUM_SHAPE,
                    K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0, stereo=False, nf=1000) > 500


def test_points_behind_and_outside(oracle_built, ref):
    """a current pose looking the other way: invzc < 0 and out-of-bounds projections reject everything"""
    cur, last, cam, bounds, scale = matchcase.make_case(seed=9)
    cur["Tcw"] = matchcase.compose(matchcase.pose(ry=np.pi), cur["Tcw"])
    want, wnm = ref.search_by_projection(cur, last, cam, bounds, scale, 15.0)
    got, nm = oracle_built.search_by_projection(cur, last, cam, bounds, scale, 15.0)
    assert nm == wnm == 0 and np.array_equal(got, want) and (got == -1).all()


# ---- the local-map matcher: ORBmatcher::SearchByProjection(F, vpMapPoints, th), src/ORBmatcher.cc:59-154 ----

@pytest.mark.parametrize("kw", [dict(seed=1), dict(seed=2, copies=3, obs_rate=0.5), dict(seed=3, copies=4, obs_rate=1.0),
                                dict(seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0,
                                     stereo=False, nf=1000)], ids=["kitti", "kitti-half-observed", "kitti-crowded", "tum-mono"])
def test_local_map_matcher(kw, oracle_built, ref):
    """RadiusByViewingCos either side of 0.998, th = 1 (no factor) / 3 / 5, two ratio thresholds, slots occupied by the
    motion-model matcher with and without observations"""
    cur, bounds, scale, mps, occ = matchcase.make_local_case(**kw)
    assert (occ == 2).sum() > 100 and (occ == 1).sum() > 10
    for th in (1.0, 3.0, 5.0):
        for ratio in (0.8, 0.6):
            want, wnm = ref.search_local_points(cur, bounds, scale, mps, occ, th, ratio)
            got, nm = oracle_built.search_local_points(cur, bounds, scale, mps, occ, th, ratio)
            assert np.array_equal(got, want) and nm == wnm, f"th={th} ratio={ratio}"
            assert not (got[occ == 2] >= 0).any()


# ---- the monocular-initialisation matcher: ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:487-632 ----

@pytest.mark.parametrize("kw", [dict(seed=1), dict(seed=2, step=3), dict(seed=4, nf=4000, step=2), dict(seed=9, copies=2),
                                dict(seed=9, copies=4), dict(seed=9, copies=48),
                                dict(seed=3, shape=synth.KITTI_SHAPE, K=matchcase.KITTI_K, dist=(0.0, 0.0, 0.0, 0.0))],
                         ids=["tum", "tum-3-later", "tum-4000", "takeovers-x2", "takeovers-x4", "takeovers-x48", "kitti"])
def test_initialisation_matcher(kw, oracle_built, ref):
    """windows 100 (the reference's), 30 and 300; ratio 0.9 / 0.7; with and without the orientation check; then the
    call chained on a later frame with the vbPrevMatched the first call updated (src/Tracking.cc:616)"""
    f1, f2, bounds, prev = matchcase.make_init_case(**kw)
    for win in (100, 30, 300):
        for ratio, ori in ((0.9, True), (0.7, False)):
            want, wprev, wnm = ref.search_for_initialization(f1, f2, bounds, prev, win, ratio, ori)
            got, gprev, nm = oracle_built.search_for_initialization(f1, f2, bounds, prev, win, ratio, ori)
            assert np.array_equal(got, want) and nm == wnm and np.array_equal(gprev.view(np.uint32), wprev.view(np.uint32))
    _, pv, _ = ref.search_for_initialization(f1, f2, bounds, prev, 100, 0.9, True)
    f3 = matchcase.make_init_case(**{**kw, "copies": 1, "step": kw.get("step", 1) + 1})[1]
    a = ref.search_for_initialization(f1, f3, bounds, pv, 100, 0.9, True)
    b = oracle_built.search_for_initialization(f1, f3, bounds, pv, 100, 0.9, True)
    assert np.array_equal(a[0], b[0]) and a[2] == b[2] and np.array_equal(a[1], b[1])


# ---- the relocalisation matcher: ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist),
# ---- src/ORBmatcher.cc:1670-1798 ----

def _compare_reloc(py, ref, tag, settings=((10.0, 100), (3.0, 64)), **kw):
    cur, cam, bounds, scale, log_scale, kf, occupied = matchcase.make_reloc_case(**kw)
    total = 0
    for th, orb_dist in settings:
        for ori in (True, False):
            for occ in (occupied, None):
                want, wnm = ref.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occ, th, orb_dist, ori)
                got, nm = py.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occ, th, orb_dist, ori)
                assert np.array_equal(got, want), f"{tag} th={th} dist={orb_dist} ori={ori}: {(got != want).sum()} slots differ"
                assert nm == wnm, f"{tag} th={th} dist={orb_dist} ori={ori}: nmatches {nm} vs {wnm}"
                if occ is not None:
                    assert not (want[occ != 0] >= 0).any()
                total += wnm
    return total


@pytest.mark.parametrize("seed,dz", [(1, 0.0), (2, 2.0), (3, -2.0)])
def test_reloc_kitti(seed, dz, oracle_built, ref):
    """the keyframe seen again from the same place, from 2 m ahead and 2 m behind (the predicted level moves)"""
    assert _compare_reloc(oracle_built, ref, f"reloc kitti {seed} dz {dz}", seed=seed, dz=dz) > 1000


def test_reloc_crowded(oracle_built, ref):
    """three copies of every keyframe point: a slot taken by an earlier query is closed to every later one (:1735-1736)"""
    _compare_reloc(oracle_built, ref, "reloc crowded", seed=4, copies=3, found_rate=0.1)


def test_reloc_monocular_distorted(oracle_built, ref):
    assert _compare_reloc(oracle_built, ref, "reloc tum", seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D,
                          bf=40.0, stereo=False, nf=1000, dist_jitter=0.4) > 500


def test_reloc_pose_looking_away(oracle_built, ref):
    """no invzc < 0 test in this matcher: points behind the camera that project inside the image are searched (:1701-1710)"""
    cur, cam, bounds, scale, log_scale, kf, occupied = matchcase.make_reloc_case(seed=9)
    cur["Tcw"] = matchcase.compose(matchcase.pose(ry=np.pi), cur["Tcw"])
    want, wnm = ref.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occupied, 10.0, 100, True)
    got, nm = oracle_built.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occupied, 10.0, 100, True)
    assert nm == wnm and np.array_equal(got, want)


# ---- ORBmatcher::Fuse(pKF, vpMapPoints, th), src/ORBmatcher.cc:968-1131 (LocalMapping::SearchInNeighbors) ----

def _compare_fuse(py, ref, tag, ths=(3.0, 6.0), **kw):
    kf, cam, bounds, scale, inv_sigma2, log_scale, mps = matchcase.make_fuse_case(**kw)
    total = 0
    for th in ths:
        want, wnf = ref.fuse(kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th)
        got, nf = py.fuse(kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th)
        assert np.array_equal(got, want), f"{tag} th={th}: {(got != want).sum()} map points differ"
        assert nf == wnf == int((got >= 0).sum())
        assert not (got[(mps["flags"] & 6) != 0] >= 0).any()
        total += nf
    return total


@pytest.mark.parametrize("seed,dz", [(1, 0.0), (2, 0.3), (3, -0.3)])
def test_fuse_kitti_stereo_keyframe(seed, dz, oracle_built, ref):
    """stereo keypoints take the 3-dof chi-square test (7.8), the others the 2-dof one (5.99)"""
    assert _compare_fuse(oracle_built, ref, f"fuse kitti {seed} dz {dz}", seed=seed, dz=dz) > 400


def test_fuse_crowded(oracle_built, ref):
    _compare_fuse(oracle_built, ref, "fuse crowded", seed=4, copies=3)


def test_fuse_monocular_distorted_keyframe(oracle_built, ref):
    """a distorted camera: the keyframe's bounds are the Frame's float bounds truncated to int"""
    assert _compare_fuse(oracle_built, ref, "fuse tum", seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D,
                         bf=40.0, stereo=False, nf=1000, dist_jitter=0.4) > 150


# ---- LoopClosing's matchers: Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) :1133-1260, SearchByProjection(pKF, Scw, vpPoints,
# ---- vpMatched, th) :353-485, SearchBySim3 :1262-1501 ----

@pytest.mark.parametrize("kw", [dict(seed=1), dict(seed=2, s=0.97, dz=0.3), dict(seed=4, copies=3),
                                dict(seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0,
                                     stereo=False, nf=1000, dist_jitter=0.4)])
def test_fuse_sim3_and_projection_sim3(kw, oracle_built, ref):
    kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occupied = matchcase.make_sim3_points_case(**kw)
    total = 0
    for th in (4.0, 10.0):
        want, wnf = ref.fuse_sim3(kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, th)
        got, nf = oracle_built.fuse_sim3(kf, cam, bounds, scale, log_scale, Scw, pts, th)
        assert np.array_equal(got, want) and nf == wnf == int((got >= 0).sum()), f"fuse sim3 th={th}: {(got != want).sum()} differ"
        total += nf
        for occ in (occupied, None):
            p = pts if occ is not None else dict(pts, flags=(pts["flags"] & 3).astype(np.uint8))
            want, wnm = ref.search_by_projection_kf_sim3(kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, p, occ, int(th))
            got, nm = oracle_built.search_by_projection_kf_sim3(kf, cam, bounds, scale, log_scale, Scw, p, occ, int(th))
            assert np.array_equal(got, want) and nm == wnm == int((got >= 0).sum()), f"projection sim3 th={th}: {(got != want).sum()} differ"
            if occ is not None:
                assert not (got[occ != 0] >= 0).any()
            total += nm
    assert total > 300


@pytest.mark.parametrize("kw", [dict(seed=1), dict(seed=2, s12=0.96, dz=0.3), dict(seed=3, dz=-0.3, pre_rate=0.3),
                                dict(seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0,
                                     stereo=False, nf=1000, dist_jitter=0.4)])
def test_search_by_sim3(kw, oracle_built, ref):
    k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12 = matchcase.make_sim3_pair_case(**kw)
    total = 0
    for th in (7.5, 3.0):
        want, wnf = ref.search_by_sim3(k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12, th)
        got, nf = oracle_built.search_by_sim3(k1, k2, cam, bounds, scale, log_scale, p1, p2, pre12, s12, R12, t12, th)
        assert np.array_equal(got, want) and nf == wnf, f"sim3 th={th}: {(got != want).sum()} differ, nFound {nf} vs {wnf}"
        assert np.array_equal(got[pre12 >= 0], pre12[pre12 >= 0])
        total += nf
    assert total > 100
