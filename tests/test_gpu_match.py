"""GPU parity for the frame-to-frame matcher (first consumer):
borb_search_by_projection against the CPU restatement of ORBmatcher::SearchByProjection(Frame, Frame)
(oracle/match_oracle.cpp, src/ORBmatcher.cc:1503-1668), which tests/test_oracle_match_vs_reference.py
pins to the reference's own ORBmatcher.cc compiled in place.

Bar: CurrentFrame.mvpMapPoints (as indices into the last frame) and nmatches identical.  The current
frame is extracted on the GPU (bit-exact with the oracle frame the scenario was built from, so the
indices mean the same keypoints); its descriptors, undistorted positions, grid, bounds and mvuRight never
leave the device."""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

import matchcase

pytestmark = pytest.mark.gpu


def _gpu_frame(shape, seed, frame, K, dist, stereo, nf, bf, mb):
    from orb_slam2_test_b200 import ORBextractor
    left = ORBextractor(nf, 1.2, 8, 20, 7)
    left.set_camera(K, dist)
    left(synth.make_frame(shape, seed=seed, frame=frame, eye=0))
    right = None
    if stereo:
        right = ORBextractor(nf, 1.2, 8, 20, 7)
        right(synth.make_frame(shape, seed=seed, frame=frame, eye=1))
        left.stereo_match(right, bf, mb)
    return left, right


def _last_for_abi(last):
    d = dict(last)
    d["octave"] = last["kps"]["octave"]
    d["angle"] = last["kps"]["angle"]
    return d


def _check(tag, oracle, ths=(7.0, 15.0, 30.0), mono_flag=False, **kw):
    cur, last, cam, bounds, scale = matchcase.make_case(**kw)
    stereo = kw.get("stereo", True)
    K = kw.get("K", matchcase.KITTI_K)
    dist = kw.get("dist", (0.0, 0.0, 0.0, 0.0))
    bf = float(cam[4]); mb = float(cam[5])
    ext, right = _gpu_frame(kw.get("shape", synth.KITTI_SHAPE), kw.get("seed", 0), kw.get("frame", 0) + 1, K, dist,
                            stereo, kw.get("nf", 2000), bf, mb)
    kps, desc = ext.result(0)
    assert kps.tobytes() == cur["kps"].tobytes() and np.array_equal(desc, cur["desc"]), f"{tag}: current frame differs"
    total = 0
    for th in ths:
        for check_ori in (True, False):
            want, wnm = oracle.search_by_projection(cur, last, cam, bounds, scale, th, mono=mono_flag, check_ori=check_ori)
            got, nm = ext.search_by_projection(_last_for_abi(last), cur["Tcw"], bf, mb, th, mono=mono_flag,
                                               check_orientation=check_ori, use_stereo=stereo)
            assert len(got) == len(want)
            bad = np.nonzero(got != want)[0]
            assert bad.size == 0, f"{tag} th={th} ori={check_ori}: {bad.size} slots differ, e.g. {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
            assert nm == wnm, f"{tag} th={th}: nmatches {nm} vs {wnm}"
            assert 1 <= ext.match_rounds() <= 64
            total += wnm
    return total


@pytest.mark.parametrize("seed,dz", [(1, 0.0), (2, 1.0), (3, -1.0)])
def test_kitti_stereo_tracking(seed, dz, oracle_built):
    """dz = 0: the +-1 octave rule; dz = +-1 m: the forward / backward rules (tlc.z against mb, :1524-1526)."""
    assert _check(f"kitti seed {seed} dz {dz}", oracle_built, seed=seed, dz=dz) > 2000


def test_crowded_queries_exercise_the_observation_rule(oracle_built):
    """Every map point three times, 90 % with observations: later copies must skip the keypoints earlier
    ones hold (:1583-1585), copies without observations get overwritten (:1610); more than one round."""
    from orb_slam2_test_b200 import ORBextractor  # noqa: F401
    _check("crowded", oracle_built, seed=4, copies=3, obs_rate=0.9)


def test_monocular_tum_with_distortion(oracle_built):
    """bMono, TUM1 camera with distortion (mvKeysUn != mvKeys, bounds beyond the image), no mvuRight."""
    assert _check("tum mono", oracle_built, ths=(15.0, 30.0), mono_flag=True, seed=5, shape=synth.TUM_SHAPE,
                  K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0, stereo=False, nf=1000) > 200


def test_no_map_points_and_argument_errors(oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    from orb_slam2_test_b200.extractor import BorbError
    cur, last, cam, bounds, scale = matchcase.make_case(seed=6)
    ext = ORBextractor(2000, 1.2, 8, 20, 7)
    ext(synth.make_frame(synth.KITTI_SHAPE, seed=6, frame=1))
    with pytest.raises(BorbError):  # no camera: no mvKeysUn / grid on the device
        ext.search_by_projection(_last_for_abi(last), cur["Tcw"], 386.0, 0.5, 7.0)
    ext.set_camera(matchcase.KITTI_K, (0.0, 0.0, 0.0, 0.0))
    ext(synth.make_frame(synth.KITTI_SHAPE, seed=6, frame=1))
    with pytest.raises(BorbError):  # no stereo result on this handle
        ext.search_by_projection(_last_for_abi(last), cur["Tcw"], 386.0, 0.5, 7.0, use_stereo=True)
    empty = _last_for_abi(last)
    empty["mp_flags"] = np.zeros_like(last["mp_flags"])
    got, nm = ext.search_by_projection(empty, cur["Tcw"], 386.0, 0.5, 7.0)
    assert nm == 0 and (got == -1).all() and ext.match_rounds() == 1
    bad = _last_for_abi(last)
    bad["octave"] = np.full(len(last["kps"]), 9, np.int32)
    with pytest.raises(BorbError):
        ext.search_by_projection(bad, cur["Tcw"], 386.0, 0.5, 7.0)


# ---- the local-map matcher: ORBmatcher::SearchByProjection(F, vpMapPoints, th), src/ORBmatcher.cc:59-154 ----

def _check_local(tag, oracle, ths=(1.0, 3.0, 5.0), ratios=(0.8, 0.6), **kw):
    cur, bounds, scale, mps, occ = matchcase.make_local_case(**kw)
    stereo = kw.get("stereo", True)
    K = kw.get("K", matchcase.KITTI_K)
    bf = kw.get("bf", matchcase.KITTI_BF)
    ext, right = _gpu_frame(kw.get("shape", synth.KITTI_SHAPE), kw.get("seed", 0), 1, K, kw.get("dist", (0.0, 0.0, 0.0, 0.0)),
                            stereo, kw.get("nf", 2000), bf, bf / K[0])
    kps, desc = ext.result(0)
    assert kps.tobytes() == cur["kps"].tobytes() and np.array_equal(desc, cur["desc"]), f"{tag}: current frame differs"
    total = 0
    for th in ths:
        for ratio in ratios:
            want, wnm = oracle.search_local_points(cur, bounds, scale, mps, occ, th, ratio)
            got, nm = ext.search_local_points(mps, occ, th, ratio, use_stereo=stereo)
            bad = np.nonzero(got != want)[0]
            assert bad.size == 0, f"{tag} th={th} ratio={ratio}: {bad.size} slots differ, e.g. {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
            assert nm == wnm, f"{tag} th={th} ratio={ratio}: nmatches {nm} vs {wnm}"
            total += wnm
    return total


def test_long_displacement_chains_are_finished_sequentially(oracle_built, capsys):
    """Crowded queries (every map point 3 / 12 / 40 times, all with observations): the round counts of the fixed point
    are reported, and with the round limit lowered to 2 / 4 the still-open suffix is walked in the reference's order by
    one thread -- same answer for the frame-to-frame and the local-map matcher."""
    from orb_slam2_test_b200.extractor import load_library
    lib = load_library()
    try:
        for copies in (3, 12, 40):
            cur, last, cam, bounds, scale = matchcase.make_case(seed=4, copies=copies, obs_rate=1.0)
            bf, mb = float(cam[4]), float(cam[5])
            ext, right = _gpu_frame(synth.KITTI_SHAPE, 4, 1, matchcase.KITTI_K, (0.0, 0.0, 0.0, 0.0), True, 2000, bf, mb)
            for th in (7.0, 30.0):
                want, wnm = oracle_built.search_by_projection(cur, last, cam, bounds, scale, th)
                for limit in (0, 4, 2):
                    lib.borb_debug_set_match_rounds(limit)
                    got, nm = ext.search_by_projection(_last_for_abi(last), cur["Tcw"], bf, mb, th, use_stereo=True)
                    assert np.array_equal(got, want) and nm == wnm, f"copies={copies} th={th} limit={limit}"
                    if limit == 0:
                        assert ext.match_rounds() <= 16
                        with capsys.disabled():
                            print(f"\n[info] frame-to-frame matcher, every map point x{copies}, th {th:.0f}: {ext.match_rounds()} rounds, "
                                  f"{ext.match_rescans()} rescans, {nm} matches")
                    else:
                        assert ext.match_rounds() == limit + 1  # the sequential finish ran
        cur, bounds, scale, mps, occ = matchcase.make_local_case(seed=3, copies=4, obs_rate=1.0)
        ext, right = _gpu_frame(synth.KITTI_SHAPE, 3, 1, matchcase.KITTI_K, (0.0, 0.0, 0.0, 0.0), True, 2000, matchcase.KITTI_BF,
                                matchcase.KITTI_BF / matchcase.KITTI_K[0])
        want, wnm = oracle_built.search_local_points(cur, bounds, scale, mps, occ, 3.0, 0.8)
        for limit in (0, 2):
            lib.borb_debug_set_match_rounds(limit)
            got, nm = ext.search_local_points(mps, occ, 3.0, 0.8, use_stereo=True)
            assert np.array_equal(got, want) and nm == wnm, f"local map, limit={limit}"
    finally:
        lib.borb_debug_set_match_rounds(0)


@pytest.mark.parametrize("seed,copies,obs", [(1, 2, 0.9), (2, 3, 0.5), (3, 4, 1.0)])
def test_local_map_kitti(seed, copies, obs, oracle_built):
    assert _check_local(f"local kitti {seed}", oracle_built, seed=seed, copies=copies, obs_rate=obs) > 2000


def test_local_map_monocular_tum(oracle_built):
    assert _check_local("local tum", oracle_built, seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D,
                        bf=40.0, stereo=False, nf=1000) > 1000


def test_local_map_without_occupancy(oracle_built):
    """occupied = NULL: every slot is free (a frame nobody matched yet)"""
    cur, bounds, scale, mps, occ = matchcase.make_local_case(seed=6)
    ext, _ = _gpu_frame(synth.KITTI_SHAPE, 6, 1, matchcase.KITTI_K, (0.0, 0.0, 0.0, 0.0), True, 2000, matchcase.KITTI_BF,
                        matchcase.KITTI_BF / matchcase.KITTI_K[0])
    want, wnm = oracle_built.search_local_points(cur, bounds, scale, mps, np.zeros(len(occ), np.uint8), 3.0, 0.8)
    got, nm = ext.search_local_points(mps, None, 3.0, 0.8, use_stereo=True)
    assert np.array_equal(got, want) and nm == wnm > 500


# ---- the monocular-initialisation matcher: ORBmatcher::SearchForInitialization, src/ORBmatcher.cc:487-632 ----

def _check_init(tag, oracle, windows=(100, 30, 300), **kw):
    from orb_slam2_test_b200 import ORBextractor
    f1, f2, bounds, prev = matchcase.make_init_case(**kw)
    shape = kw.get("shape", synth.TUM_SHAPE)
    ext = ORBextractor(kw.get("nf", 2000), 1.2, 8, 20, 7)
    ext.set_camera(kw.get("K", matchcase.TUM1_K), kw.get("dist", matchcase.TUM1_D))
    kps, desc = ext(synth.make_frame(shape, seed=kw.get("seed", 0), frame=kw.get("frame", 0) + kw.get("step", 1)))
    assert kps.tobytes() == f2["kps"].tobytes() and np.array_equal(desc, f2["desc"]), f"{tag}: current frame differs"
    total = 0
    for win in windows:
        for ratio, ori in ((0.9, True), (0.7, False)):
            want, wprev, wnm = oracle.search_for_initialization(f1, f2, bounds, prev, win, ratio, ori)
            got, gprev, nm = ext.search_for_initialization(f1, prev, win, ratio, check_orientation=ori)
            bad = np.nonzero(got != want)[0]
            assert bad.size == 0, f"{tag} win={win} ratio={ratio}: {bad.size} entries differ, e.g. {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
            assert nm == wnm and np.array_equal(gprev.view(np.uint32), wprev.view(np.uint32)), f"{tag} win={win}: nmatches {nm} vs {wnm}"
            total += wnm
    return total


@pytest.mark.parametrize("kw", [dict(seed=1), dict(seed=2, step=3), dict(seed=4, nf=4000, step=2),
                                dict(seed=3, shape=synth.KITTI_SHAPE, K=matchcase.KITTI_K, dist=(0.0, 0.0, 0.0, 0.0))],
                         ids=["tum", "tum-3-frames-later", "tum-4000", "kitti"])
def test_initialisation_matcher(kw, oracle_built):
    assert _check_init("init", oracle_built, **kw) > 800


@pytest.mark.parametrize("copies", [2, 4, 48])
def test_initialisation_matcher_take_overs(copies, oracle_built):
    """several initial keypoints want the same current keypoint at different distances: a later, strictly better
    one takes it over (:552-558), an equal or worse one is skipped (:526-527).  With 48 copies a keypoint has up to
    48 choosers in a round (repetitive texture): the chooser lists have no length limit."""
    _check_init(f"init x{copies}", oracle_built, windows=(100, 40), seed=9, copies=copies)


# ---- the relocalisation matcher: ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist),
# ---- src/ORBmatcher.cc:1670-1798 (borb_search_by_projection_keyframe) ----

def _check_reloc(tag, oracle, settings=((10.0, 100), (3.0, 64)), look_away=False, **kw):
    cur, cam, bounds, scale, log_scale, kf, occupied = matchcase.make_reloc_case(**kw)
    if look_away:
        cur["Tcw"] = matchcase.compose(matchcase.pose(ry=np.pi), cur["Tcw"])
    ext, _ = _gpu_frame(kw.get("shape", synth.KITTI_SHAPE), kw.get("seed", 0), kw.get("frame", 0) + 1,
                        kw.get("K", matchcase.KITTI_K), kw.get("dist", (0.0, 0.0, 0.0, 0.0)), False, kw.get("nf", 2000), 0.0, 0.0)
    kps, desc = ext.result(0)
    assert kps.tobytes() == cur["kps"].tobytes() and np.array_equal(desc, cur["desc"]), f"{tag}: current frame differs"
    total = 0
    for th, od in settings:
        for ori in (True, False):
            for occ in (occupied, None):
                want, wnm = oracle.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occ, th, od, ori)
                got, nm = ext.search_by_projection_keyframe(kf, cur["Tcw"], occ, th, od, check_orientation=ori)
                assert len(got) == len(want)
                bad = np.nonzero(got != want)[0]
                assert bad.size == 0, f"{tag} th={th} d={od} ori={ori}: {bad.size} slots differ, e.g. {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
                assert nm == wnm, f"{tag} th={th} d={od}: nmatches {nm} vs {wnm}"
                total += nm
    return total, ext


@pytest.mark.parametrize("seed,dz", [(1, 0.0), (2, 2.0), (3, -2.0), (6, 0.7)])
def test_relocalisation_matcher_kitti(seed, dz, oracle_built):
    assert _check_reloc(f"reloc kitti {seed} dz {dz}", oracle_built, seed=seed, dz=dz)[0] > 1000


def test_relocalisation_matcher_crowded_and_the_sequential_finish(oracle_built):
    """three and twelve copies of every keyframe point (a taken slot is closed to every later query, :1735-1736); with the
    round limit lowered the open suffix is walked by one thread in the reference's order"""
    from orb_slam2_test_b200.extractor import load_library
    lib = load_library()
    _check_reloc("reloc crowded x3", oracle_built, seed=4, copies=3, found_rate=0.1)
    try:
        lib.borb_debug_set_match_rounds(2)
        _, ext = _check_reloc("reloc crowded x12, 2 rounds", oracle_built, settings=((10.0, 100),), seed=5, copies=12,
                              found_rate=0.0, bad_rate=0.0)
        assert ext.match_rounds() == 3
    finally:
        lib.borb_debug_set_match_rounds(0)


def test_relocalisation_matcher_monocular_tum_with_distortion(oracle_built):
    assert _check_reloc("reloc tum", oracle_built, seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D,
                        bf=40.0, stereo=False, nf=1000, dist_jitter=0.4)[0] > 500


def test_relocalisation_matcher_pose_looking_away(oracle_built):
    """this matcher has no invzc < 0 test: points behind the camera that project inside the image are searched"""
    _check_reloc("reloc away", oracle_built, settings=((10.0, 100),), look_away=True, seed=9)


def test_relocalisation_matcher_empty_and_argument_errors(oracle_built):
    from orb_slam2_test_b200 import BorbError
    cur, cam, bounds, scale, log_scale, kf, occupied = matchcase.make_reloc_case(seed=1)
    ext, _ = _gpu_frame(synth.KITTI_SHAPE, 1, 1, matchcase.KITTI_K, (0.0, 0.0, 0.0, 0.0), False, 2000, 0.0, 0.0)
    none = {k: v[:0] for k, v in kf.items()}
    m, nm = ext.search_by_projection_keyframe(none, cur["Tcw"], occupied)
    assert nm == 0 and (m == -1).all() and len(m) == len(cur["kps"])
    nothing = dict(kf)
    nothing["flags"] = kf["flags"] | 2  # every point bad
    m, nm = ext.search_by_projection_keyframe(nothing, cur["Tcw"], None)
    assert nm == 0 and (m == -1).all()
    everything_taken = np.ones(len(cur["kps"]), np.uint8)
    m, nm = ext.search_by_projection_keyframe(kf, cur["Tcw"], everything_taken)
    assert nm == 0 and (m == -1).all()
    for bad_dist in (-1, 256):
        with pytest.raises(BorbError):
            ext.search_by_projection_keyframe(kf, cur["Tcw"], None, 10.0, bad_dist)
    with pytest.raises(BorbError):
        ext.search_by_projection_keyframe(kf, cur["Tcw"], None, frame=3)
    no_cam = __import__("orb_slam2_test_b200").ORBextractor(2000, 1.2, 8, 20, 7)
    no_cam(synth.make_frame(synth.KITTI_SHAPE, seed=1, frame=1, eye=0))
    with pytest.raises(BorbError):
        no_cam.search_by_projection_keyframe(kf, cur["Tcw"], None)


# ---- keyframes on the device (borb_load_frame) and ORBmatcher::Fuse(pKF, vpMapPoints, th), src/ORBmatcher.cc:968-1131 ----

def test_a_loaded_keyframe_is_the_extracted_frame(oracle_built):
    """borb_load_frame: the records and descriptors of a frame extracted earlier, put back into a handle, give the same
    mvKeysUn / mGrid / bounds and the same matcher results as the frame the extract call left there"""
    from orb_slam2_test_b200 import ORBextractor
    cur, last, cam, bounds, scale = matchcase.make_case(seed=6, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D,
                                                        bf=40.0, stereo=False, nf=1000)
    ext, _ = _gpu_frame(synth.TUM_SHAPE, 6, 1, matchcase.TUM1_K, matchcase.TUM1_D, False, 1000, 0.0, 0.0)
    kps, desc = ext.result(0)
    want = ext.search_by_projection(_last_for_abi(last), cur["Tcw"], 40.0, float(cam[5]), 15.0, mono=True)
    grid = ext.frame_grid(0)
    kf = ORBextractor(1000, 1.2, 8, 20, 7)
    kf.set_camera(matchcase.TUM1_K, matchcase.TUM1_D)
    kf.load_frame(kps, desc, 640, 480)
    got = kf.search_by_projection(_last_for_abi(last), cur["Tcw"], 40.0, float(cam[5]), 15.0, mono=True)
    assert np.array_equal(got[0], want[0]) and got[1] == want[1] and got[1] > 200
    kf.fetch()
    k2, d2 = kf.result(0)
    assert k2.tobytes() == kps.tobytes() and np.array_equal(d2, desc)
    g2 = kf.frame_grid(0)
    for a, b in zip(grid, g2):
        assert np.asarray(a).tobytes() == np.asarray(b).tobytes()
    ext(synth.make_frame(synth.TUM_SHAPE, seed=6, frame=3))  # the handle takes extract calls again afterwards, and loads again
    kf.load_frame(kps[:10], desc[:10], 640, 480)
    kf.fetch()
    assert kf.count(0) == 10


def _check_fuse(tag, oracle, ths=(3.0, 6.0), **kw):
    from orb_slam2_test_b200 import ORBextractor
    kf, cam, bounds, scale, inv_sigma2, log_scale, mps = matchcase.make_fuse_case(**kw)
    stereo = kw.get("stereo", True)
    h = ORBextractor(kw.get("nf", 2000), 1.2, 8, 20, 7)
    h.set_camera(kw.get("K", matchcase.KITTI_K), kw.get("dist", (0.0, 0.0, 0.0, 0.0)))
    shape = kw.get("shape", synth.KITTI_SHAPE)
    h.load_frame(kf["kps"], kf["desc"], shape[1], shape[0], kf["u_right"] if stereo else None)
    Ow = oracle.camera_centre(kf["Tcw"])
    total = 0
    for th in ths:
        want, wnf = oracle.fuse(kf, cam, bounds, scale, inv_sigma2, log_scale, mps, th)
        got, nf = h.fuse_search(mps, kf["Tcw"], Ow, float(cam[4]), th, use_stereo=stereo)
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0 and nf == wnf, f"{tag} th={th}: {bad.size} map points differ, e.g. {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
        total += nf
    return total


@pytest.mark.parametrize("seed,dz", [(1, 0.0), (2, 0.3), (3, -0.3)])
def test_fuse_search_kitti_stereo_keyframe(seed, dz, oracle_built):
    assert _check_fuse(f"fuse kitti {seed} dz {dz}", oracle_built, seed=seed, dz=dz) > 400


def test_fuse_search_crowded(oracle_built):
    assert _check_fuse("fuse crowded", oracle_built, seed=4, copies=3) > 2000


def test_fuse_search_monocular_distorted_keyframe(oracle_built):
    """the keyframe's bounds are the frame's float bounds truncated to int"""
    assert _check_fuse("fuse tum", oracle_built, seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D,
                       bf=40.0, stereo=False, nf=1000, dist_jitter=0.4) > 300


# ---- LoopClosing's matchers on the device: Fuse(pKF, Scw, ...) :1133-1260, SearchByProjection(pKF, Scw, ...) :353-485,
# ---- SearchBySim3 :1262-1501 ----

def _loaded(kf, shape, K, dist, nf):
    from orb_slam2_test_b200 import ORBextractor
    h = ORBextractor(nf, 1.2, 8, 20, 7)
    h.set_camera(K, dist)
    h.load_frame(kf["kps"], kf["desc"], shape[1], shape[0])
    return h


@pytest.mark.parametrize("kw", [dict(seed=1), dict(seed=2, s=0.97, dz=0.3), dict(seed=4, copies=3),
                                dict(seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0,
                                     stereo=False, nf=1000, dist_jitter=0.4)])
def test_fuse_sim3_and_projection_sim3_on_the_device(kw, oracle_built):
    kf, cam, bounds, scale, inv_sigma2, log_scale, Scw, pts, occupied = matchcase.make_sim3_points_case(**kw)
    h = _loaded(kf, kw.get("shape", synth.KITTI_SHAPE), kw.get("K", matchcase.KITTI_K), kw.get("dist", (0.0, 0.0, 0.0, 0.0)),
                kw.get("nf", 2000))
    T, Ow = oracle_built.sim3_decompose(Scw)
    total = 0
    for th in (4.0, 10.0):
        want, wnf = oracle_built.fuse_sim3(kf, cam, bounds, scale, log_scale, Scw, pts, th)
        got, nf = h.project_search(pts, T, None, Ow, th, 50)
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0 and nf == wnf, f"fuse sim3 th={th}: {bad.size} differ, e.g. {bad[:5]}: {got[bad[:5]]} vs {want[bad[:5]]}"
        total += nf
        for occ in (occupied, None):
            p = pts if occ is not None else dict(pts, flags=(pts["flags"] & 3).astype(np.uint8))
            want, wnm = oracle_built.search_by_projection_kf_sim3(kf, cam, bounds, scale, log_scale, Scw, p, occ, int(th))
            got, nm = h.search_by_projection_sim3(p, T, Ow, occ, int(th))
            bad = np.nonzero(got != want)[0]
            assert bad.size == 0 and nm == wnm, f"projection sim3 th={th}: {bad.size} differ, nm {nm} vs {wnm}"
            total += nm
    assert total > 300


@pytest.mark.parametrize("kw", [dict(seed=1), dict(seed=2, s12=0.96, dz=0.3), dict(seed=3, dz=-0.3, pre_rate=0.3),
                                dict(seed=5, shape=synth.TUM_SHAPE, K=matchcase.TUM1_K, dist=matchcase.TUM1_D, bf=40.0,
                                     stereo=False, nf=1000, dist_jitter=0.4)])
def test_search_by_sim3_on_the_device(kw, oracle_built):
    k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12 = matchcase.make_sim3_pair_case(**kw)
    args = (kw.get("shape", synth.KITTI_SHAPE), kw.get("K", matchcase.KITTI_K), kw.get("dist", (0.0, 0.0, 0.0, 0.0)), kw.get("nf", 2000))
    h1, h2 = _loaded(k1, *args), _loaded(k2, *args)
    T12, T21 = oracle_built.sim3_relative(s12, R12, t12)
    total = 0
    for th in (7.5, 3.0):
        want, wnf = oracle_built.search_by_sim3(k1, k2, cam, bounds, scale, log_scale, p1, p2, pre12, s12, R12, t12, th)
        got, nf = h1.search_by_sim3(h2, p1, p2, k1["Tcw"], k2["Tcw"], T12, T21, th, pre12)
        bad = np.nonzero(got != want)[0]
        assert bad.size == 0 and nf == wnf, f"sim3 th={th}: {bad.size} differ, nFound {nf} vs {wnf}"
        total += nf
    assert total > 100


@pytest.mark.parametrize("seed", [11, 12, 13, 14, 15, 16])
def test_back_end_matchers_over_more_scenes(seed, oracle_built):
    """Relocalisation, Fuse and SearchBySim3 once more, on scenes and pose offsets the cases above do not use (the
    offsets sit between theirs, so scale prediction lands on other sides of the octave boundaries): still identical."""
    dz = (-1.3, -0.45, 0.15, 0.55, 1.2, 1.7)[seed - 11]
    assert _check_reloc(f"reloc sweep {seed}", oracle_built, settings=((10.0, 100),), seed=seed, dz=dz)[0] > 300
    assert _check_fuse(f"fuse sweep {seed}", oracle_built, ths=(3.0,), seed=seed, dz=0.25 * dz) > 50
    k1, k2, cam, bounds, scale, inv_sigma2, log_scale, p1, p2, pre12, s12, R12, t12 = matchcase.make_sim3_pair_case(
        seed=seed, s12=1.0 + 0.02 * dz, dz=0.2 * dz)
    args = (synth.KITTI_SHAPE, matchcase.KITTI_K, (0.0, 0.0, 0.0, 0.0), 2000)
    h1, h2 = _loaded(k1, *args), _loaded(k2, *args)
    T12, T21 = oracle_built.sim3_relative(s12, R12, t12)
    want, wnf = oracle_built.search_by_sim3(k1, k2, cam, bounds, scale, log_scale, p1, p2, pre12, s12, R12, t12, 7.5)
    got, nf = h1.search_by_sim3(h2, p1, p2, k1["Tcw"], k2["Tcw"], T12, T21, 7.5, pre12)
    bad = np.nonzero(got != want)[0]
    assert bad.size == 0 and nf == wnf, f"sim3 sweep {seed}: {bad.size} differ, nFound {nf} vs {wnf}"
