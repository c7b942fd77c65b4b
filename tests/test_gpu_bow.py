"""Bag of words on the device against the CPU restatement (itself equal to the reference's own DBoW2 + Frame.cc +
ORBmatcher.cc object code, tests/test_oracle_bow_vs_reference.py): Frame::ComputeBoW (src/Frame.cc:532-539) and
ORBmatcher::SearchByBoW(KeyFrame*, Frame&) (src/ORBmatcher.cc:195-351), through the C ABI.  Word values are compared
as float64 BIT PATTERNS."""
import numpy as np
import pytest

import bowcase
from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def vocdir(tmp_path_factory):
    return tmp_path_factory.mktemp("voc")


@pytest.mark.parametrize("name", [v[0] for v in bowcase.VOCABULARIES])
def test_transform_of_host_descriptors(name, vocdir, oracle_built):
    from oracle import pyoracle
    from orb_slam2_test_b200 import ORBVocabulary
    path = bowcase.vocabulary(vocdir, name)
    ours, want = ORBVocabulary(path), pyoracle.BowVocabulary(path)
    assert (ours.k, ours.L, ours.n_nodes, ours.n_words, ours.scoring, ours.weighting) == \
           (want.k, want.L, want.n_nodes, want.n_words, want.scoring, want.weighting)
    sets = [bowcase.frame_features(synth.TUM_SHAPE, 1, 0, 1000)[1], bowcase.frame_features(synth.KITTI_SHAPE, 2, 3, 2000)[1],
            bowcase.word_like(name, 3000, 5), bowcase.word_like(name, 500, 6, noise_bits=0), np.zeros((0, 32), np.uint8),
            np.zeros((7, 32), np.uint8), bowcase.word_like(name, 16384, 7)]
    for d in sets:
        for levelsup in (4, 2, 0, 12):
            assert bowcase.same_bow(ours.transform(d, levelsup), want.transform(d, levelsup)), f"{name} n={len(d)} levelsup={levelsup}"
    # the vocabulary from arrays is the vocabulary from the file
    parent, leaf, desc, weight = want.arrays()
    again = ORBVocabulary(arrays=(want.k, want.L, want.scoring, want.weighting, parent, leaf, desc, weight))
    assert bowcase.same_bow(again.transform(sets[2], 4), want.transform(sets[2], 4))


@pytest.mark.parametrize("scoring", range(6))
@pytest.mark.parametrize("weighting", range(4))
def test_every_scoring_and_weighting_type(scoring, weighting, vocdir, oracle_built):
    from oracle import pyoracle
    from orb_slam2_test_b200 import ORBVocabulary
    path = bowcase.vocabulary(vocdir, "k10L3", scoring, weighting)
    ours, want = ORBVocabulary(path), pyoracle.BowVocabulary(path)
    d = bowcase.word_like("k10L3", 2500, 11)
    assert bowcase.same_bow(ours.transform(d, 4), want.transform(d, 4))
    assert bowcase.same_bow(ours.transform(d, 1), want.transform(d, 1))


def test_errors(vocdir, tmp_path):
    from orb_slam2_test_b200 import BorbError, ORBVocabulary, ORBextractor
    bad = tmp_path / "bad.txt"
    bad.write_text("30 6 0 0\n0 1 " + " ".join(["0"] * 32) + " 1.0")
    with pytest.raises(BorbError):
        ORBVocabulary(str(bad))  # k > 20: the reference's own header check
    with pytest.raises(BorbError):
        ORBVocabulary(str(tmp_path / "missing.txt"))
    voc = ORBVocabulary(bowcase.vocabulary(vocdir, "k10L3"))
    ext = ORBextractor(500, 1.2, 8, 20, 7)
    with pytest.raises(BorbError):
        ext.compute_bow(voc)  # nothing extracted yet
    ext(synth.make_frame(synth.TUM_SHAPE, seed=1, frame=0))
    with pytest.raises(BorbError):
        ext.search_by_bow({"desc": np.zeros((0, 32), np.uint8), "angle": np.zeros(0, np.float32), "mp_flags": np.zeros(0, np.uint8),
                           "fv": {"node_id": np.zeros(0, np.uint32), "node_start": np.zeros(1, np.int32), "feat_idx": np.zeros(0, np.uint32)}})
    ext.compute_bow(voc)
    ext(synth.make_frame(synth.TUM_SHAPE, seed=1, frame=1))
    with pytest.raises(BorbError):  # the bow vectors belonged to the previous extraction
        ext.search_by_bow({"desc": np.zeros((0, 32), np.uint8), "angle": np.zeros(0, np.float32), "mp_flags": np.zeros(0, np.uint8),
                           "fv": {"node_id": np.zeros(0, np.uint32), "node_start": np.zeros(1, np.int32), "feat_idx": np.zeros(0, np.uint32)}})


@pytest.mark.parametrize("name,shape,nf", [("k10L4", synth.TUM_SHAPE, 1000), ("k4L6", synth.KITTI_SHAPE, 2000),
                                           ("k9L5ragged", synth.TUM_SHAPE, 2000), ("k10L3", synth.TUM_SHAPE, 500)])
def test_compute_bow_and_search_by_bow_on_extracted_frames(name, shape, nf, vocdir, oracle_built):
    """the frame's descriptors never leave the device between the extraction, ComputeBoW and SearchByBoW"""
    from oracle import pyoracle
    from orb_slam2_test_b200 import ORBVocabulary, ORBextractor
    path = bowcase.vocabulary(vocdir, name)
    voc, want = ORBVocabulary(path), pyoracle.BowVocabulary(path)
    ext = ORBextractor(nf, 1.2, 8, 20, 7)
    total = 0
    for t in (0, 2):
        kk, kd = bowcase.frame_features(shape, 3, t, nf)          # keyframe: frame t (host side, as a KeyFrame is)
        fk, fd = bowcase.frame_features(shape, 3, t + 1, nf)      # current frame: frame t + 1, extracted on the device
        gk, gd = ext(synth.make_frame(shape, seed=3, frame=t + 1))
        assert gk.tobytes() == fk.tobytes() and np.array_equal(gd, fd)
        bow = ext.compute_bow(voc)
        assert bowcase.same_bow(bow, want.transform(fd, 4)), f"{name} t={t}: ComputeBoW differs"
        mp = bowcase.keyframe_flags(len(kk), 40 + t)
        kf = {"desc": kd, "angle": kk["angle"], "mp_flags": mp, "fv": voc.transform(kd, 4)}
        assert bowcase.same_bow(kf["fv"], want.transform(kd, 4))
        for ratio, ori in ((0.7, True), (0.9, True), (0.6, False)):
            w, wnm = want.search_by_bow(kk, kd, mp, fk, fd, ratio, ori)
            g, nm = ext.search_by_bow(kf, ratio, check_orientation=ori)
            bad = np.nonzero(g != w)[0]
            assert bad.size == 0 and nm == wnm, f"{name} t={t} ratio={ratio} ori={ori}: {bad.size} differ, nm {nm} vs {wnm}, e.g. {bad[:5]}"
            total += nm
    assert total > 100
    # an empty keyframe matches nothing
    e = {"desc": np.zeros((0, 32), np.uint8), "angle": np.zeros(0, np.float32), "mp_flags": np.zeros(0, np.uint8),
         "fv": {"node_id": np.zeros(0, np.uint32), "node_start": np.zeros(1, np.int32), "feat_idx": np.zeros(0, np.uint32)}}
    g, nm = ext.search_by_bow(e)
    assert nm == 0 and (g == -1).all()


def test_batch_frames_and_replay(vocdir, oracle_built):
    """ComputeBoW on frame 2 of a 4-frame batch, twice (same answer), then on frame 0"""
    from oracle import pyoracle
    from orb_slam2_test_b200 import ORBVocabulary, ORBextractor
    path = bowcase.vocabulary(vocdir, "k10L4")
    voc, want = ORBVocabulary(path), pyoracle.BowVocabulary(path)
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=4)
    imgs = np.stack([synth.make_frame(synth.TUM_SHAPE, seed=5, frame=f) for f in range(4)])
    res = ext.extract_batch(imgs)
    for f in (2, 2, 0, 3):
        assert bowcase.same_bow(ext.compute_bow(voc, frame=f), want.transform(res[f][1], 4)), f"frame {f}"


@pytest.mark.parametrize("name,shape,nf", [("k10L4", synth.TUM_SHAPE, 1000), ("k4L6", synth.KITTI_SHAPE, 2000),
                                           ("k9L5ragged", synth.TUM_SHAPE, 2000), ("k10L3", synth.TUM_SHAPE, 500)])
def test_search_by_bow_between_keyframes(name, shape, nf, vocdir, oracle_built):
    """ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (src/ORBmatcher.cc:634-777) through borb_voc_search_by_bow: both
    keyframes as flat arrays, their feature vectors from borb_voc_transform"""
    from oracle import pyoracle
    from orb_slam2_test_b200 import BorbError, ORBVocabulary
    path = bowcase.vocabulary(vocdir, name)
    voc, want = ORBVocabulary(path), pyoracle.BowVocabulary(path)
    total = 0
    kfs = None
    for t in (0, 2):
        k1, d1 = bowcase.frame_features(shape, 3, t, nf)
        k2, d2 = bowcase.frame_features(shape, 3, t + 1, nf)
        m1, m2 = bowcase.keyframe_flags(len(k1), 40 + t), bowcase.keyframe_flags(len(k2), 60 + t)
        kf1 = {"desc": d1, "angle": k1["angle"], "mp_flags": m1, "fv": voc.transform(d1, 4)}
        kf2 = {"desc": d2, "angle": k2["angle"], "mp_flags": m2, "fv": voc.transform(d2, 4)}
        kfs = (kf1, kf2)
        for ratio, ori in ((0.75, True), (0.9, True), (0.6, False)):
            w, wnm = want.search_by_bow_kf(k1, d1, m1, k2, d2, m2, ratio, ori)
            for rep in range(2):  # the second call reuses the scratch
                g, nm = voc.search_by_bow(kf1, kf2, ratio, check_orientation=ori)
                bad = np.nonzero(g != w)[0]
                assert bad.size == 0 and nm == wnm, f"{name} t={t} ratio={ratio} ori={ori}: {bad.size} differ, nm {nm} vs {wnm}, e.g. {bad[:5]}"
            total += nm
        # the sides swapped (another pair of sizes through the same scratch)
        w, wnm = want.search_by_bow_kf(k2, d2, m2, k1, d1, m1, 0.75, True)
        g, nm = voc.search_by_bow(kf2, kf1, 0.75)
        assert np.array_equal(g, w) and nm == wnm
    assert total > 100
    e = {"desc": np.zeros((0, 32), np.uint8), "angle": np.zeros(0, np.float32), "mp_flags": np.zeros(0, np.uint8),
         "fv": {"node_id": np.zeros(0, np.uint32), "node_start": np.zeros(1, np.int32), "feat_idx": np.zeros(0, np.uint32)}}
    g, nm = voc.search_by_bow(e, kfs[1])
    assert nm == 0 and len(g) == 0
    g, nm = voc.search_by_bow(kfs[0], e)
    assert nm == 0 and (g == -1).all() and len(g) == len(kfs[0]["desc"])
    nobody = dict(kfs[1])
    nobody["mp_flags"] = np.zeros_like(kfs[1]["mp_flags"])  # no pKF2 keypoint has a map point
    g, nm = voc.search_by_bow(kfs[0], nobody)
    assert nm == 0 and (g == -1).all()
    broken = dict(kfs[0])
    broken["fv"] = dict(kfs[0]["fv"])
    broken["fv"]["feat_idx"] = kfs[0]["fv"]["feat_idx"] + np.uint32(100000)
    with pytest.raises(BorbError):
        voc.search_by_bow(broken, kfs[1])


def _tri_abi(kf, voc):
    """a keyframe of bowcase.triangulation_case as ORBVocabulary.search_for_triangulation takes it"""
    k = kf["kps"]
    return {"desc": kf["desc"], "angle": k["angle"], "mp_flags": kf["mp"], "fv": voc.transform(kf["desc"], 4),
            "xy": np.stack([k["x"], k["y"]], 1).astype(np.float32), "octave": k["octave"].astype(np.int32),
            "u_right": kf.get("u_right")}


@pytest.mark.parametrize("name,shape,nf,kw", [
    ("k10L4", synth.TUM_SHAPE, 1000, dict()),
    ("k4L6", synth.KITTI_SHAPE, 2000, dict(baseline=(0.02, 0.01, 0.25), K=(718.856, 718.856, 607.1928, 185.2157))),
    ("k9L5ragged", synth.TUM_SHAPE, 2000, dict(stereo_rate=0.5)),
    ("k10L3", synth.TUM_SHAPE, 500, dict(baseline=(0.0, 0.0, 0.3), has=0.2)),
])
def test_search_for_triangulation(name, shape, nf, kw, vocdir, oracle_built):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:779-966) through borb_voc_search_for_triangulation against the
    restatement (equal to the reference's object code, tests/test_oracle_bow_vs_reference.py)"""
    from oracle import pyoracle
    from orb_slam2_test_b200 import BorbError, ORBVocabulary
    path = bowcase.vocabulary(vocdir, name)
    voc, want = ORBVocabulary(path), pyoracle.BowVocabulary(path)
    total = 0
    for t in (0, 2):
        kf1, kf2, geo = bowcase.triangulation_case(shape=shape, t=t, nf=nf, **kw)
        a, b = _tri_abi(kf1, voc), _tri_abi(kf2, voc)
        g = dict(geo)
        g["Ow1"], g["Tcw2"] = pyoracle.camera_centre(kf1["Tcw"]), kf2["Tcw"]
        for only_stereo, ori in ((False, True), (False, False), (True, True)):
            w, wnm = want.search_for_triangulation(kf1, kf2, geo, only_stereo, ori)
            for rep in range(2):
                got, nm = voc.search_for_triangulation(a, b, g, only_stereo, ori)
                bad = np.nonzero(got != w)[0]
                assert bad.size == 0 and nm == wnm, f"{name} t={t} stereo={only_stereo} ori={ori}: {bad.size} differ, nm {nm} vs {wnm}, e.g. {bad[:5]}"
            total += nm
        # interleaved with the keyframe-to-keyframe BoW matcher on the same scratch
        voc.search_by_bow(a, b, 0.75)
        got, nm = voc.search_for_triangulation(a, b, g)
        w, wnm = want.search_for_triangulation(kf1, kf2, geo)
        assert np.array_equal(got, w) and nm == wnm
    assert total > 60
    bad_oct = dict(b)
    bad_oct["octave"] = b["octave"] + 8
    with pytest.raises(BorbError):
        voc.search_for_triangulation(a, bad_oct, g)
