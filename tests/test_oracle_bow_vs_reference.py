"""The bag-of-words restatement (oracle/bow_oracle.cpp) against the reference's OWN object code
(oracle/_ref/librefbow.so: DBoW2's TemplatedVocabulary / FORB / BowVector / FeatureVector, Frame::ComputeBoW
src/Frame.cc:532-539 and ORBmatcher::SearchByBoW(KeyFrame*, Frame&) src/ORBmatcher.cc:195-351, compiled where they
lie).  Word values are compared as float64 BIT PATTERNS."""
import numpy as np
import pytest

import bowcase
from orb_slam2_test_b200 import synth

pyoracle = pytest.importorskip("oracle.pyoracle")
pytestmark = pytest.mark.skipif(not pyoracle.RefBow.available(), reason="oracle/_ref/librefbow.so not built (no reference tree)")


@pytest.fixture(scope="module")
def vocdir(tmp_path_factory):
    return tmp_path_factory.mktemp("voc")


@pytest.mark.parametrize("name", [v[0] for v in bowcase.VOCABULARIES])
def test_transform_equals_the_reference(name, vocdir, oracle_built):
    path = bowcase.vocabulary(vocdir, name)
    ref, ours = pyoracle.RefBow(path), pyoracle.BowVocabulary(path)
    assert (ref.k, ref.L, ref.n_words, ref.scoring, ref.weighting) == (ours.k, ours.L, ours.n_words, ours.scoring, ours.weighting)
    sets = [bowcase.frame_features(synth.TUM_SHAPE, 1, 0, 1000)[1], bowcase.frame_features(synth.KITTI_SHAPE, 2, 3, 2000)[1],
            bowcase.word_like(name, 3000, 5), bowcase.word_like(name, 500, 6, noise_bits=0), np.zeros((0, 32), np.uint8),
            np.zeros((7, 32), np.uint8)]
    for d in sets:
        for levelsup in (4, 2, 0, 12):
            a, b = ref.transform(d, levelsup), ours.transform(d, levelsup)
            assert bowcase.same_bow(a, b), f"{name} n={len(d)} levelsup={levelsup}"
    if len(sets[0]):  # Frame::ComputeBoW's output is L1-normalised (L1Scoring::mustNormalize)
        v = ours.transform(sets[0], 4)
        assert abs(v["word_val"].sum() - 1.0) < 1e-12 and len(v["feat_idx"]) <= len(sets[0])
        assert np.array_equal(np.sort(v["feat_idx"]), np.unique(v["feat_idx"]))


@pytest.mark.parametrize("scoring", range(6))
@pytest.mark.parametrize("weighting", range(4))
def test_every_scoring_and_weighting_type(scoring, weighting, vocdir, oracle_built):
    """TF-IDF / TF accumulate (BowVector::addWeight), IDF / BINARY set once (addIfNotExist); L1 / L2 / no
    normalisation by scoring type (ScoringObject.h:74-89), the divide by the number of words when not normalising."""
    path = bowcase.vocabulary(vocdir, "k10L3", scoring, weighting)
    ref, ours = pyoracle.RefBow(path), pyoracle.BowVocabulary(path)
    d = bowcase.word_like("k10L3", 2500, 11)
    assert bowcase.same_bow(ref.transform(d, 4), ours.transform(d, 4))
    assert bowcase.same_bow(ref.transform(d, 1), ours.transform(d, 1))


def test_l1_score_equals_the_reference(vocdir, oracle_built):
    path = bowcase.vocabulary(vocdir, "k10L4")
    ref, ours = pyoracle.RefBow(path), pyoracle.BowVocabulary(path)
    vs = [ours.transform(bowcase.word_like("k10L4", 1500, s), 4) for s in (1, 2)]
    vs.append(ours.transform(bowcase.frame_features(synth.TUM_SHAPE, 1, 0, 1000)[1], 4))
    for a in vs:
        for b in vs:
            assert np.float64(ref.score(a, b)).view(np.uint64) == np.float64(ours.score(a, b)).view(np.uint64)
    assert abs(ours.score(vs[0], vs[0]) - 1.0) < 1e-12


@pytest.mark.parametrize("name,shape,nf", [("k10L4", synth.TUM_SHAPE, 1000), ("k4L6", synth.KITTI_SHAPE, 2000),
                                           ("k9L5ragged", synth.TUM_SHAPE, 2000), ("k10L3", synth.TUM_SHAPE, 500)])
def test_search_by_bow_equals_the_reference(name, shape, nf, vocdir, oracle_built):
    """keyframe = frame t, current frame = frame t + 1 of the synthetic sequence; 70 % of the keyframe's keypoints
    carry a map point, 5 % of those are bad.  L = 4 puts every feature under the root (one node: the whole
    sequential rule in one list); k = 4, L = 6 gives 16 nodes."""
    path = bowcase.vocabulary(vocdir, name)
    ref, ours = pyoracle.RefBow(path), pyoracle.BowVocabulary(path)
    total = 0
    for t in (0, 2):
        kk, kd = bowcase.frame_features(shape, 3, t, nf)
        fk, fd = bowcase.frame_features(shape, 3, t + 1, nf)
        mp = bowcase.keyframe_flags(len(kk), 40 + t)
        for ratio, ori in ((0.7, True), (0.9, True), (0.6, False)):
            want, wnm = ref.search_by_bow(kk, kd, mp, fk, fd, ratio, ori)
            got, nm = ours.search_by_bow(kk, kd, mp, fk, fd, ratio, ori)
            assert np.array_equal(got, want) and nm == wnm, f"{name} t={t} ratio={ratio} ori={ori}"
            assert nm == int((got >= 0).sum()) and not (mp[got[got >= 0]] != 1).any()
            total += nm
    assert total > 100
    # degenerate sides
    e = np.zeros((0, 32), np.uint8)
    ek = np.zeros(0, kk.dtype)
    assert ref.search_by_bow(ek, e, np.zeros(0, np.uint8), fk, fd)[1] == ours.search_by_bow(ek, e, np.zeros(0, np.uint8), fk, fd)[1] == 0
    assert ref.search_by_bow(kk, kd, mp, ek, e)[1] == ours.search_by_bow(kk, kd, mp, ek, e)[1] == 0


@pytest.mark.parametrize("name,shape,nf", [("k10L4", synth.TUM_SHAPE, 1000), ("k4L6", synth.KITTI_SHAPE, 2000),
                                           ("k9L5ragged", synth.TUM_SHAPE, 2000), ("k10L3", synth.TUM_SHAPE, 500)])
def test_search_by_bow_between_keyframes_equals_the_reference(name, shape, nf, vocdir, oracle_built):
    """ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (src/ORBmatcher.cc:634-777, LoopClosing::ComputeSim3): keyframe 1 =
    frame t, keyframe 2 = frame t + 1 of the synthetic sequence, 70 % of the keypoints of either carry a map point, 5 % of
    those are bad -- the pKF2 side skips them too (:686-690) and the distance test is strict (:712)."""
    path = bowcase.vocabulary(vocdir, name)
    ref, ours = pyoracle.RefBow(path), pyoracle.BowVocabulary(path)
    total = 0
    for t in (0, 2):
        k1, d1 = bowcase.frame_features(shape, 3, t, nf)
        k2, d2 = bowcase.frame_features(shape, 3, t + 1, nf)
        m1, m2 = bowcase.keyframe_flags(len(k1), 40 + t), bowcase.keyframe_flags(len(k2), 60 + t)
        for ratio, ori in ((0.75, True), (0.9, True), (0.6, False)):
            want, wnm = ref.search_by_bow_kf(k1, d1, m1, k2, d2, m2, ratio, ori)
            got, nm = ours.search_by_bow_kf(k1, d1, m1, k2, d2, m2, ratio, ori)
            assert np.array_equal(got, want) and nm == wnm, f"{name} t={t} ratio={ratio} ori={ori}"
            hit = got >= 0
            assert nm == int(hit.sum()) and (m1[hit] == 1).all() and (m2[got[hit]] == 1).all()
            assert len(np.unique(got[hit])) == nm  # a pKF2 keypoint is matched once
            total += nm
    assert total > 100
    e = np.zeros((0, 32), np.uint8)
    ek = np.zeros(0, k1.dtype)
    z = np.zeros(0, np.uint8)
    assert ref.search_by_bow_kf(ek, e, z, k2, d2, m2)[1] == ours.search_by_bow_kf(ek, e, z, k2, d2, m2)[1] == 0
    assert ref.search_by_bow_kf(k1, d1, m1, ek, e, z)[1] == ours.search_by_bow_kf(k1, d1, m1, ek, e, z)[1] == 0


@pytest.mark.parametrize("name,shape,nf,kw", [
    ("k10L4", synth.TUM_SHAPE, 1000, dict()),
    ("k4L6", synth.KITTI_SHAPE, 2000, dict(baseline=(0.02, 0.01, 0.25), K=(718.856, 718.856, 607.1928, 185.2157))),
    ("k9L5ragged", synth.TUM_SHAPE, 2000, dict(stereo_rate=0.5)),
    ("k10L3", synth.TUM_SHAPE, 500, dict(baseline=(0.0, 0.0, 0.3), has=0.2)),
])
def test_search_for_triangulation_equals_the_reference(name, shape, nf, kw, vocdir, oracle_built):
    """ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:779-966) + CheckDistEpipolarLine (:165-182): sideways and
    forward baselines (the epipole inside the image: the 100 * scale exclusion zone, :874-879), half of the keypoints
    stereo, bOnlyStereo, orientation check on and off"""
    path = bowcase.vocabulary(vocdir, name)
    ref, ours = pyoracle.RefBow(path), pyoracle.BowVocabulary(path)
    total = 0
    for t in (0, 2):
        kf1, kf2, geo = bowcase.triangulation_case(shape=shape, t=t, nf=nf, **kw)
        for only_stereo, ori in ((False, True), (False, False), (True, True)):
            want, wnm = ref.search_for_triangulation(kf1, kf2, geo, only_stereo, ori)
            got, nm = ours.search_for_triangulation(kf1, kf2, geo, only_stereo, ori)
            assert np.array_equal(got, want) and nm == wnm, f"{name} t={t} stereo={only_stereo} ori={ori}: {(got != want).sum()} differ"
            hit = got >= 0
            assert nm == int(hit.sum()) and not (kf1["mp"][hit] & 1).any() and not (kf2["mp"][got[hit]] & 1).any()
            total += nm
    assert total > 60
