"""GPU parity for the stereo association: borb_stereo_match against the CPU
restatement of Frame::ComputeStereoMatches (oracle/stereo_oracle.cpp, src/Frame.cc:619-834).

Bar: mvuRight and mvDepth bit-exact (float32 bit patterns), same matched count.  The oracle for
this row is pinned to the reference's own src/Frame.cc compiled in place (oracle/_ref/libreframe.so,
tests/test_oracle_frame_vs_reference.py); tests/test_golden_frame.py compares the CUDA path with the
reference Frame's outputs directly.
"""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu

BF = 386.1448          # KITTI00-02.yaml Camera.bf
FX = 718.856
MB = BF / FX           # Frame.cc:87 mb = mbf / fx


def _extractors(nf=2000, max_batch=1):
    from orb_slam2_test_b200 import ORBextractor
    return (ORBextractor(nf, 1.2, 8, 20, 7, max_batch=max_batch),
            ORBextractor(nf, 1.2, 8, 20, 7, max_batch=max_batch))


def _oracle_for(left, right, frame, mbf, mb):
    from oracle import pyoracle
    kl, dl = left.result(frame)
    kr, dr = right.result(frame)
    pl = [left.pyramid_level(frame, l) for l in range(8)]
    pr = [right.pyramid_level(frame, l) for l in range(8)]
    return pyoracle.stereo_match(kl, dl, kr, dr, pl, pr, left.GetScaleFactors(), left.GetInverseScaleFactors(),
                                 mbf, mb)


def _check(left, right, n, mbf, mb, tag):
    left.stereo_match(right, mbf, mb)
    total = 0
    for f in range(n):
        u, z, kept = left.stereo_result(f)
        wu, wz, _, wkept = _oracle_for(left, right, f, mbf, mb)
        assert len(u) == len(wu)
        bad = np.nonzero(u.view(np.uint32) != wu.view(np.uint32))[0]
        assert bad.size == 0, f"{tag} frame {f}: uRight differs at {bad[:5]}: {u[bad[:5]]} vs {wu[bad[:5]]}"
        bad = np.nonzero(z.view(np.uint32) != wz.view(np.uint32))[0]
        assert bad.size == 0, f"{tag} frame {f}: depth differs at {bad[:5]}: {z[bad[:5]]} vs {wz[bad[:5]]}"
        assert kept == wkept == int((z > 0).sum()), f"{tag} frame {f}: kept {kept} vs {wkept}"
        total += kept
    return total


@pytest.mark.parametrize("shape,nf", [(synth.KITTI_SHAPE, 2000), (synth.TUM_SHAPE, 1000)])
def test_stereo_pairs(shape, nf, oracle_built):
    n = 3
    left, right = _extractors(nf, max_batch=n)
    li = synth.make_sequence(shape, seed=31, start=0, count=n, eye=0)
    ri = synth.make_sequence(shape, seed=31, start=0, count=n, eye=1)
    left.extract_batch(li, want_pyramid=True)
    right.extract_batch(ri, want_pyramid=True)
    matched = _check(left, right, n, BF, MB, "pairs")
    assert matched > 100 * n  # the synthetic right eye is a shifted left eye: most points match


def test_identical_images_are_all_cut(oracle_built):
    """All SADs are 0, so the median threshold is 0 and nothing passes `dist < thDist`."""
    left, right = _extractors()
    img = synth.make_frame(synth.KITTI_SHAPE, seed=8)
    left.extract_batch(img[None], want_pyramid=True)
    right.extract_batch(img[None], want_pyramid=True)
    assert _check(left, right, 1, BF, MB, "identical") == 0


def test_zero_disparity_takes_the_epsilon_branch(oracle_built):
    """disparity 0 takes the reference's `disparity <= 0 -> 0.01` branch (src/Frame.cc:771-775)."""
    from test_oracle_stereo import _half_noisy, _mirror_tiles
    left, right = _extractors()
    img = _mirror_tiles(synth.make_frame(synth.KITTI_SHAPE, seed=9))
    left.extract_batch(img[None], want_pyramid=True)
    right.extract_batch(_half_noisy(img, 5)[None], want_pyramid=True)
    _check(left, right, 1, BF, MB, "half noisy")
    u, z, kept = left.stereo_result(0)
    k, _ = left.result(0)
    eps = z == np.float32(BF) / np.float32(0.01)
    assert eps.sum() >= 10
    assert np.all(u[eps] == (k["x"][eps].astype(np.float64) - 0.01).astype(np.float32))


def test_right_image_without_keypoints(oracle_built):
    left, right = _extractors()
    img = synth.make_frame(synth.KITTI_SHAPE, seed=9)
    left.extract_batch(img[None], want_pyramid=True)
    right.extract_batch(np.full((1,) + synth.KITTI_SHAPE, 90, np.uint8), want_pyramid=True)
    _check(left, right, 1, BF, MB, "flat right")
    u, z, kept = left.stereo_result(0)
    assert kept == 0 and np.all(u == -1) and np.all(z == -1)


@pytest.mark.parametrize("mb", [MB / 4, MB * 8])
def test_search_range(mb, oracle_built):
    """maxD = mbf / mb: a long baseline-in-pixels cuts the search range below the true disparity."""
    left, right = _extractors()
    left.extract_batch(synth.make_frame(synth.KITTI_SHAPE, seed=12, eye=0)[None], want_pyramid=True)
    right.extract_batch(synth.make_frame(synth.KITTI_SHAPE, seed=12, eye=1)[None], want_pyramid=True)
    _check(left, right, 1, BF, mb, f"mb {mb}")


def test_after_device_extract_without_pyramid_download(oracle_built):
    """The match reads the device pyramids: no BORB_WANT_PYRAMID is needed for it."""
    left, right = _extractors()
    li = synth.make_frame(synth.KITTI_SHAPE, seed=14, eye=0)
    ri = synth.make_frame(synth.KITTI_SHAPE, seed=14, eye=1)
    left.extract_batch(li[None])
    right.extract_batch(ri[None])
    left.stereo_match(right, BF, MB)
    u, z, kept = left.stereo_result(0)
    left.extract_batch(li[None], want_pyramid=True)
    right.extract_batch(ri[None], want_pyramid=True)
    wu, wz, _, wkept = _oracle_for(left, right, 0, BF, MB)
    assert np.array_equal(u.view(np.uint32), wu.view(np.uint32)) and np.array_equal(z.view(np.uint32), wz.view(np.uint32))
    assert kept == wkept


def test_argument_errors():
    from orb_slam2_test_b200 import ORBextractor
    from orb_slam2_test_b200.extractor import BorbError
    a = ORBextractor(1000, 1.2, 8, 20, 7)
    b = ORBextractor(1000, 1.2, 8, 20, 7)
    with pytest.raises(BorbError):
        a.stereo_match(b, BF, MB)            # nothing extracted yet
    a(synth.make_frame(synth.TUM_SHAPE, seed=1))
    b(synth.make_frame(synth.KITTI_SHAPE, seed=1))
    with pytest.raises(BorbError):
        a.stereo_match(b, BF, MB)            # different geometry
    with pytest.raises(BorbError):
        a.stereo_match(a, BF, MB)            # same handle
    with pytest.raises(BorbError):
        a.stereo_result(0)
