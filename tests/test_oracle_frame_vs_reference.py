"""Pin the restatements of the rows around the extractor to the reference's OWN src/Frame.cc.

oracle/_ref/libreframe.so is /root/reference/src/Frame.cc compiled where it lies (oracle/Makefile,
oracle/frame_driver.cpp): its stereo constructor runs the reference's extractor on both eyes and then
UndistortKeyPoints, ComputeStereoMatches, ComputeImageBounds and AssignFeaturesToGrid.  Fed with the
reference's own keypoints and descriptors, oracle/stereo_oracle.cpp and oracle/frame_oracle.c must
reproduce mvuRight, mvDepth, mvKeysUn, mGrid and the bounds bit for bit.  (The extractor inside the
Frame runs on freshly spawned threads with the stock allocator, so in rare quadtree ties its keypoints
differ from the canonical ones -- SURVEY.md B.4 -- which is why the comparison feeds its keypoints in.)
"""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth
from test_oracle_frame import CAMS
from test_oracle_stereo import _half_noisy, _mirror_tiles

KITTI_K = (718.856, 718.856, 607.1928, 185.2157)
BF = 386.1448


@pytest.fixture(scope="module")
def ref(oracle_built):
    if not oracle_built.RefFrame.available():
        pytest.skip("oracle/_ref/libreframe.so is not built (no /root/reference here)")
    return oracle_built


def _pyramids(py, img, nf):
    o = py.Oracle(nf, 1.2, 8, 20, 7)
    o.extract(img)
    return o, [o.pyramid(l, 0) for l in range(8)]


def _compare(py, left, right, K, dist, bf, mb, nf, tag):
    r = py.RefFrame.stereo(left, right, K, dist, bf, mb, nfeatures=nf)
    o, pl = _pyramids(py, left, nf)
    _, pr = _pyramids(py, right, nf)
    u, z, _, kept = py.stereo_match(r["kps_l"], r["desc_l"], r["kps_r"], r["desc_r"], pl, pr, o.table(0), o.table(1), bf, mb)
    assert np.array_equal(u.view(np.uint32), r["u_right"].view(np.uint32)), f"{tag}: mvuRight"
    assert np.array_equal(z.view(np.uint32), r["depth"].view(np.uint32)), f"{tag}: mvDepth"
    assert kept == int((r["depth"] > 0).sum())
    un, start, idx, b = py.frame_post(r["kps_l"], left.shape[1], left.shape[0], K, dist)
    assert np.array_equal(b.view(np.uint32), r["bounds"].view(np.uint32)), f"{tag}: bounds {b} vs {r['bounds']}"
    assert np.array_equal(un.view(np.uint32), r["un_xy"].view(np.uint32)), f"{tag}: mvKeysUn"
    assert np.array_equal(start, r["cell_start"]) and np.array_equal(idx, r["cell_idx"]), f"{tag}: mGrid"
    return r, kept


@pytest.mark.parametrize("seed", [31, 32, 33])
def test_kitti_pairs(seed, ref):
    left = synth.make_frame(synth.KITTI_SHAPE, seed=seed, eye=0)
    right = synth.make_frame(synth.KITTI_SHAPE, seed=seed, eye=1)
    _, kept = _compare(ref, left, right, KITTI_K, (0, 0, 0, 0), BF, BF / KITTI_K[0], 2000, f"kitti {seed}")
    assert kept > 800


@pytest.mark.parametrize("cam", ["TUM1", "TUM2", "strong"])
def test_distorted_camera(cam, ref):
    """UndistortKeyPoints / ComputeImageBounds take their cv::undistortPoints branch."""
    K, dist = CAMS[cam]
    left = synth.make_frame(synth.TUM_SHAPE, seed=51, eye=0)
    right = synth.make_frame(synth.TUM_SHAPE, seed=51, eye=1)
    r, _ = _compare(ref, left, right, K, dist, 40.0, 40.0 / K[0], 1000, cam)
    assert not np.array_equal(r["un_xy"][:, 0], r["kps_l"]["x"])


@pytest.mark.parametrize("mb_scale", [0.25, 8.0, 64.0])
def test_search_range(mb_scale, ref):
    left = synth.make_frame(synth.KITTI_SHAPE, seed=34, eye=0)
    right = synth.make_frame(synth.KITTI_SHAPE, seed=34, eye=1)
    _compare(ref, left, right, KITTI_K, (0, 0, 0, 0), BF, BF / KITTI_K[0] * mb_scale, 2000, f"mb x{mb_scale}")


def test_zero_disparity_and_all_cut(ref):
    img = _mirror_tiles(synth.make_frame(synth.KITTI_SHAPE, seed=9))
    r, kept = _compare(ref, img, _half_noisy(img, 5), KITTI_K, (0, 0, 0, 0), BF, BF / KITTI_K[0], 2000, "half noisy")
    assert (r["depth"] == np.float32(BF) / np.float32(0.01)).sum() >= 5      # the `disparity <= 0 -> 0.01` branch ran
    _, kept = _compare(ref, img, img, KITTI_K, (0, 0, 0, 0), BF, BF / KITTI_K[0], 2000, "identical")
    assert kept == 0                                                          # median SAD 0 cuts everything


def test_right_eye_without_keypoints(ref):
    left = synth.make_frame(synth.KITTI_SHAPE, seed=35)
    flat = np.full(synth.KITTI_SHAPE, 90, np.uint8)
    r, kept = _compare(ref, left, flat, KITTI_K, (0, 0, 0, 0), BF, BF / KITTI_K[0], 2000, "flat right")
    assert kept == 0 and len(r["kps_r"]) == 0
