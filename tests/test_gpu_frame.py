"""GPU parity for the per-keypoint step after the extractor: keypoint
undistortion (Frame::UndistortKeyPoints, src/Frame.cc:542-572), image bounds (:574-609) and the
64 x 48 grid (:292-307, :510-520), against oracle/frame_oracle.c (undistortPoints pinned bit-exact to
cv2 4.13.0 by tests/test_oracle_frame.py).  Bar: float32 bit patterns and the grid lists equal."""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth
from test_oracle_frame import CAMS

pytestmark = pytest.mark.gpu


def _check(ext, oracle, f, cols, rows, K, dist, tag):
    kps, _ = ext.result(f)
    un, start, idx, bounds = ext.frame_grid(f)
    wun, wstart, widx, wbounds = oracle.frame_post(kps, cols, rows, K, dist)
    assert np.array_equal(bounds.view(np.uint32), wbounds.view(np.uint32)), f"{tag}: bounds {bounds} vs {wbounds}"
    bad = np.nonzero((un.view(np.uint32) != wun.view(np.uint32)).any(axis=1))[0]
    assert bad.size == 0, f"{tag}: {bad.size} undistorted points differ, e.g. {un[bad[:3]]} vs {wun[bad[:3]]}"
    assert np.array_equal(start, wstart), f"{tag}: cell_start"
    assert np.array_equal(idx, widx), f"{tag}: cell_idx"
    return int(start[-1]), len(kps)


@pytest.mark.parametrize("name", ["TUM1", "TUM2", "strong"])
def test_tum_distortion(name, oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    K, dist = CAMS[name]
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=3)
    ext.set_camera(K, dist)
    imgs = synth.make_sequence(synth.TUM_SHAPE, seed=51, start=0, count=3)
    ext.extract_batch(imgs)
    for f in range(3):
        in_grid, n = _check(ext, oracle_built, f, 640, 480, K, dist, f"{name} frame {f}")
        assert 0.5 * n <= in_grid <= n


def test_no_distortion_kitti(oracle_built):
    """k1 == 0: mvKeysUn = mvKeys, bounds = the image rectangle (src/Frame.cc:544-548, :601-607)."""
    from orb_slam2_test_b200 import ORBextractor
    K, dist = (718.856, 718.856, 607.1928, 185.2157), (0.0, 0.0, 0.0, 0.0)
    ext = ORBextractor(2000, 1.2, 8, 20, 7)
    ext.set_camera(K, dist)
    kps, _ = ext(synth.make_frame(synth.KITTI_SHAPE, seed=52))
    _check(ext, oracle_built, 0, 1241, 376, K, dist, "kitti")
    un, _, _, bounds = ext.frame_grid(0)
    assert np.array_equal(un[:, 0], kps["x"]) and np.array_equal(un[:, 1], kps["y"])
    assert bounds.tolist() == [0.0, 1241.0, 0.0, 376.0]


def test_four_coefficients_and_dense_cells(oracle_built):
    """EuRoC-style 4-coefficient model; a noise frame with many keypoints per grid cell."""
    from orb_slam2_test_b200 import ORBextractor
    K, dist = CAMS["EuRoC"]
    ext = ORBextractor(8000, 1.2, 8, 20, 7)
    ext.set_camera(K, dist)
    ext(synth.make_noise_frame(synth.EUROC_SHAPE, seed=2))
    in_grid, n = _check(ext, oracle_built, 0, 752, 480, K, dist, "noise")
    assert n > 6000 and in_grid > 5000


def test_device_path_and_switching_off(oracle_built):
    import torch
    from orb_slam2_test_b200 import ORBextractor
    from orb_slam2_test_b200.extractor import BorbError
    K, dist = CAMS["TUM1"]
    imgs = synth.make_sequence(synth.TUM_SHAPE, seed=53, start=0, count=4)
    dev = torch.from_numpy(imgs).cuda()
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=4)
    ext.set_camera(K, dist)
    for _ in range(2):  # second round replays the cached graph
        ext.extract_device(dev.data_ptr(), 4, 640, 480)
        ext.fetch()
        for f in range(4):
            _check(ext, oracle_built, f, 640, 480, K, dist, f"device frame {f}")
    ext.set_camera(None)
    ext.extract_batch(imgs)
    with pytest.raises(BorbError):
        ext.frame_grid(0)
