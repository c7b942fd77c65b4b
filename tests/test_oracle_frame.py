"""Pin the oracle's restatement of cv::undistortPoints (the call of Frame::UndistortKeyPoints,
src/Frame.cc:557, and Frame::ComputeImageBounds, :594) bit-exact to the cv2 4.13.0 wheel, and check
the grid assignment (src/Frame.cc:292-307, :510-520) against a direct transcription of its rule."""
import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

# Examples/Monocular/TUM1.yaml, TUM2.yaml, Examples/Stereo/EuRoC.yaml (LEFT), and a strong-distortion case
CAMS = {
    "TUM1": ((517.306408, 516.469215, 318.643040, 255.313989), (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)),
    "TUM2": ((520.908620, 521.007327, 325.141442, 249.701764), (0.231222, -0.784899, -0.003257, -0.000105, 0.917205)),
    "EuRoC": ((458.654, 457.296, 367.215, 248.375), (-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05)),
    "strong": ((300.0, 310.0, 320.0, 240.0), (-0.45, 0.25, 0.01, -0.02, -0.08)),
}


def _cv2(xy, K, dist):
    Km = np.array([[K[0], 0, K[2]], [0, K[1], K[3]], [0, 0, 1]], np.float32)
    D = np.array(dist, np.float32).reshape(-1, 1)
    return cv2.undistortPoints(xy.reshape(-1, 1, 2).copy(), Km, D, None, Km).reshape(-1, 2)


@pytest.mark.parametrize("name", list(CAMS))
def test_undistort_points_matches_cv2(name, oracle_built):
    K, dist = CAMS[name]
    rng = np.random.default_rng(5)
    xy = np.stack([rng.uniform(-20, 780, 30000), rng.uniform(-20, 520, 30000)], 1).astype(np.float32)
    xy[:4] = [[0, 0], [640, 0], [0, 480], [640, 480]]
    got = oracle_built.undistort_points(xy, K, dist)
    want = _cv2(xy, K, dist)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def test_frame_post_semantics(oracle_built):
    K, dist = CAMS["TUM1"]
    rng = np.random.default_rng(6)
    kps = np.zeros(3000, [("x", "<f4"), ("y", "<f4")])
    kps["x"] = rng.uniform(0, 640, 3000).astype(np.float32)
    kps["y"] = rng.uniform(0, 480, 3000).astype(np.float32)
    un, start, idx, bounds = oracle_built.frame_post(kps, 640, 480, K, dist)
    xy = np.stack([kps["x"], kps["y"]], 1)
    assert np.array_equal(un, _cv2(xy, K, dist))
    corners = _cv2(np.array([[0, 0], [640, 0], [0, 480], [640, 480]], np.float32), K, dist)
    want_b = [min(corners[0, 0], corners[2, 0]), max(corners[1, 0], corners[3, 0]),
              min(corners[0, 1], corners[1, 1]), max(corners[2, 1], corners[3, 1])]
    assert bounds.tolist() == [float(v) for v in want_b]
    # the grid, transcribed: posX = round((x - minX) * 64/(maxX-minX)), cell lists in index order
    inv_w = np.float32(64) / (bounds[1] - bounds[0])
    inv_h = np.float32(48) / (bounds[3] - bounds[2])
    cells = {}
    for i, (x, y) in enumerate(un):
        px = int(np.floor(abs(v := (x - bounds[0]) * inv_w) + np.float32(0.5)) * (1 if v >= 0 else -1))
        py = int(np.floor(abs(w := (y - bounds[2]) * inv_h) + np.float32(0.5)) * (1 if w >= 0 else -1))
        if 0 <= px < 64 and 0 <= py < 48:
            cells.setdefault(px * 48 + py, []).append(i)
    assert start[-1] == sum(len(v) for v in cells.values()) == len(idx)
    for c in range(64 * 48):
        assert idx[start[c]:start[c + 1]].tolist() == cells.get(c, [])
    # keypoints undistorted to outside the bounds are dropped from the grid, as PosInGrid does
    assert start[-1] <= 3000


def test_no_distortion_is_the_identity(oracle_built):
    kps = np.zeros(100, [("x", "<f4"), ("y", "<f4")])
    kps["x"] = np.linspace(0, 1240, 100, dtype=np.float32)
    kps["y"] = np.linspace(0, 375, 100, dtype=np.float32)
    un, start, idx, bounds = oracle_built.frame_post(kps, 1241, 376, (718.856, 718.856, 607.1928, 185.2157), (0, 0, 0, 0))
    assert np.array_equal(un[:, 0], kps["x"]) and np.array_equal(un[:, 1], kps["y"])
    assert bounds.tolist() == [0.0, 1241.0, 0.0, 376.0]
    assert start[-1] == 99  # x = 1240 rounds to column 64 -> outside (the reference drops it too)
