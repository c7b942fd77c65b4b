"""Deterministic synthetic grayscale frames (no datasets are available offline).

SURVEY.md section 8(d): frames are a pure function of (shape, sequence seed, frame index, eye)
through a counter-based Philox generator, so every rank / host regenerates identical bytes.
Texture = three octaves of value noise + random filled polygons (corner sources) + pixel noise,
with a low-contrast band (forces the per-cell FAST threshold-7 retry of the reference,
/root/reference/src/ORBextractor.cc:1069-1075) and a flat patch (cells with no corner at all).
"""
from __future__ import annotations

import functools

import numpy as np

TUM_SHAPE = (480, 640)     # rows, cols   Examples/Monocular/TUM1.yaml
KITTI_SHAPE = (376, 1241)  # rows, cols   Examples/Stereo/KITTI00-02.yaml:18-19


def _rng(seed: int, frame: int, eye: int) -> np.random.Generator:
    return np.random.Generator(np.random.Philox(key=[(seed << 32) ^ (frame << 1) ^ eye, 0x0B200]))


def _value_noise(rng: np.random.Generator, rows: int, cols: int, cell: int, amp: float) -> np.ndarray:
    gh, gw = rows // cell + 2, cols // cell + 2
    g = rng.uniform(-amp, amp, size=(gh, gw)).astype(np.float32)
    ys = np.arange(rows, dtype=np.float32) / cell
    xs = np.arange(cols, dtype=np.float32) / cell
    y0 = ys.astype(np.int32)
    x0 = xs.astype(np.int32)
    fy = (ys - y0)[:, None]
    fx = (xs - x0)[None, :]
    a = g[y0][:, x0]
    b = g[y0][:, x0 + 1]
    c = g[y0 + 1][:, x0]
    d = g[y0 + 1][:, x0 + 1]
    return (a * (1 - fx) + b * fx) * (1 - fy) + (c * (1 - fx) + d * fx) * fy


def _polygons(rng: np.random.Generator, img: np.ndarray, n: int) -> None:
    rows, cols = img.shape
    yy, xx = np.mgrid[0:rows, 0:cols]
    for _ in range(n):
        cx, cy = rng.uniform(0, cols), rng.uniform(0, rows)
        w, h = rng.uniform(6, 60), rng.uniform(6, 60)
        th = rng.uniform(0, np.pi)
        contrast = rng.uniform(15, 120) * (1 if rng.uniform() < 0.5 else -1)
        x0, x1 = int(max(0, cx - 45)), int(min(cols, cx + 45))
        y0, y1 = int(max(0, cy - 45)), int(min(rows, cy + 45))
        if x1 <= x0 or y1 <= y0:
            continue
        u = (xx[y0:y1, x0:x1] - cx) * np.cos(th) + (yy[y0:y1, x0:x1] - cy) * np.sin(th)
        v = -(xx[y0:y1, x0:x1] - cx) * np.sin(th) + (yy[y0:y1, x0:x1] - cy) * np.cos(th)
        if rng.uniform() < 0.6:
            m = (np.abs(u) < w / 2) & (np.abs(v) < h / 2)              # rotated rectangle
        else:
            m = (v > -h / 2) & (v < h / 2) & (np.abs(u) < (v + h / 2) * (w / (2 * h)))  # triangle
        img[y0:y1, x0:x1][m] += contrast


@functools.lru_cache(maxsize=8)
def _scene(rows: int, cols: int, seed: int, n_shapes: int) -> np.ndarray:
    """Static scene of a sequence, (rows+64) x (cols+128) float32; frames are crops of it."""
    rng = _rng(seed, 0, 0)
    img = np.full((rows + 64, cols + 128), 128, np.float32)
    R, C = img.shape
    for cell, amp in ((64, 60.0), (16, 40.0), (4, 25.0)):
        img += _value_noise(rng, R, C, cell, amp)
    _polygons(rng, img, n_shapes)
    # low-contrast band (15 % of the height) and a flat patch (5 % of the area)
    b0 = int(R * 0.55)
    b1 = b0 + int(R * 0.15)
    img[b0:b1] = 128 + (img[b0:b1] - 128) * 0.12
    ph, pw = int(R * 0.22), int(C * 0.22)
    img[8:8 + ph, C - pw - 8:C - 8] = 90.0
    img.setflags(write=False)
    return img


def make_frame(shape=KITTI_SHAPE, seed: int = 0, frame: int = 0, eye: int = 0,
               n_shapes: int = 300) -> np.ndarray:
    """One uint8 frame of `shape` = (rows, cols)."""
    rows, cols = int(shape[0]), int(shape[1])
    img = _scene(rows, cols, seed, n_shapes)  # scene content is fixed per sequence ...
    # ... the camera moves (3,1) px per frame; the right eye is shifted by a disparity
    dx = (3 * frame + (17 if eye else 0)) % 128
    dy = frame % 64
    view = img[dy:dy + rows, dx:dx + cols]
    nrng = _rng(seed, frame + 1, eye)
    view = view + nrng.uniform(-4, 4, size=view.shape).astype(np.float32)
    return np.clip(np.rint(view), 0, 255).astype(np.uint8)


def make_noise_frame(shape, seed: int = 0) -> np.ndarray:
    """Pure i.i.d. noise: the candidate-count stress case (~70 k FAST candidates at KITTI shape)."""
    rng = _rng(seed, 0, 7)
    return rng.integers(0, 256, size=shape, dtype=np.uint8)


def make_sequence(shape=KITTI_SHAPE, seed: int = 0, start: int = 0, count: int = 1, eye: int = 0) -> np.ndarray:
    """[count, rows, cols] uint8, frames start .. start+count-1 of the sequence `seed`."""
    return np.stack([make_frame(shape, seed, start + i, eye) for i in range(count)])


# ---- BASELINE.json configs[4]: the KITTI-00-shaped stereo sequence ------------------------------

KSEQ_PAIRS = 4541   # frames of KITTI odometry sequence 00 (the reference's Examples/Stereo/stereo_kitti.cc loop)
KSEQ_SEED = 4541
KSEQ_K = (718.856, 718.856, 607.1928, 185.2157)  # Examples/Stereo/KITTI00-02.yaml: fx, fy, cx, cy (no distortion)
KSEQ_BF = 386.1448                               # Camera.bf


def kseq_pair(f: int):
    """(left, right) of pair f of the synthetic Kseq sequence: a pure function of f."""
    return make_frame(KITTI_SHAPE, KSEQ_SEED, f, 0), make_frame(KITTI_SHAPE, KSEQ_SEED, f, 1)


def kseq_fill(frames, left: np.ndarray, right: np.ndarray, threads: int = 0) -> None:
    """Render pairs `frames` into left[i], right[i] (e.g. page-locked arrays), on a thread pool: the noise
    generator and the array arithmetic release the GIL."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    frames = list(frames)
    _scene(KITTI_SHAPE[0], KITTI_SHAPE[1], KSEQ_SEED, 300)  # build the shared scene once, before the threads

    def one(i):
        left[i] = make_frame(KITTI_SHAPE, KSEQ_SEED, frames[i], 0)
        right[i] = make_frame(KITTI_SHAPE, KSEQ_SEED, frames[i], 1)
    with ThreadPoolExecutor(threads or max(1, len(os.sched_getaffinity(0)))) as ex:
        list(ex.map(one, range(len(frames))))


# ---- inputs for the step before the extractor (colour frames, stereo rectification maps) -------

EUROC_SHAPE = (480, 752)
# (K, D, R, P) of the left and right camera, Examples/Stereo/EuRoC.yaml (LEFT.* / RIGHT.*)
EUROC = (
    ([458.654, 0.0, 367.215, 0.0, 457.296, 248.375, 0.0, 0.0, 1.0],
     [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05, 0.0],
     [0.999966347530033, -0.001422739138722922, 0.008079580483432283, 0.001365741834644127, 0.9999741760894847,
      0.007055629199258132, -0.008089410156878961, -0.007044357138835809, 0.9999424675829176],
     [435.2046959714599, 0, 367.4517211914062, 0, 0, 435.2046959714599, 252.2008514404297, 0, 0, 0, 1, 0]),
    ([457.587, 0.0, 379.999, 0.0, 456.134, 255.238, 0.0, 0.0, 1],
     [-0.28368365, 0.07451284, -0.00010473, -3.555907e-05, 0.0],
     [0.9999633526194376, -0.003625811871560086, 0.007755443660172947, 0.003680398547259526, 0.9999684752771629,
      -0.007035845251224894, -0.007729688520722713, 0.007064130529506649, 0.999945173484644],
     [435.2046959714599, 0, 367.4517211914062, -47.90639384423901, 0, 435.2046959714599, 252.2008514404297, 0,
      0, 0, 1, 0]),
)


def rectify_maps(K, D, R, P, size):
    """float32 (map_x, map_y) of shape (rows, cols): where each rectified pixel samples the raw
    image -- the quantity cv::initUndistortRectifyMap(K, D, R, P, size, CV_32F, M1, M2) produces
    (Examples/Stereo/stereo_euroc.cc:107-108).  The maps are INPUT DATA of the remap step; this is a
    plain float64 evaluation of the pinhole + radial/tangential model, not a bit-level restatement."""
    cols, rows = size
    K = np.asarray(K, np.float64).reshape(3, 3)
    R = np.asarray(R, np.float64).reshape(3, 3)
    P = np.asarray(P, np.float64).reshape(3, 4)[:, :3]
    k1, k2, p1, p2, k3 = (list(D) + [0.0] * 5)[:5]
    ir = np.linalg.inv(P @ R)
    u, v = np.meshgrid(np.arange(cols, dtype=np.float64), np.arange(rows, dtype=np.float64))
    X = ir[0, 0] * u + ir[0, 1] * v + ir[0, 2]
    Y = ir[1, 0] * u + ir[1, 1] * v + ir[1, 2]
    Wd = ir[2, 0] * u + ir[2, 1] * v + ir[2, 2]
    x, y = X / Wd, Y / Wd
    r2 = x * x + y * y
    kr = 1 + ((k3 * r2 + k2) * r2 + k1) * r2
    xd = x * kr + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
    yd = y * kr + p1 * (r2 + 2 * y * y) + 2 * p2 * x * y
    return ((K[0, 0] * xd + K[0, 2]).astype(np.float32), (K[1, 1] * yd + K[1, 2]).astype(np.float32))


def make_color_frame(shape=TUM_SHAPE, seed: int = 0, frame: int = 0, eye: int = 0, channels: int = 3) -> np.ndarray:
    """[rows, cols, channels] uint8: the gray synthetic frame with a smooth per-channel tint and
    independent per-channel noise (so the three channels differ and the weights matter)."""
    g = make_frame(shape, seed, frame, eye).astype(np.float32)
    rng = _rng(seed, frame + 7, eye + 2)
    rows, cols = g.shape
    out = np.empty((rows, cols, channels), np.uint8)
    for c in range(channels):
        tint = _value_noise(rng, rows, cols, 64, 30.0)
        out[..., c] = np.clip(np.rint(g * (0.8 + 0.1 * c) + tint + rng.uniform(-3, 3, g.shape)), 0, 255).astype(np.uint8)
    return out
