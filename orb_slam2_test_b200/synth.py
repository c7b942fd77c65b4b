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
