"""CPU checks of the stereo oracle (oracle/stereo_oracle.cpp, a restatement of
Frame::ComputeStereoMatches, src/Frame.cc:619-834): the properties the algorithm guarantees.  The pin
against the reference's own compiled Frame.cc is tests/test_oracle_frame_vs_reference.py, the golden
vectors generated from it are checked by tests/test_golden_frame.py.
"""
import numpy as np

from orb_slam2_test_b200 import synth

BF, FX = 386.1448, 718.856
MB = BF / FX


def _extract(ora, img):
    k, d = ora.extract(img)
    pyr = [ora.pyramid(l, 0) for l in range(8)]
    return k, d, pyr


def _run(seed, eye_r=1, mb=MB):
    from oracle import pyoracle
    ora = pyoracle.Oracle(2000, 1.2, 8, 20, 7)
    kl, dl, pl = _extract(ora, synth.make_frame(synth.KITTI_SHAPE, seed=seed, eye=0))
    kr, dr, pr = _extract(ora, synth.make_frame(synth.KITTI_SHAPE, seed=seed, eye=eye_r))
    sc, inv = ora.table(0), ora.table(1)
    return kl, pyoracle.stereo_match(kl, dl, kr, dr, pl, pr, sc, inv, BF, mb)


def test_shifted_pair_recovers_the_shift(oracle_built):
    kl, (u, z, sad, kept) = _run(31)
    m = z > 0
    assert kept == int(m.sum()) and kept > 800
    disp = kl["x"][m] - u[m]
    assert abs(float(np.median(disp)) - 17.0) < 0.25      # synth: the right eye is shifted by 17 px
    assert np.all(disp > 0) and np.all(disp < np.float32(BF) / np.float32(MB))
    assert np.all(z[m] == np.float32(BF) / disp.astype(np.float32))
    assert np.all(u[~m] == -1) and np.all(z[~m] == -1)
    # the outlier cut: everything kept has SAD below 1.5*1.4*median of the matched SADs
    s = np.sort(sad[sad >= 0])
    th = np.float32(1.5) * np.float32(1.4) * np.float32(s[len(s) // 2])
    assert np.all(sad[m].astype(np.float32) < th)
    assert np.all(sad[(sad >= 0) & ~m].astype(np.float32) >= th)


def _half_noisy(img, seed):
    """A right image equal to the left in its top 35 % and lightly noised below: zero-disparity,
    zero-SAD matches on top, a positive median SAD from the rest (so the cut keeps the top ones)."""
    rng = np.random.default_rng(seed)
    out = img.astype(np.int16)
    r0 = int(img.shape[0] * 0.35)
    out[r0:] += rng.integers(-3, 4, size=out[r0:].shape, dtype=np.int16)
    return np.clip(out, 0, 255).astype(np.uint8)


def _mirror_tiles(img, tw=31):
    """Every 61-column tile mirrored about its centre column: for a keypoint on such an axis the
    SADs at shifts -1 and +1 are equal, the parabola offset is exactly 0 and the disparity of an
    identical right image is exactly 0."""
    out = img.copy()
    for x0 in range(0, img.shape[1] - (2 * tw - 1) + 1, 2 * tw - 1):
        half = img[:, x0:x0 + tw]
        out[:, x0:x0 + tw] = half
        out[:, x0 + tw - 1:x0 + 2 * tw - 1] = half[:, ::-1]
    return out


def test_identical_images_are_all_cut(oracle_built):
    """All SADs 0 -> median 0 -> threshold 0 -> `dist < thDist` never holds (src/Frame.cc:792-803)."""
    _, (u, z, sad, kept) = _run(8, eye_r=0)
    assert kept == 0 and np.all(z == -1) and np.all(u == -1)
    assert (sad == 0).sum() > 500 and np.all(sad <= 0)


def test_zero_disparity_takes_the_epsilon_branch(oracle_built):
    from oracle import pyoracle
    ora = pyoracle.Oracle(2000, 1.2, 8, 20, 7)
    img = _mirror_tiles(synth.make_frame(synth.KITTI_SHAPE, seed=9))
    kl, dl, pl = _extract(ora, img)
    kr, dr, pr = _extract(ora, _half_noisy(img, 5))
    u, z, sad, kept = pyoracle.stereo_match(kl, dl, kr, dr, pl, pr, ora.table(0), ora.table(1), BF, MB)
    eps = z == np.float32(BF) / np.float32(0.01)
    assert eps.sum() >= 10
    assert np.all(u[eps] == (kl["x"][eps].astype(np.float64) - 0.01).astype(np.float32))


def test_search_range_excludes_the_true_disparity(oracle_built):
    _, (u, z, sad, kept) = _run(31, mb=MB * 64)      # maxD = fx / 64 = 11.2 px < 17 px
    _, (_, _, _, kept_full) = _run(31)
    assert kept < kept_full / 4
