"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded frames.

Bar (BASELINE.json north_star): keypoint coordinates, octave, size, response and the 32-byte
descriptors bit-exact, in the reference's output order; angle within 1e-3 degrees (asserted
bit-exact here, which is stricter).
"""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu

CONFIGS = {
    "T1k": (synth.TUM_SHAPE, 1000),    # TUM1.yaml
    "T2k": (synth.TUM_SHAPE, 2000),    # mono-init extractor, 2 x nFeatures
    "K2k": (synth.KITTI_SHAPE, 2000),  # KITTI00-02.yaml
}


def _pair(nf, max_batch=1):
    from oracle.pyoracle import Oracle
    from orb_slam2_test_b200 import ORBextractor
    return ORBextractor(nf, 1.2, 8, 20, 7, max_batch=max_batch), Oracle(nf, 1.2, 8, 20, 7)


def _assert_same(got, want, tag=""):
    gk, gd = got
    wk, wd = want
    assert len(gk) == len(wk), f"{tag}: {len(gk)} keypoints, oracle {len(wk)}"
    for f in ("x", "y", "octave", "size", "response", "class_id"):
        bad = np.nonzero(gk[f] != wk[f])[0]
        assert bad.size == 0, f"{tag}: field {f} differs at {bad[:5]} ({bad.size} total)"
    dang = np.abs(gk["angle"].astype(np.float64) - wk["angle"].astype(np.float64))
    assert dang.max(initial=0) <= 1e-3, f"{tag}: angle off by {dang.max()}"
    assert np.array_equal(gk["angle"], wk["angle"]), f"{tag}: angle not bit-exact"
    nbits = int(np.unpackbits(gd ^ wd).sum())
    assert nbits == 0, f"{tag}: {nbits} descriptor bits differ in {(gd != wd).any(axis=1).sum()} keypoints"


@pytest.mark.parametrize("cfg", list(CONFIGS))
def test_stages_and_end_to_end(cfg, oracle_built):
    shape, nf = CONFIGS[cfg]
    ext, ora = _pair(nf)
    for frame in range(2):
        img = synth.make_frame(shape, seed=3, frame=frame)
        got = ext(img)
        want = ora.extract(img)
        for l in range(8):
            assert ext.level_size(l) == ora.level_size(l)
            assert np.array_equal(ext.pyramid_level(0, l, border=19), ora.pyramid(l, 19)), f"pyramid L{l}"
            xs, ys, sc = ext.debug_candidates(0, l)
            c = ora.candidates(l)
            a = sorted(zip(xs.tolist(), ys.tolist(), sc.tolist()))
            b = sorted(zip(c["x"].tolist(), c["y"].tolist(), c["score"].tolist()))
            assert a == b, f"FAST candidates L{l}: {len(a)} vs {len(b)}"
            blur = ora.blurred(l)
            if blur is not None:
                assert np.array_equal(ext.debug_blurred(0, l), blur), f"blur L{l}"
        _assert_same(got, want, f"{cfg} frame {frame}")


def test_batch_equals_single(oracle_built):
    shape, nf = CONFIGS["K2k"]
    ext, ora = _pair(nf, max_batch=4)
    imgs = synth.make_sequence(shape, seed=5, start=0, count=4)
    res = ext.extract_batch(imgs)
    for f in range(4):
        _assert_same(res[f], ora.extract(imgs[f]), f"batch frame {f}")


def test_noise_frame_many_candidates(oracle_built):
    shape, nf = CONFIGS["T1k"]
    ext, ora = _pair(nf)
    img = synth.make_noise_frame(shape, seed=1)
    _assert_same(ext(img), ora.extract(img), "noise")


def test_flat_image_gives_nothing(oracle_built):
    ext, ora = _pair(1000)
    img = np.full(synth.TUM_SHAPE, 77, np.uint8)
    k, d = ext(img)
    wk, wd = ora.extract(img)
    assert len(k) == 0 and len(wk) == 0 and d.shape == (0, 32)


@pytest.mark.parametrize("shape", [synth.KITTI_SHAPE, synth.TUM_SHAPE, (251, 397), (263, 1001), (1080, 1920)])
def test_pyramid_planes_in_both_launch_shapes(shape, oracle_built):
    """K1 has two launch shapes: one launch for the whole level chain when the call is a frame or two
    (k_pyr_chain), one launch per level otherwise.  Both must give the oracle's planes, frame included."""
    from oracle.pyoracle import Oracle
    from orb_slam2_test_b200 import ORBextractor
    nb = 8
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=nb)
    ora = Oracle(1000, 1.2, 8, 20, 7)
    imgs = np.stack([synth.make_noise_frame(shape, seed=11 + f) if f % 2 else synth.make_frame(shape, seed=11, frame=f)
                     for f in range(nb)])
    want = {}
    for f in (0, nb - 1):
        ora.extract(imgs[f])
        want[f] = [ora.pyramid(l, 19) for l in range(8)]
    ext.extract_batch(imgs, want_pyramid=True)  # per-level kernels
    for f in (0, nb - 1):
        for l in range(8):
            assert np.array_equal(ext.pyramid_level(f, l, border=19), want[f][l]), f"per-level kernels, frame {f} L{l}"
    for f in (0, nb - 1):            # the chain kernel
        ext(imgs[f])
        for l in range(8):
            assert np.array_equal(ext.pyramid_level(0, l, border=19), want[f][l]), f"chain kernel, frame {f} L{l}"


@pytest.mark.parametrize("shape,scale,nlevels", [((263, 401), 1.1, 12), ((300, 400), 1.5, 4), ((376, 1241), 2.0, 3),
                                                 ((211, 333), 1.3, 5), ((480, 640), 1.05, 8)])
def test_batch_shape_pyramid_with_other_scale_factors(shape, scale, nlevels, oracle_built):
    """The per-level pyramid kernels of the batch shape (k_pyr_stream: source-row driven, staged pitch 384 B or, from a
    scale factor of about 1.5, 640 B) on other scale factors and level counts: every plane, frame included, and the
    results of the whole path against the oracle."""
    from oracle.pyoracle import Oracle
    from orb_slam2_test_b200 import ORBextractor
    nb = 48
    ext = ORBextractor(500, scale, nlevels, 20, 7, max_batch=nb)
    ora = Oracle(500, scale, nlevels, 20, 7)
    imgs = np.stack([synth.make_noise_frame(shape, seed=21 + f) if f % 3 == 1 else synth.make_frame(shape, seed=21, frame=f)
                     for f in range(nb)])
    res = ext.extract_batch(imgs, want_pyramid=True)
    assert ext.launches_per_call() != 5, "expected the batch launch shape"
    for f in (0, 1, nb - 1):
        want = ora.extract(imgs[f])
        for l in range(nlevels):
            assert np.array_equal(ext.pyramid_level(f, l, border=19), ora.pyramid(l, 19)), f"scale {scale}, frame {f} L{l}"
        _assert_same(res[f], want, f"scale {scale} frame {f}")


@pytest.mark.parametrize("cfg,count", [("K2k", 2), ("T1k", 3)])
def test_chain_shape_with_several_frames(cfg, count, oracle_built):
    """The single-launch level chain also serves calls of two or three frames (grid.y > 1): every frame's planes and
    results against the oracle."""
    shape, nf = CONFIGS[cfg]
    ext, ora = _pair(nf, max_batch=count)
    imgs = synth.make_sequence(shape, seed=9, start=2, count=count)
    res = ext.extract_batch(imgs, want_pyramid=True)
    assert ext.launches_per_call() == 5, "expected the single-frame launch shape"
    for f in range(count):
        want = ora.extract(imgs[f])
        for l in range(8):
            assert np.array_equal(ext.pyramid_level(f, l, border=19), ora.pyramid(l, 19)), f"frame {f} L{l}"
        _assert_same(res[f], want, f"{cfg} frame {f} of {count}")
