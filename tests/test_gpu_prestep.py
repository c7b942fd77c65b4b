"""GPU parity for the step before the extractor: colour -> gray
(cv::cvtColor, src/Tracking.cc:176-201) and stereo rectification (cv::remap,
Examples/Stereo/stereo_euroc.cc:136-137) fused in front of the pyramid.

Oracle: oracle/cvprim_pre.c, pinned bit-exact against cv2 4.13.0 by tests/test_oracle_primitives.py.
Bar: the gray / rectified image (= pyramid level 0) byte-exact, and keypoints + descriptors equal
to the oracle extractor run on the oracle's gray / rectified image.
"""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu


def _same(got, want, tag):
    gk, gd = got
    wk, wd = want
    assert gk.tobytes() == wk.tobytes(), f"{tag}: keypoints differ ({len(gk)} vs {len(wk)})"
    assert gd.tobytes() == wd.tobytes(), f"{tag}: descriptors differ"


@pytest.mark.parametrize("fmt", ["bgr", "rgb", "bgra", "rgba"])
def test_colour_frames(fmt, oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    ch = 3 if len(fmt) == 3 else 4
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=2)
    ext.set_input_format(fmt)
    ora = oracle_built.Oracle(1000, 1.2, 8, 20, 7)
    imgs = np.stack([synth.make_color_frame(synth.TUM_SHAPE, seed=21, frame=f, channels=ch) for f in range(2)])
    res = ext.extract_batch(imgs, want_pyramid=True)
    for f in range(2):
        gray = oracle_built.cvt_gray(imgs[f], fmt.startswith("rgb"))
        assert np.array_equal(ext.pyramid_level(f, 0), gray), f"{fmt} frame {f}: gray image"
        _same(res[f], ora.extract(gray), f"{fmt} frame {f}")
    # single frame through operator(), odd width, non-contiguous view (ROI of a wider frame)
    wide = synth.make_color_frame((240, 331), seed=22, channels=ch)
    roi = wide[5:230, 7:320]
    got = ext(roi)
    _same(got, ora.extract(oracle_built.cvt_gray(np.ascontiguousarray(roi), fmt.startswith("rgb"))), f"{fmt} roi")


@pytest.mark.parametrize("eye", [0, 1])
def test_euroc_rectification(eye, oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    mx, my = synth.rectify_maps(*synth.EUROC[eye], size=(752, 480))
    ext = ORBextractor(1200, 1.2, 8, 20, 7, max_batch=3)        # EuRoC.yaml: 1200 features
    ext.set_rectify_map(mx, my)
    ora = oracle_built.Oracle(1200, 1.2, 8, 20, 7)
    imgs = synth.make_sequence(synth.EUROC_SHAPE, seed=40, start=0, count=3, eye=eye)
    res = ext.extract_batch(imgs, want_pyramid=True)
    for f in range(3):
        rect = oracle_built.remap_linear(imgs[f], mx, my)
        assert np.array_equal(ext.pyramid_level(f, 0), rect), f"eye {eye} frame {f}: rectified image"
        _same(res[f], ora.extract(rect), f"eye {eye} frame {f}")
    # removing the map restores the plain path
    ext.set_rectify_map(None, None)
    _same(ext.extract_batch(imgs[:1])[0], ora.extract(imgs[0]), "map removed")


def test_rectify_changes_geometry_and_handles_outside_taps(oracle_built):
    """Map smaller than the source, with samples falling off every edge (constant border 0)."""
    from orb_slam2_test_b200 import ORBextractor
    rng = np.random.default_rng(3)
    src = synth.make_frame((300, 420), seed=23)
    h, w = 260, 381
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    mx = (xx * 1.15 - 12 + 2 * np.sin(yy / 17)).astype(np.float32)
    my = (yy * 1.2 - 9 + 2 * np.cos(xx / 23)).astype(np.float32)
    mx[10:14, 20:40] = np.nan
    my[30, 50:60] = 1e20
    mx[40, 60:70] = -1e20
    mx[50:60, 100:140] += rng.uniform(-0.5, 0.5, (10, 40)).astype(np.float32)
    ext = ORBextractor(800, 1.2, 8, 20, 7)
    ext.set_rectify_map(mx, my)
    got = ext(src)
    safe_x = np.where(np.isfinite(mx) & (np.abs(mx) < 1e6), mx, np.float32(-1000))
    safe_y = np.where(np.isfinite(my) & (np.abs(my) < 1e6), my, np.float32(-1000))
    rect = oracle_built.remap_linear(src, safe_x, safe_y)
    assert ext.level_size(0) == (w, h)
    assert np.array_equal(ext.pyramid_level(0, 0), rect)
    _same(got, oracle_built.Oracle(800, 1.2, 8, 20, 7).extract(rect), "odd map")


def test_colour_and_rectification_together(oracle_built):
    """Raw colour stereo frame: remap per channel, then gray (the reference's order)."""
    from orb_slam2_test_b200 import ORBextractor
    mx, my = synth.rectify_maps(*synth.EUROC[0], size=(752, 480))
    img = synth.make_color_frame(synth.EUROC_SHAPE, seed=41, channels=3)
    ext = ORBextractor(1200, 1.2, 8, 20, 7)
    ext.set_input_format("rgb")
    ext.set_rectify_map(mx, my)
    got = ext(img)
    want_img = oracle_built.cvt_gray(oracle_built.remap_linear(img, mx, my), True)
    assert np.array_equal(ext.pyramid_level(0, 0), want_img)
    _same(got, oracle_built.Oracle(1200, 1.2, 8, 20, 7).extract(want_img), "rgb + rectify")


def test_device_resident_colour_batch(oracle_built):
    import torch
    from orb_slam2_test_b200 import ORBextractor
    imgs = np.stack([synth.make_color_frame(synth.TUM_SHAPE, seed=24, frame=f) for f in range(4)])
    dev = torch.from_numpy(imgs).cuda()
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=4)
    ext.set_input_format("bgr")
    ext.extract_device(dev.data_ptr(), 4, 640, 480)
    res = ext.fetch()
    ora = oracle_built.Oracle(1000, 1.2, 8, 20, 7)
    for f in range(4):
        _same(res[f], ora.extract(oracle_built.cvt_gray(imgs[f], False)), f"device frame {f}")


def test_format_errors():
    from orb_slam2_test_b200 import ORBextractor
    from orb_slam2_test_b200.extractor import BorbError, _check
    ext = ORBextractor(500, 1.2, 8, 20, 7)
    with pytest.raises(BorbError):
        _check(ext._lib, ext._lib.borb_set_input_format(ext._h, 9))
    ext.set_input_format("bgr")
    with pytest.raises(TypeError):
        ext(synth.make_frame(synth.TUM_SHAPE, seed=1))      # gray frame handed to a BGR extractor
