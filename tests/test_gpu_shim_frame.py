"""The reference's OWN, UNMODIFIED src/Frame.cc over the drop-in ORBextractor (north_star: "Frame.cc ... link it as a
drop-in").

tests/shim/libshimframe.so = /root/reference/src/Frame.cc compiled where it lies with shim/ORBextractor.h in place of
the reference's include/ORBextractor.h, linked with the drop-in class (shim/ORBextractor.cc) and liborb_b200.so
(tests/shim/Makefile).  oracle/_ref/libreframe.so = the same Frame.cc over the reference's own ORBextractor.cc.  Both
run the reference's stereo Frame constructor (src/Frame.cc:79-151): two ExtractORB threads (:110-113), which call
operator() and read mvImagePyramid of the extractor they are given, UndistortKeyPoints, ComputeStereoMatches (:619-834,
on the HOST here, from the pyramids the drop-in brought back), ComputeImageBounds, AssignFeaturesToGrid.

Bar: mvKeys, mDescriptors, mvKeysRight, mDescriptorsRight, mvKeysUn, mvuRight, mvDepth, mGrid, bounds identical.
Neither library can be rebuilt on the GPU box (no /root/reference there): the prebuilt files travel with the snapshot.
"""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu

SHIMLIB = Path(__file__).parent / "shim" / "libshimframe.so"
CASES = {
    "kitti": (synth.KITTI_SHAPE, 301, (718.856, 718.856, 607.1928, 185.2157), (0.0, 0.0, 0.0, 0.0), 386.1448, 2000),
    "tum1_distorted": (synth.TUM_SHAPE, 302, (517.306408, 516.469215, 318.643040, 255.313989),
                       (0.262383, -0.953104, -0.005358, 0.002628, 1.163314), 40.0, 1000),
    "small_4coeff": ((240, 416), 303, (320.0, 320.0, 208.0, 100.0), (-0.28, 0.07, 0.0002, 0.00002), 170.0, 600),
}


@pytest.fixture(scope="module")
def libs(oracle_built):
    if not SHIMLIB.exists():
        pytest.fail(f"{SHIMLIB} is missing: build it with `make -C tests/shim libshimframe.so` where /root/reference exists")
    if not oracle_built.RefFrame.available():
        pytest.fail("oracle/_ref/libreframe.so is missing")
    return oracle_built.RefFrame, oracle_built.RefFrame.bind(C.CDLL(str(SHIMLIB)))


@pytest.mark.parametrize("name", list(CASES))
def test_reference_frame_constructor_over_the_drop_in_extractor(libs, name):
    RefFrame, shim = libs
    shape, seed, K, dist, bf, nf = CASES[name]
    for frame in (0, 3):
        left = synth.make_frame(shape, seed=seed, frame=frame, eye=0)
        right = synth.make_frame(shape, seed=seed, frame=frame, eye=1)
        mb = bf / K[0]
        want = RefFrame.stereo(left, right, K, dist, bf, mb, nfeatures=nf, canonical=True)
        got = RefFrame.stereo(left, right, K, dist, bf, mb, nfeatures=nf, lib=shim)
        assert len(want["kps_l"]) >= nf
        for k in ("kps_l", "desc_l", "kps_r", "desc_r", "un_xy", "u_right", "depth", "cell_start", "cell_idx", "bounds"):
            assert got[k].shape == want[k].shape, f"{name} frame {frame}: {k} shape {got[k].shape} vs {want[k].shape}"
            assert got[k].tobytes() == want[k].tobytes(), f"{name} frame {frame}: {k} differs"
        assert int((want["depth"] > 0).sum()) > 50
