"""Soak parity: many distinct full-size frames of the BASELINE shapes against the CPU oracle, counted.
north_star's bar is a rate: keypoint records (pixel coordinates, octave, size, angle, response) and
descriptors bit-exact, any descriptor bit difference reported as a rate with target 0.  This test holds
the rate at exactly 0 over 96 KITTI-shape and 96 TUM-shape frames (mono and stereo eyes, three seeds),
i.e. ~390 k keypoints and 100 M descriptor bits."""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("shape,nf", [(synth.KITTI_SHAPE, 2000), (synth.TUM_SHAPE, 1000), (synth.TUM_SHAPE, 2000)],
                         ids=["K2k", "T1k", "T2k"])
def test_mismatch_rate_is_zero(shape, nf, oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    frames = [synth.make_frame(shape, seed=s, frame=f, eye=e) for s in (101, 102, 103) for f in range(16) for e in (0, 1)]
    batch = np.stack(frames)
    ext = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=len(frames), want_pyramid=False)
    res = ext.extract_batch(batch)
    o = oracle_built.Oracle(nf, 1.2, 8, 20, 7)
    n_kp = bad_kp = bad_angle = bad_bits = bad_frames = 0
    for img, (k, d) in zip(frames, res):
        ok, od = o.extract(img)
        n_kp += len(ok)
        if len(k) != len(ok):
            bad_frames += 1
            continue
        same = k.tobytes() == ok.tobytes() and d.tobytes() == od.tobytes()
        if not same:
            bad_frames += 1
            for name in ("x", "y", "size", "response", "octave"):
                bad_kp += int((k[name] != ok[name]).sum())
            bad_angle += int((np.abs(k["angle"] - ok["angle"]) > 1e-3).sum())
            bad_bits += int(np.unpackbits(d ^ od).sum())
    assert n_kp > 90 * nf
    assert (bad_frames, bad_kp, bad_angle, bad_bits) == (0, 0, 0, 0), \
        f"{bad_frames} frames differ: {bad_kp} keypoint fields, {bad_angle} angles > 1e-3 deg, {bad_bits} descriptor bits of {n_kp * 256}"


@pytest.mark.parametrize("shape,nf", [(synth.KITTI_SHAPE, 2000), (synth.TUM_SHAPE, 1000), (synth.TUM_SHAPE, 2000)],
                         ids=["K2k", "T1k", "T2k"])
def test_single_frame_shape_equals_the_batched_shape(shape, nf):
    """The same 96 frames, one per call (single-launch level chain, 1024-thread quadtree with the candidates in
    registers, blur on a graph branch, one-copy download) and as one batch (per-level kernels, 256-thread quadtree):
    byte-identical records and descriptors.  The batched results are the ones held against the oracle above."""
    from orb_slam2_test_b200 import ORBextractor
    frames = [synth.make_frame(shape, seed=s, frame=f, eye=e) for s in (101, 102, 103) for f in range(16) for e in (0, 1)]
    ext = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=len(frames), want_pyramid=False)
    res = ext.extract_batch(np.stack(frames))
    one = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
    differ = 0
    for img, (k, d) in zip(frames, res):
        k1, d1 = one(img)
        assert one.launches_per_call() == 5
        differ += not (k1.tobytes() == k.tobytes() and d1.tobytes() == d.tobytes())
    assert differ == 0, f"{differ} of {len(frames)} frames differ between the two launch shapes"
