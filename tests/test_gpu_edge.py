"""GPU edge cases the reference never tested: determinism, concurrent instances,
strided input, odd geometries and parameters, error paths.  All through the C ABI."""
import threading

import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu


def _same(a, b):
    return a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()


def test_determinism_100_runs_one_instance():
    from orb_slam2_test_b200 import ORBextractor
    ext = ORBextractor(2000, 1.2, 8, 20, 7, want_pyramid=False)
    img = synth.make_frame(synth.KITTI_SHAPE, seed=9, frame=0)
    first = ext(img)
    for _ in range(100):
        assert _same(ext(img), first)


def test_two_concurrent_instances_like_stereo(oracle_built):
    """src/Frame.cc:110-113: left and right extractors run on two threads."""
    from orb_slam2_test_b200 import ORBextractor
    imgs = [synth.make_frame(synth.KITTI_SHAPE, seed=9, frame=0, eye=e) for e in (0, 1)]
    exts = [ORBextractor(2000, 1.2, 8, 20, 7) for _ in (0, 1)]
    want = [oracle_built.Oracle(2000, 1.2, 8, 20, 7).extract(im) for im in imgs]
    out = [None, None]

    def run(i):
        for _ in range(25):
            out[i] = exts[i](imgs[i])

    th = [threading.Thread(target=run, args=(i,)) for i in (0, 1)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert _same(out[0], want[0]) and _same(out[1], want[1])


def test_non_contiguous_roi_input(oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    big = synth.make_frame((420, 700), seed=4, frame=0)
    roi = big[17:390, 33:677]  # step 700 > cols 644, as a cv::Mat ROI
    got = ORBextractor(1000, 1.2, 8, 20, 7)(roi)
    want = oracle_built.Oracle(1000, 1.2, 8, 20, 7).extract(np.ascontiguousarray(roi))
    assert _same(got, want)


PARAMS = [((480, 752), (1200, 1.2, 8, 20, 7)),      # EuRoC.yaml
          (synth.KITTI_SHAPE, (2000, 1.2, 8, 12, 7)),  # KITTI04-12.yaml
          ((211, 333), (700, 1.3, 5, 30, 5)),
          ((240, 320), (300, 1.1, 12, 20, 7)),
          ((300, 400), (500, 1.5, 4, 25, 10)),
          ((200, 260), (50, 2.0, 2, 20, 7)),
          ((240, 320), (1000, 1.2, 8, 7, 20)),        # ini < min: a second pass can add nothing
          ((360, 1100), (1500, 1.2, 6, 130, 100))]    # thresholds above 127


@pytest.mark.parametrize("shape,p", PARAMS)
def test_other_parameters_and_shapes(shape, p, oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    ext, ora = ORBextractor(*p), oracle_built.Oracle(*p)
    for img in (synth.make_frame(shape, seed=17, frame=0), synth.make_noise_frame(shape, seed=3)):
        got, want = ext(img), ora.extract(img)
        assert len(got[0]) == len(want[0])
        assert _same(got, want)


def test_geometry_change_between_calls(oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    ext, ora = ORBextractor(1000, 1.2, 8, 20, 7), oracle_built.Oracle(1000, 1.2, 8, 20, 7)
    for shape in (synth.TUM_SHAPE, (240, 320), synth.KITTI_SHAPE, synth.TUM_SHAPE):
        img = synth.make_frame(shape, seed=2, frame=1)
        assert _same(ext(img), ora.extract(img))


def test_half_flat_half_textured_and_low_contrast(oracle_built):
    """cells with nothing at 20 (retry at 7), cells with nothing at all, saturated regions"""
    from orb_slam2_test_b200 import ORBextractor
    img = synth.make_frame(synth.TUM_SHAPE, seed=5, frame=0).astype(np.int32)
    img[:, :200] = 128 + (img[:, :200] - 128) // 6   # low contrast: only the min threshold fires
    img[300:, :] = 40                                 # flat
    img[:60, 400:] = 255                              # saturated
    img = np.clip(img, 0, 255).astype(np.uint8)
    got = ORBextractor(1000, 1.2, 8, 20, 7)(img)
    want = oracle_built.Oracle(1000, 1.2, 8, 20, 7).extract(img)
    assert len(want[0]) > 100 and _same(got, want)


def test_errors_are_reported_not_hidden():
    from orb_slam2_test_b200 import BorbError, ORBextractor
    ext = ORBextractor(1000, 1.2, 8, 20, 7)
    with pytest.raises(BorbError) as e:   # level 7 smaller than one FAST cell
        ext(np.zeros((100, 120), np.uint8))
    assert e.value.code == -3
    with pytest.raises(BorbError) as e:   # aspect ratio < 0.5: the reference divides by zero (nIni == 0)
        ext(np.zeros((900, 300), np.uint8))
    assert e.value.code == -3
    with pytest.raises(TypeError):
        ext(np.zeros((240, 320), np.float32))
    assert ext(np.zeros((0, 0), np.uint8)) is None  # empty image: silent return, src/ORBextractor.cc:1333
    with pytest.raises(BorbError) as e:   # more frames than max_batch
        ext.extract_batch(np.zeros((2, 240, 320), np.uint8))
    assert e.value.code == -4
    img = synth.make_frame((240, 320), seed=1, frame=0)
    assert len(ext(img)[0]) > 0           # the handle is still usable after errors


def test_batch_tum_and_mixed_content(oracle_built):
    from orb_slam2_test_b200 import ORBextractor
    ext, ora = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=6), oracle_built.Oracle(1000, 1.2, 8, 20, 7)
    imgs = np.stack([synth.make_frame(synth.TUM_SHAPE, seed=8, frame=0), synth.make_noise_frame(synth.TUM_SHAPE, seed=8),
                     np.full(synth.TUM_SHAPE, 99, np.uint8), synth.make_frame(synth.TUM_SHAPE, seed=8, frame=3),
                     synth.make_frame(synth.TUM_SHAPE, seed=11, frame=0)])
    res = ext.extract_batch(imgs)
    for f in range(len(imgs)):
        assert _same(res[f], ora.extract(imgs[f])), f
    # async halves give the same
    ext.submit_batch(imgs)
    ext.wait()
    for f in range(len(imgs)):
        assert _same(ext.result(f), res[f])


def test_sequence_shard_runner_on_gpu(oracle_built):
    """BASELINE config 5 in miniature: a stereo sequence through sequence.run_shard with the real
    extractor (batched calls), every (frame, eye) digest equal to the oracle's."""
    from orb_slam2_test_b200 import ORBextractor, sequence
    n_frames, world = 7, 2
    ext = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=4, want_pyramid=False)
    ora = oracle_built.Oracle(2000, 1.2, 8, 20, 7)

    def load(f, e):
        return synth.make_frame(synth.KITTI_SHAPE, seed=41, frame=f, eye=e)

    allr = []
    for rank in range(world):  # the two shards one after the other on this one GPU
        allr += sequence.run_shard(sequence.shard_indices(n_frames, rank, world), load, ext.extract_batch, 4, True)
    allr = sequence.gather_to_rank0(allr, 0, 1)
    sequence.check_complete(allr, n_frames, True)
    for r in allr:
        k, d = ora.extract(load(r.frame, r.eye))
        assert sequence.FrameResult.of(r.frame, r.eye, k, d).digest == r.digest


@pytest.mark.parametrize("shape,nlevels", [((120, 91), 1), ((100, 77), 1), ((97, 152), 1), ((200, 121), 2)])
def test_wide_cells_use_the_wider_tma_boxes(shape, nlevels, oracle_built):
    """Narrow images give one or two FAST cells per row, up to 59 px wide (window 65 px): the
    window loads then need the 80- and 96-byte TMA boxes instead of the 64-byte one."""
    from orb_slam2_test_b200 import ORBextractor
    ext = ORBextractor(300, 1.2, nlevels, 20, 7)
    ora = oracle_built.Oracle(300, 1.2, nlevels, 20, 7)
    for seed in (1, 2):
        img = synth.make_frame(shape, seed=seed)
        assert _same(ext(img), ora.extract(img)), f"{shape} seed {seed}"
    img = synth.make_noise_frame(shape, seed=3)
    assert _same(ext(img), ora.extract(img)), f"{shape} noise"


@pytest.mark.parametrize("count", [1, 6])
def test_device_input_with_odd_row_step_and_unaligned_base(count, oracle_built):
    """borb_extract_device on a region of a larger device buffer: row step 701 (odd), base address 1 mod 4, frames
    421 rows apart.  count = 1 runs the single-launch level chain, count = 6 the per-level kernels; both read the
    input through aligned words at whatever alignment each row has."""
    import torch
    from orb_slam2_test_b200 import ORBextractor
    rows, cols, step, frows = 373, 641, 701, 421
    big = np.stack([synth.make_frame((frows, step), seed=21, frame=f) for f in range(count)])
    dev = torch.from_numpy(big).cuda()
    y0, x0 = 17, 32
    base = dev.data_ptr() + y0 * step + x0
    assert base % 4 == 1
    ext = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=count)
    ext.extract_device(base, count, cols, rows, step=step, frame_stride=frows * step)
    got = ext.fetch()
    ora = oracle_built.Oracle(1000, 1.2, 8, 20, 7)
    for f in range(count):
        want = ora.extract(np.ascontiguousarray(big[f, y0:y0 + rows, x0:x0 + cols]))
        assert _same(got[f], want), f"frame {f}"


def test_a_new_input_buffer_on_every_call(oracle_built):
    """Device input at a different address on every call: the graph cache cannot hit; after a few misses the
    calls are launched directly instead of being captured each time, and a buffer that comes back is captured
    again.  Results stay the oracle's throughout."""
    import torch
    from orb_slam2_test_b200 import ORBextractor
    shape = synth.TUM_SHAPE
    imgs = [synth.make_frame(shape, seed=31, frame=f) for f in range(3)]
    ora = oracle_built.Oracle(1000, 1.2, 8, 20, 7)
    want = [ora.extract(i) for i in imgs]
    ext = ORBextractor(1000, 1.2, 8, 20, 7)
    keep = []
    for k in range(24):
        dev = torch.from_numpy(imgs[k % 3]).cuda()
        keep.append(dev)  # keep them alive so that every address is new
        ext.extract_device(dev.data_ptr(), 1, shape[1], shape[0])
        assert _same(ext.fetch()[0], want[k % 3]), f"call {k}"
    for k in range(6):  # a stable buffer again
        ext.extract_device(keep[0].data_ptr(), 1, shape[1], shape[0])
        assert _same(ext.fetch()[0], want[0])


def test_alternating_frames_in_the_single_frame_shape(oracle_built):
    """The single-frame shape launches FAST, the quadtree and the describe stage programmatically behind their
    predecessors (griddepcontrol, borb_cta.cuh): a kernel that read its input before the predecessor had finished
    would see the PREVIOUS call's candidates / survivors, which a repeated image cannot show.  Three very different
    images in rotation (textured, noise, nearly flat), one and two frames per call, every result against the oracle's."""
    from orb_slam2_test_b200 import ORBextractor
    shape = synth.KITTI_SHAPE
    flat = np.full(shape, 90, np.uint8)
    flat[100:140, 300:360] = 160
    imgs = [synth.make_frame(shape, seed=21, frame=0), synth.make_noise_frame(shape, seed=4), flat,
            synth.make_frame(shape, seed=22, frame=5)]
    ora = oracle_built.Oracle(2000, 1.2, 8, 20, 7)
    want = [ora.extract(im) for im in imgs]
    assert len({len(w[0]) for w in want}) >= 3 # the counts differ, so a stale count shows
    one = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=2, want_pyramid=False)
    for it in range(120):
        i = (it * 7 + it // 5) % len(imgs)
        assert _same(one(imgs[i]), want[i]), f"call {it}, image {i}"
        if it % 4 == 3:
            j = (i + 1 + it % 3) % len(imgs)
            res = one.extract_batch(np.stack([imgs[i], imgs[j]]))
            assert _same(res[0], want[i]) and _same(res[1], want[j]), f"pair call {it}"


def test_very_wide_image_uses_the_64_bit_winner_keys(oracle_built):
    """More than 64 cell columns on level 0 (2200 px / 30): the quadtree's winner keys no longer fit 32 bits on that
    level (k_quadtree.cuh) while the smaller levels still take the 32-bit path; textured and noise frame, one frame per
    call and as a batch, against the oracle."""
    from orb_slam2_test_b200 import ORBextractor
    shape = (300, 2200)
    ora = oracle_built.Oracle(1500, 1.2, 8, 20, 7)
    imgs = [synth.make_frame(shape, seed=41, frame=0), synth.make_noise_frame(shape, seed=6), synth.make_frame(shape, seed=42, frame=3)]
    want = [ora.extract(im) for im in imgs]
    one = ORBextractor(1500, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
    for im, w in zip(imgs, want):
        assert _same(one(im), w)
    big = ORBextractor(1500, 1.2, 8, 20, 7, max_batch=3, want_pyramid=False)
    for r, w in zip(big.extract_batch(np.stack(imgs)), want):
        assert _same(r, w)
