"""GPU parity for the pipelined sequence feeder (include/borb.h borb_stream_*).

A stream re-points the extractor's result pointers at per-slot slabs and runs uploads, kernels and
downloads on three CUDA streams; none of that may change a byte.  Bar: every item that comes out of
a stream -- keypoints, descriptors, mvuRight / mvDepth, mvKeysUn + grid, pyramid levels -- equals the
CPU oracle (oracle/pyoracle.py, pinned to the reference's own ORBextractor.cc / Frame.cc) bit for bit,
in push order, for full and partial batches, staged and in-place uploads, the caller-driven loop and
the library's own feeder thread.  Reference loop: Examples/Stereo/stereo_kitti.cc:66-108,
src/Frame.cc:79-151.
"""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu

BF, FX = 386.1448, 718.856
MB = BF / FX
KITTI_K = (718.856, 718.856, 607.1928, 185.2157)


def _same(kg, dg, ko, do, tag):
    assert len(kg) == len(ko), f"{tag}: {len(kg)} vs {len(ko)} keypoints"
    assert kg.tobytes() == ko.tobytes(), f"{tag}: keypoint records differ"
    assert np.array_equal(dg, do), f"{tag}: descriptors differ"


def _oracle_stereo(ko_l, do_l, ko_r, do_r, pl, pr, scale, inv):
    from oracle import pyoracle
    return pyoracle.stereo_match(ko_l, do_l, ko_r, do_r, pl, pr, scale, inv, BF, MB)


def test_mono_stream_equals_the_oracle_in_order(oracle_built):
    """11 TUM frames through a 4-item ring of 3 slots: full batches, a partial last batch, slots re-used."""
    from orb_slam2_test_b200 import OrbStream
    O = oracle_built.Oracle(1000, 1.2, 8, 20, 7)
    frames = synth.make_sequence(synth.TUM_SHAPE, seed=91, start=0, count=11)
    st = OrbStream(1000, 1.2, 8, 20, 7, 640, 480, batch=4, slots=3)
    got, i = {}, 0
    while i < len(frames) or st.in_flight():
        while i < len(frames) and st.slots_free() > 0:
            m = min(4, len(frames) - i)
            st.push(frames[i:i + m], tag=i)
            i += m
        tag, m = st.pop()
        for j in range(m):
            got[tag + j] = st.result(j)
        st.release()
    assert sorted(got) == list(range(11))
    for f in range(11):
        ko, do = O.extract(frames[f])
        _same(*got[f], ko, do, f"frame {f}")
    assert st.launches_per_push() >= 5


def test_ring_bookkeeping_errors():
    from orb_slam2_test_b200 import BorbError, OrbStream
    st = OrbStream(500, 1.2, 8, 20, 7, 640, 480, batch=2, slots=2)
    with pytest.raises(BorbError):
        st.pop()                                  # nothing in flight
    img = synth.make_sequence(synth.TUM_SHAPE, seed=92, start=0, count=2)
    st.push(img, tag=7)
    st.push(img, tag=8)
    assert st.slots_free() == 0 and st.in_flight() == 2
    with pytest.raises(BorbError):
        st.push(img, tag=9)                       # ring full
    with pytest.raises(BorbError):
        st.input_buffer(0)
    assert st.pop() == (7, 2)
    assert st.slots_free() == 0                   # the popped slot is held for reading
    a = st.result(0)
    st.release()
    assert st.slots_free() == 1
    st.push(img, tag=9)
    assert st.pop() == (8, 2)
    b = st.result(0)
    assert st.pop() == (9, 2)
    _same(*a, *b, "same frame, different slots")
    with pytest.raises(BorbError):
        st.push(np.zeros((3, 480, 640), np.uint8))  # more than a batch
    with pytest.raises(BorbError):
        st.result(5)


def test_stereo_stream_equals_the_oracle(oracle_built):
    """KITTI pairs: both eyes, the stereo association, undistortion + grid and the pyramids through one stream."""
    from oracle import pyoracle
    from orb_slam2_test_b200 import OrbStream
    n = 5
    li = synth.make_sequence(synth.KITTI_SHAPE, seed=93, start=0, count=n, eye=0)
    ri = synth.make_sequence(synth.KITTI_SHAPE, seed=93, start=0, count=n, eye=1)
    st = OrbStream(2000, 1.2, 8, 20, 7, 1241, 376, batch=2, slots=2, stereo=True, want_pyramid=True)
    st.set_stereo(BF, MB)
    st.set_camera(KITTI_K, (0.0, 0.0, 0.0, 0.0))
    O = oracle_built.Oracle(2000, 1.2, 8, 20, 7)
    from orb_slam2_test_b200 import ORBextractor
    tab = ORBextractor(2000, 1.2, 8, 20, 7)
    scale, inv = tab.GetScaleFactors(), tab.GetInverseScaleFactors()
    i, seen, matched = 0, 0, 0
    while i < n or st.in_flight():
        while i < n and st.slots_free() > 0:
            m = min(2, n - i)
            st.push(li[i:i + m], ri[i:i + m], tag=i)
            i += m
        tag, m = st.pop()
        for j in range(m):
            f = tag + j
            kl, dl = st.result(j, 0)
            kr, dr = st.result(j, 1)
            ko, do = O.extract(li[f])
            pl_o = [O.pyramid(l, border=0) for l in range(8)]
            _same(kl, dl, ko, do, f"pair {f} left")
            ko_r, do_r = O.extract(ri[f])
            pr_o = [O.pyramid(l, border=0) for l in range(8)]
            _same(kr, dr, ko_r, do_r, f"pair {f} right")
            for l in range(8):
                assert np.array_equal(st.pyramid_level(j, l, 0), pl_o[l]), f"pair {f} left pyramid level {l}"
                assert np.array_equal(st.pyramid_level(j, l, 1), pr_o[l]), f"pair {f} right pyramid level {l}"
            u, z, kept = st.stereo_result(j)
            wu, wz, _, wkept = _oracle_stereo(ko, do, ko_r, do_r, pl_o, pr_o, scale, inv)
            assert u.view(np.uint32).tobytes() == wu.view(np.uint32).tobytes(), f"pair {f}: mvuRight differs"
            assert z.view(np.uint32).tobytes() == wz.view(np.uint32).tobytes(), f"pair {f}: mvDepth differs"
            assert kept == wkept
            un, start, idx, b = st.frame_grid(j)
            wun, wstart, widx, wb = pyoracle.frame_post(kl, 1241, 376, KITTI_K, (0.0, 0.0, 0.0, 0.0))
            assert un.tobytes() == wun.tobytes() and np.array_equal(start, wstart) and np.array_equal(idx, widx)
            assert b.tobytes() == np.asarray(wb, np.float32).tobytes()
            matched += kept
            seen += 1
        st.release()
    assert seen == n and matched > 100 * n


def test_in_place_upload_equals_staged_upload_and_the_feeder_thread():
    """Page-locked caller memory is uploaded from where it lies; pageable memory is staged through the slot; the
    library's own feeder thread produces into the slot buffers.  All three give the same bytes."""
    from orb_slam2_test_b200 import OrbStream, sequence
    from orb_slam2_test_b200.extractor import pinned_empty
    n, B = 13, 4
    li = synth.make_sequence(synth.TUM_SHAPE, seed=94, start=0, count=n, eye=0)
    ri = synth.make_sequence(synth.TUM_SHAPE, seed=94, start=0, count=n, eye=1)

    def stream():
        s = OrbStream(1000, 1.2, 8, 20, 7, 640, 480, batch=B, slots=3, stereo=True)
        s.set_stereo(BF, MB)
        return s

    pl, pr = pinned_empty(li.shape), pinned_empty(ri.shape)
    pl[:], pr[:] = li, ri
    a = sequence.run_stream_shard(stream(), range(n), pl, pr)                       # in place
    b = sequence.run_stream_shard(stream(), range(n), li, ri)                       # staged by the library
    def fill(f, l, r):
        l[:] = li[f]
        r[:] = ri[f]
    c = sequence.run_stream_shard(stream(), range(n), None, None, fill=fill)        # producer writes the slot buffers
    # the feeder thread
    out = {}
    s = stream()

    def produce(index, left, right):
        lo = index * B
        m = max(0, min(B, n - lo))
        left[:m], right[:m] = li[lo:lo + m], ri[lo:lo + m]
        return m

    def consume(stm, tag, m):
        for j in range(m):
            out[tag * B + j] = sequence.item_digest(stm, j)
    s.run(produce, consume)
    assert [r.frame for r in a] == list(range(n))
    da = [r.digest for r in a]
    assert da == [r.digest for r in b] == [r.digest for r in c] == [out[f] for f in range(n)]
    assert len(set(da)) == n
    assert all(r.n_matched > 50 for r in a)
    assert sequence.sequence_digest(a, n) == sequence.sequence_digest(c, n)


def test_sharded_sequence_digest_equals_the_single_stream_digest():
    """BASELINE configs[4] in miniature: 12 KITTI pairs through one stream, and as two shards (pair f -> shard
    f mod 2) through two streams; re-ordered by frame index the two runs give one digest."""
    from orb_slam2_test_b200 import OrbStream, sequence
    n = 12
    li = synth.make_sequence(synth.KITTI_SHAPE, seed=95, start=0, count=n, eye=0)
    ri = synth.make_sequence(synth.KITTI_SHAPE, seed=95, start=0, count=n, eye=1)

    def run(frames, batch):
        s = OrbStream(2000, 1.2, 8, 20, 7, 1241, 376, batch=batch, slots=2, stereo=True)
        s.set_stereo(BF, MB)
        idx = list(frames)
        return sequence.run_stream_shard(s, idx, np.ascontiguousarray(li[idx]), np.ascontiguousarray(ri[idx]))

    whole = run(range(n), 5)
    parts = [run(sequence.shard_indices(n, r, 2), 4) for r in range(2)]
    merged = sorted(parts[0] + parts[1], key=lambda r: r.frame)
    assert sequence.sequence_digest(merged, n) == sequence.sequence_digest(whole, n)
    with pytest.raises(RuntimeError):
        sequence.sequence_digest(merged[:-1], n)


def test_advice_r1_state_guards():
    """a submit with another geometry while a call is pending must not tear the pending call's
    buffers down; a stereo result must not outlive the extract call it belongs to."""
    from orb_slam2_test_b200 import BorbError, ORBextractor
    a = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=2)
    t = synth.make_sequence(synth.TUM_SHAPE, seed=96, start=0, count=2)
    a.submit_batch(t)
    with pytest.raises(BorbError) as ei:
        a.submit_batch(np.zeros((1, 376, 1241), np.uint8))
    assert ei.value.code == -5
    a.wait()
    want = ORBextractor(1000, 1.2, 8, 20, 7, max_batch=2).extract_batch(t)
    for f in range(2):
        _same(*a.result(f), *want[f], f"pending call survived, frame {f}")
    l, r = ORBextractor(1000, 1.2, 8, 20, 7), ORBextractor(1000, 1.2, 8, 20, 7)
    l.set_camera((517.3, 516.5, 318.6, 255.3), (0.0, 0.0, 0.0, 0.0))
    l(synth.make_frame(synth.TUM_SHAPE, seed=97, eye=0))
    r(synth.make_frame(synth.TUM_SHAPE, seed=97, eye=1))
    l.stereo_match(r, BF, MB)
    l.stereo_result(0)
    l(synth.make_frame(synth.TUM_SHAPE, seed=97, frame=1, eye=0))
    with pytest.raises(BorbError):
        l.stereo_result(0)                        # belongs to the previous extract call
    last = {"mp_xyz": np.zeros((1, 3), np.float32), "mp_desc": np.zeros((1, 32), np.uint8), "mp_flags": np.zeros(1, np.uint8),
            "octave": np.zeros(1, np.int32), "angle": np.zeros(1, np.float32), "Tcw": np.eye(4, dtype=np.float32)[:3]}
    with pytest.raises(BorbError):
        l.search_by_projection(last, np.eye(4, dtype=np.float32)[:3], BF, MB, 7.0, use_stereo=True)
    l.search_by_projection(last, np.eye(4, dtype=np.float32)[:3], BF, MB, 7.0, use_stereo=False)


def test_kseq_prefix_equals_the_reference_frame_digests():
    """BASELINE configs[4]: the first 64 pairs of the synthetic KITTI-00-shaped sequence through a stereo stream give,
    pair by pair, the digests the reference's own src/Frame.cc + ORBextractor.cc produced for them
    (tests/golden/kseq_digests.json, generator make_golden_kseq.py), and the prefix digest of digests."""
    import json
    from pathlib import Path
    from orb_slam2_test_b200 import OrbStream, sequence
    gold = json.loads((Path(__file__).parent / "golden" / "kseq_digests.json").read_text())
    n = 64
    left = np.empty((n,) + synth.KITTI_SHAPE, np.uint8)
    right = np.empty_like(left)
    synth.kseq_fill(range(n), left, right)
    st = OrbStream(2000, 1.2, 8, 20, 7, 1241, 376, batch=24, slots=3, stereo=True)
    st.set_stereo(synth.KSEQ_BF, synth.KSEQ_BF / synth.KSEQ_K[0])
    st.set_camera(synth.KSEQ_K, (0.0, 0.0, 0.0, 0.0))
    res = sequence.run_stream_shard(st, range(n), left, right)
    bad = [r.frame for r in res if r.digest != gold["first_64"][r.frame]]
    assert not bad, f"pairs {bad[:8]} differ from the reference Frame's digests"
    assert sequence.sequence_digest(res, n) == gold["prefix"]["64"]
    assert sequence.sequence_digest(res[:16], 16) == gold["prefix"]["16"]
