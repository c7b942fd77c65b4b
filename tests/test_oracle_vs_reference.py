"""Pin the oracle restatement (oracle/orb_oracle.cpp) to the reference's own src/ORBextractor.cc,
compiled in place from /root/reference into oracle/_ref (prebuilt files travel to the GPU box).

libreforb_mono.so = the reference with a monotone node allocator, so that its heap-address
tie-break (src/ORBextractor.cc:864-869) is the canonical rule; libreforb.so = stock malloc.
"""
import numpy as np
import pytest

from orb_slam2_test_b200 import synth

PARAMS = [(1000, 1.2, 8, 20, 7), (2000, 1.2, 8, 20, 7), (2000, 1.2, 8, 12, 7), (1200, 1.2, 8, 20, 7),
          (500, 1.5, 4, 25, 10), (300, 1.1, 12, 20, 7), (1, 2.0, 2, 20, 7)]


@pytest.fixture(scope="module")
def ref(oracle_built):
    if not oracle_built.Reference.available(mono=True):
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return oracle_built


@pytest.mark.parametrize("p", PARAMS)
def test_constructor_tables(p, ref):
    """src/ORBextractor.cc:437-520: scale tables, per-level quotas, umax."""
    o, r = ref.Oracle(*p), ref.Reference(*p)
    for which in range(4):
        assert o.table(which).tobytes() == r.table(which).tobytes(), which
    assert np.array_equal(o.quotas(), r.quotas())
    assert np.array_equal(o.umax(), r.umax())
    assert o.umax().tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]


def test_quota_known_values(ref):
    assert ref.Oracle(1000, 1.2, 8, 20, 7).quotas().tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert ref.Oracle(2000, 1.2, 8, 20, 7).quotas().tolist() == [434, 362, 302, 251, 209, 175, 145, 122]


CASES = [("T1k", synth.TUM_SHAPE, (1000, 1.2, 8, 20, 7)), ("T2k", synth.TUM_SHAPE, (2000, 1.2, 8, 20, 7)),
         ("K2k", synth.KITTI_SHAPE, (2000, 1.2, 8, 20, 7)), ("K04", synth.KITTI_SHAPE, (2000, 1.2, 8, 12, 7)),
         ("EuRoC", (480, 752), (1200, 1.2, 8, 20, 7)), ("odd", (211, 333), (700, 1.3, 5, 30, 5))]


@pytest.mark.parametrize("name,shape,p", CASES, ids=[c[0] for c in CASES])
def test_end_to_end_bytes(name, shape, p, ref):
    o, r = ref.Oracle(*p), ref.Reference(*p, mono=True)
    for frame in range(2):
        img = synth.make_frame(shape, seed=13, frame=frame)
        ok, od = o.extract(img)
        rk, rd = r.extract(img)
        assert len(ok) == len(rk) and len(ok) > 0
        assert ok.tobytes() == rk.tobytes(), f"{name}: keypoints differ"
        assert od.tobytes() == rd.tobytes(), f"{name}: descriptors differ"
        for l in range(p[2]):
            assert o.level_size(l) == r.level_size(l)
            assert np.array_equal(o.pyramid(l, 19), r.pyramid(l, 19)), f"{name}: pyramid level {l}"


def test_noise_and_flat_and_strided(ref):
    p = (1000, 1.2, 8, 20, 7)
    o, r = ref.Oracle(*p), ref.Reference(*p, mono=True)
    img = synth.make_noise_frame(synth.TUM_SHAPE, seed=2)
    a, b = o.extract(img), r.extract(img)
    assert a[0].tobytes() == b[0].tobytes() and a[1].tobytes() == b[1].tobytes()
    flat = np.full((240, 320), 90, np.uint8)
    a, b = o.extract(flat), r.extract(flat)
    assert len(a[0]) == 0 and len(b[0]) == 0
    big = synth.make_frame((300, 500), seed=4, frame=0)
    roi = big[10:250, 30:400]  # step > cols, as a cv::Mat ROI
    a, b, c = o.extract(roi), r.extract(roi), o.extract(np.ascontiguousarray(roi))
    assert a[0].tobytes() == b[0].tobytes() == c[0].tobytes()
    assert a[1].tobytes() == b[1].tobytes() == c[1].tobytes()


def test_empty_image_is_the_early_return(ref):
    """src/ORBextractor.cc:1333-1334."""
    assert ref.Oracle(1000, 1.2, 8, 20, 7).extract(None) == -1
    assert ref.Reference(1000, 1.2, 8, 20, 7).extract(None) == -1


def _random_points(rng, n, w, h, max_resp):
    """distinct integer points (FAST + NMS never yields duplicates), responses with many ties"""
    idx = rng.choice(w * h, size=min(n, w * h), replace=False)
    xs, ys = (idx % w).astype(np.int32), (idx // w).astype(np.int32)
    order = np.lexsort((xs, ys))
    return xs[order], ys[order], rng.integers(7, max_resp, size=len(idx)).astype(np.int32)


QT = [(w, h, n, N, mr) for (w, h) in [(608, 448), (1209, 344), (300, 100), (97, 64), (120, 121), (200, 80)]
      for (n, N, mr) in [(5000, 434, 60), (600, 217, 9), (50, 100, 30), (3000, 10, 200), (1, 5, 20), (2000, 1, 8)]]


@pytest.mark.parametrize("w,h,n,N,mr", QT)
def test_quadtree_vs_reference(w, h, n, N, mr, ref):
    """DistributeOctTree (src/ORBextractor.cc:668-951) run on its own through the reference object
    (monotone allocator): same survivors in the same order, incl. response ties, size ties at the
    break point, fewer candidates than N, nIni 1..4, points on node boundaries."""
    rng = np.random.default_rng(w * 131 + h * 7 + n + N)
    xs, ys, rs = _random_points(rng, n, w, h, mr)
    want = ref.Reference(1000, 1.2, 8, 20, 7, mono=True).distribute(xs, ys, rs, 16, 16 + w, 16, 16 + h, N)
    got = ref.distribute(xs, ys, rs, 16, 16 + w, 16, 16 + h, N)
    assert np.array_equal(got, want)


def test_quadtree_boundary_points(ref):
    """points exactly on root boundaries x == hX*i and on split lines"""
    w, h = 1209, 344
    hx = w / 4.0
    xs, ys = [], []
    for i in range(1, 4):
        for d in (-1, 0, 1):
            for y in range(0, h, 7):
                xs.append(int(hx * i) + d); ys.append(y)
    for x in range(0, w, 5):
        for y in (171, 172, 173, 85, 86, 87):
            xs.append(x); ys.append(y)
    pts = sorted(set(zip(ys, xs)))
    ys = np.array([p[0] for p in pts], np.int32); xs = np.array([p[1] for p in pts], np.int32)
    rs = np.full(len(xs), 20, np.int32)
    for N in (50, 434, 5000):
        want = ref.Reference(1000, 1.2, 8, 20, 7, mono=True).distribute(xs, ys, rs, 16, 16 + w, 16, 16 + h, N)
        assert np.array_equal(ref.distribute(xs, ys, rs, 16, 16 + w, 16, 16 + h, N), want)


def test_raw_pointer_tiebreak_noise_floor(ref, capsys):
    """The stock reference breaks quadtree ties by heap address (src/ORBextractor.cc:864-869: sort of
    pair<int, ExtractorNode*>), so its keypoint set depends on the allocator's state.  Two numbers side by side
    (SURVEY.md B.4): canonical (creation-order tie-break) against the stock build, and the stock build against ITSELF
    after the heap has been perturbed (blocks of random sizes allocated and every other one freed between the two
    runs) -- informative, loose bounds."""
    import ctypes as C
    if not ref.Reference.available(mono=False):
        pytest.skip("no raw build")
    o = ref.Oracle(2000, 1.2, 8, 20, 7)
    r = ref.Reference(2000, 1.2, 8, 20, 7, mono=False)
    libc = C.CDLL("libc.so.6")
    libc.malloc.restype = C.c_void_p
    libc.malloc.argtypes = [C.c_size_t]
    libc.free.argtypes = [C.c_void_p]
    rng = np.random.default_rng(12)
    held = []

    def perturb():
        blocks = [libc.malloc(int(s)) for s in rng.integers(16, 4096, 4000)]
        for b in blocks[::2]:
            libc.free(b)
        held.extend(blocks[1::2])

    def keyset(k):
        return set(zip(k["x"].tolist(), k["y"].tolist(), k["octave"].tolist()))

    inter = union = inter_s = union_s = 0
    for f in range(3):
        img = synth.make_frame(synth.KITTI_SHAPE, seed=3, frame=f)
        a = keyset(o.extract(img)[0])
        b = keyset(r.extract(img)[0])
        perturb()
        b2 = keyset(r.extract(img)[0])
        inter += len(a & b); union += len(a | b)
        inter_s += len(b & b2); union_s += len(b | b2)
    for b in held:
        libc.free(b)
    j, js = inter / union, inter_s / union_s
    with capsys.disabled():
        print(f"\n[info] canonical vs stock-malloc reference keypoint-set Jaccard: {j:.4f} ({inter}/{union}); "
              f"stock-malloc reference vs itself after heap perturbation: {js:.4f} ({inter_s}/{union_s})")
    assert j > 0.97 and js > 0.97
