"""The C++ drop-in class (orb_slam2_test_b200/shim/ORBextractor.{h,cc}) driven the way the
reference's Frame.cc / Tracking.cc drive theirs (tests/shim/frame_like.cc)."""
import re
import struct
import subprocess
from pathlib import Path

import numpy as np
import pytest

from orb_slam2_test_b200 import synth

ROOT = Path(__file__).resolve().parents[1]
SHIM = ROOT / "orb_slam2_test_b200" / "shim"
EXE = ROOT / "tests" / "shim" / "frame_like"
KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
               ("octave", "<i4"), ("class_id", "<i4")])

# the public interface of the reference class, include/ORBextractor.h:50-101 (whitespace-normalised)
REFERENCE_PUBLIC = [
    "enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };",
    "ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);",
    "void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint> &keypoints, cv::OutputArray descriptors);",
    "int inline GetLevels()",
    "float inline GetScaleFactor()",
    "std::vector<float> inline GetScaleFactors()",
    "std::vector<float> inline GetInverseScaleFactors()",
    "std::vector<float> inline GetScaleSigmaSquares()",
    "std::vector<float> inline GetInverseScaleSigmaSquares()",
    "std::vector<cv::Mat> mvImagePyramid;",
]


def _norm(s):
    s = re.sub(r"//[^\n]*", "", s)
    s = re.sub(r"\s+", " ", s)
    return re.sub(r"\s*([(),;&{}=])\s*", r"\1", s)


def test_public_interface_is_the_reference_interface():
    hdr = _norm((SHIM / "ORBextractor.h").read_text())
    assert "namespace ORB_SLAM2" in hdr and "class ORBextractor" in hdr and "class ExtractorNode" in hdr
    for sig in REFERENCE_PUBLIC:
        assert _norm(sig) in hdr, sig
    ref = Path("/root/reference/include/ORBextractor.h")
    if ref.exists():  # in the build container also check against the reference header itself
        rh = _norm(ref.read_text())
        for sig in REFERENCE_PUBLIC[1:]:
            assert _norm(sig) in rh, f"test list drifted from the reference: {sig}"


def test_shim_has_no_pixel_code():
    """the shim is a host wrapper: no image arithmetic, nothing from oracle/"""
    src = (SHIM / "ORBextractor.cc").read_text()
    for bad in ("cv::resize", "GaussianBlur", "cv::FAST", "fastAtan2", "copyMakeBorder", "oracle"):
        assert bad not in src


def _build():
    subprocess.run(["make", "-s", "-C", str(SHIM)], check=True)
    subprocess.run(["make", "-s", "-C", str(ROOT / "tests" / "shim")], check=True)


def test_fails_loudly_without_a_gpu(tmp_path):
    from conftest import HAVE_GPU
    if HAVE_GPU:
        pytest.skip("a GPU is present")
    _build()
    img = synth.make_frame((240, 320), seed=1, frame=0)
    (tmp_path / "l.raw").write_bytes(img.tobytes())
    r = subprocess.run([str(EXE), str(tmp_path / "l.raw"), str(tmp_path / "l.raw"), "320", "240", "500",
                        str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode != 0
    assert "no CPU path" in r.stderr or "runtime_error" in r.stderr


def _read_block(buf, off):
    n, = struct.unpack_from("<i", buf, off); off += 4
    k = np.frombuffer(buf, KP, n, off); off += 28 * n
    d = np.frombuffer(buf, np.uint8, 32 * n, off).reshape(n, 32); off += 32 * n
    lv = []
    for _ in range(8):
        c, r, s = struct.unpack_from("<iiq", buf, off); off += 16
        lv.append((c, r, s))
    return k, d, lv, off


def _level_checksum(L):
    rows, cols = L.shape
    w, s = 5, 0
    wt = (1 + (np.arange(121) % 7)).reshape(11, 11).astype(np.int64)
    for y in range(w, rows - w, 11):
        for x in range(w, cols - w, 11):
            s += int((L[y - w:y + w + 1, x - w:x + w + 1].astype(np.int64) * wt).sum())
    return s


@pytest.mark.gpu
@pytest.mark.parametrize("shape,nf", [(synth.KITTI_SHAPE, 2000), (synth.TUM_SHAPE, 1000)])
def test_frame_like_caller_matches_oracle(shape, nf, tmp_path, oracle_built):
    _build()
    left = synth.make_frame(shape, seed=31, frame=0, eye=0)
    right = synth.make_frame(shape, seed=31, frame=0, eye=1)
    (tmp_path / "l.raw").write_bytes(left.tobytes())
    (tmp_path / "r.raw").write_bytes(right.tobytes())
    r = subprocess.run([str(EXE), str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(shape[1]), str(shape[0]),
                        str(nf), str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    buf = (tmp_path / "o.bin").read_bytes()
    off = 0
    for img, n_feat in ((left, nf), (right, nf), (left, 2 * nf)):
        k, d, lv, off = _read_block(buf, off)
        o = oracle_built.Oracle(n_feat, 1.2, 8, 20, 7)
        ok, od = o.extract(img)
        assert k.tobytes() == ok.tobytes() and d.tobytes() == od.tobytes()
        for l in range(8):
            c, rws = o.level_size(l)
            assert (lv[l][0], lv[l][1]) == (c, rws)
            assert lv[l][2] == _level_checksum(o.pyramid(l, 0))
    untouched, = struct.unpack_from("<i", buf, off)
    assert untouched == 1
    # the stereo association the Frame-like caller ran on the device
    ns, = struct.unpack_from("<i", buf, off + 4)
    u = np.frombuffer(buf, "<f4", ns, off + 8)
    z = np.frombuffer(buf, "<f4", ns, off + 8 + 4 * ns)
    py = oracle_built
    ol, orr = py.Oracle(nf, 1.2, 8, 20, 7), py.Oracle(nf, 1.2, 8, 20, 7)
    kl, dl = ol.extract(left)
    kr, dr = orr.extract(right)
    wu, wz, _, kept = py.stereo_match(kl, dl, kr, dr, [ol.pyramid(l, 0) for l in range(8)],
                                      [orr.pyramid(l, 0) for l in range(8)], ol.table(0), ol.table(1),
                                      np.float32(386.1448), np.float32(386.1448) / np.float32(718.856))
    assert ns == len(kl) and kept > 100
    assert u.tobytes() == wu.tobytes() and z.tobytes() == wz.tobytes()
    # mvKeysUn, mGrid and the image bounds the caller got from FillFrame (camera: TUM1.yaml)
    off += 8 + 8 * ns
    nu, = struct.unpack_from("<i", buf, off); off += 4
    un = np.frombuffer(buf, "<f4", 2 * nu, off).reshape(nu, 2); off += 8 * nu
    gcount = np.frombuffer(buf, "<i4", 64 * 48, off); off += 4 * 64 * 48
    ng, = struct.unpack_from("<i", buf, off); off += 4
    gidx = np.frombuffer(buf, "<i4", ng, off); off += 4 * ng
    bnd = np.frombuffer(buf, "<f4", 4, off)
    from test_oracle_frame import CAMS
    wun, wstart, widx, wb = py.frame_post(kl, shape[1], shape[0], *CAMS["TUM1"])
    assert nu == len(kl) and un.tobytes() == wun.tobytes() and bnd.tobytes() == wb.tobytes()
    assert np.array_equal(gcount, np.diff(wstart)) and np.array_equal(gidx, widx)


@pytest.mark.gpu
def test_prestep_through_the_shim(tmp_path, oracle_built):
    """Raw RGB frame + rectification maps handed to the C++ class (SetColorInput / SetRectifyMap)."""
    _build()
    rows, cols = synth.EUROC_SHAPE
    img = synth.make_color_frame(synth.EUROC_SHAPE, seed=43, channels=3)
    mx, my = synth.rectify_maps(*synth.EUROC[0], size=(cols, rows))
    (tmp_path / "c.raw").write_bytes(img.tobytes())
    (tmp_path / "mx.f32").write_bytes(mx.tobytes())
    (tmp_path / "my.f32").write_bytes(my.tobytes())
    r = subprocess.run([str(EXE), "pre", str(tmp_path / "c.raw"), str(cols), str(rows), str(tmp_path / "mx.f32"),
                        str(tmp_path / "my.f32"), "1200", str(tmp_path / "o.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    k, d, lv, _ = _read_block((tmp_path / "o.bin").read_bytes(), 0)
    gray = oracle_built.cvt_gray(oracle_built.remap_linear(img, mx, my), True)
    ok, od = oracle_built.Oracle(1200, 1.2, 8, 20, 7).extract(gray)
    assert k.tobytes() == ok.tobytes() and d.tobytes() == od.tobytes()
    assert lv[0][2] == _level_checksum(gray)


@pytest.mark.gpu
def test_frame_to_frame_matcher_through_the_shim(tmp_path, oracle_built):
    """ORBextractor::SearchByProjection (shim) = ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th,
    bMono) restated (oracle/match_oracle.cpp, pinned to the reference's ORBmatcher.cc): the Frame-like caller
    builds the current stereo frame on the device and matches the last frame's map points against it."""
    import matchcase
    _build()
    bf = np.float32(386.1448)
    mb = bf / np.float32(718.856)  # as tests/shim/frame_like.cc computes it
    cur, last, cam, bounds, scale = matchcase.make_case(seed=33, bf=float(bf), mb=float(mb), dz=1.0)
    shape = synth.KITTI_SHAPE
    (tmp_path / "l.raw").write_bytes(synth.make_frame(shape, seed=33, frame=1, eye=0).tobytes())
    (tmp_path / "r.raw").write_bytes(synth.make_frame(shape, seed=33, frame=1, eye=1).tobytes())
    th = 7.0
    n = len(last["kps"])
    T4 = lambda T: np.vstack([T, [0, 0, 0, 1]]).astype("<f4")
    blob = struct.pack("<i4ffffi", n, *matchcase.KITTI_K, float(bf), float(mb), th, 0)
    blob += T4(cur["Tcw"]).tobytes() + T4(last["Tcw"]).tobytes()
    blob += last["mp_xyz"].astype("<f4").tobytes() + last["mp_desc"].tobytes() + last["mp_flags"].tobytes()
    blob += last["kps"]["octave"].astype("<i4").tobytes() + last["kps"]["angle"].astype("<f4").tobytes()
    (tmp_path / "last.bin").write_bytes(blob)
    r = subprocess.run([str(EXE), "match", str(tmp_path / "l.raw"), str(tmp_path / "r.raw"), str(shape[1]), str(shape[0]),
                        "2000", str(tmp_path / "last.bin"), str(tmp_path / "m.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    buf = (tmp_path / "m.bin").read_bytes()
    nm, nc = struct.unpack_from("<ii", buf, 0)
    got = np.frombuffer(buf, "<i4", nc, 8)
    want, wnm = oracle_built.search_by_projection(cur, last, cam, bounds, scale, th)
    assert nc == len(want) and np.array_equal(got, want) and nm == wnm and nm > 300
    # ORBextractor::SearchLocalPoints on the same frame: the caller wrote the projections it used next to its result
    off = 8 + 4 * nc
    nlocal, nl = struct.unpack_from("<ii", buf, off); off += 8
    gl = np.frombuffer(buf, "<i4", nl, off); off += 4 * nl
    proj = np.frombuffer(buf, "<f4", 4 * n, off).reshape(n, 4); off += 16 * n
    lflags = np.frombuffer(buf, "u1", n, off)
    mps = {"proj": proj, "desc": last["mp_desc"], "flags": lflags, "level": last["kps"]["octave"].astype(np.int32)}
    occ = np.where(want >= 0, 2, 0).astype(np.uint8)
    wl, wlnm = oracle_built.search_local_points(cur, bounds, scale, mps, occ, 3.0, 0.8)
    assert nl == len(wl) and np.array_equal(gl, wl) and nlocal == wlnm and nlocal > 50


@pytest.mark.gpu
def test_relocalisation_matcher_through_the_shim(tmp_path, oracle_built):
    """ORBextractor::SearchByProjectionKF (shim) = ORBmatcher::SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th,
    ORBdist) restated (oracle/match_oracle.cpp, pinned to the reference's ORBmatcher.cc): the Frame-like caller extracts the
    current frame and matches a keyframe's map points against it."""
    import matchcase
    _build()
    cur, cam, bounds, scale, log_scale, kf, occ = matchcase.make_reloc_case(seed=35, copies=2)
    shape = synth.KITTI_SHAPE
    (tmp_path / "cur.raw").write_bytes(synth.make_frame(shape, seed=35, frame=1, eye=0).tobytes())
    n = len(kf["flags"])
    T4 = np.vstack([cur["Tcw"], [0, 0, 0, 1]]).astype("<f4")
    blob = struct.pack("<ii4ffi", n, len(occ), *matchcase.KITTI_K, 10.0, 100) + T4.tobytes()
    blob += kf["xyz"].astype("<f4").tobytes() + kf["desc"].tobytes() + kf["flags"].tobytes()
    blob += kf["dist"].astype("<f4").tobytes() + kf["angle"].astype("<f4").tobytes() + occ.tobytes()
    (tmp_path / "kf.bin").write_bytes(blob)
    r = subprocess.run([str(EXE), "reloc", str(tmp_path / "cur.raw"), str(shape[1]), str(shape[0]), "2000",
                        str(tmp_path / "kf.bin"), str(tmp_path / "r.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr + r.stdout
    buf = (tmp_path / "r.bin").read_bytes()
    nm, nc = struct.unpack_from("<ii", buf, 0)
    got = np.frombuffer(buf, "<i4", nc, 8)
    want, wnm = oracle_built.search_by_projection_kf(cur, cam, bounds, scale, log_scale, kf, occ, 10.0, 100, True)
    assert nc == len(want) and np.array_equal(got, want) and nm == wnm and nm > 300


@pytest.mark.gpu
def test_initialisation_matcher_through_the_shim(tmp_path, oracle_built):
    """ORBextractor::SearchForInitialization (shim) = ORBmatcher::SearchForInitialization restated (pinned to the
    reference's ORBmatcher.cc): the Frame-like caller extracts the current frame with the initialisation extractor and
    matches the initial frame's level-0 keypoints against it."""
    import matchcase
    _build()
    f1, f2, bounds, prev = matchcase.make_init_case(seed=34)
    (tmp_path / "cur.raw").write_bytes(synth.make_frame(synth.TUM_SHAPE, seed=34, frame=1).tobytes())
    n = len(f1["kps"])
    blob = struct.pack("<iif", n, 100, 0.9) + f1["kps"]["octave"].astype("<i4").tobytes()
    blob += f1["kps"]["angle"].astype("<f4").tobytes() + f1["desc"].tobytes() + prev.astype("<f4").tobytes()
    (tmp_path / "f1.bin").write_bytes(blob)
    r = subprocess.run([str(EXE), "init", str(tmp_path / "cur.raw"), "640", "480", "2000", str(tmp_path / "f1.bin"),
                        str(tmp_path / "i.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    buf = (tmp_path / "i.bin").read_bytes()
    nm, nn = struct.unpack_from("<ii", buf, 0)
    got = np.frombuffer(buf, "<i4", nn, 8)
    gprev = np.frombuffer(buf, "<f4", 2 * nn, 8 + 4 * nn).reshape(nn, 2)
    want, wprev, wnm = oracle_built.search_for_initialization(f1, f2, bounds, prev, 100, np.float32(0.9), True)
    assert nn == n and np.array_equal(got, want) and nm == wnm > 150
    assert np.array_equal(gprev.view(np.uint32), wprev.view(np.uint32))


@pytest.mark.gpu
def test_bag_of_words_through_the_shim(tmp_path, oracle_built):
    """DeviceVocabulary / ORBextractor::ComputeBoW / ORBextractor::SearchByBoW (shim) = Frame::ComputeBoW (src/Frame.cc:532-539)
    and ORBmatcher::SearchByBoW(KeyFrame*, Frame&) (src/ORBmatcher.cc:195-351) restated (pinned to the reference's own
    DBoW2 + Frame.cc + ORBmatcher.cc): std::map contents identical, word values as float64 bit patterns."""
    import bowcase
    from oracle import pyoracle
    _build()
    path = bowcase.vocabulary(tmp_path, "k10L4")
    want = pyoracle.BowVocabulary(path)
    kk, kd = bowcase.frame_features(synth.TUM_SHAPE, 3, 0, 1000)
    fk, fd = bowcase.frame_features(synth.TUM_SHAPE, 3, 1, 1000)
    mp = bowcase.keyframe_flags(len(kk), 44)
    (tmp_path / "cur.raw").write_bytes(synth.make_frame(synth.TUM_SHAPE, seed=3, frame=1).tobytes())
    (tmp_path / "kf.bin").write_bytes(struct.pack("<if", len(kk), 0.7) + kk["angle"].astype("<f4").tobytes() + kd.tobytes() + mp.tobytes())
    r = subprocess.run([str(EXE), "bow", str(tmp_path / "cur.raw"), "640", "480", "1000", path, str(tmp_path / "kf.bin"),
                        str(tmp_path / "b.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    buf = (tmp_path / "b.bin").read_bytes()
    nm, nf = struct.unpack_from("<ii", buf, 0)
    got = np.frombuffer(buf, "<i4", nf, 8)
    off = 8 + 4 * nf

    def read_bow(off):
        words, nodes, entries = struct.unpack_from("<iii", buf, off)
        off += 12
        wid = np.frombuffer(buf, "<u4", words, off); off += 4 * words
        wv = np.frombuffer(buf, "<f8", words, off); off += 8 * words
        nid = np.frombuffer(buf, "<u4", nodes, off); off += 4 * nodes
        cnt = np.frombuffer(buf, "<i4", nodes, off); off += 4 * nodes
        fi = np.frombuffer(buf, "<u4", entries, off); off += 4 * entries
        return {"word_id": wid, "word_val": wv, "node_id": nid, "node_start": np.concatenate([[0], np.cumsum(cnt)]).astype(np.int32),
                "feat_idx": fi}, off
    fbow, off = read_bow(off)
    kbow, off = read_bow(off)
    assert bowcase.same_bow(fbow, want.transform(fd, 4)) and bowcase.same_bow(kbow, want.transform(kd, 4))
    w, wnm = want.search_by_bow(kk, kd, mp, fk, fd, np.float32(0.7), True)
    assert nf == len(w) and np.array_equal(got, w) and nm == wnm > 50
    # DeviceVocabulary::SearchByBoW(keyframe, keyframe) = ORBmatcher::SearchByBoW(pKF1, pKF2, vpMatches12) (:634-777)
    nm12, n1 = struct.unpack_from("<ii", buf, off)
    got12 = np.frombuffer(buf, "<i4", n1, off + 8)
    i = np.arange(len(fk))
    mp2 = ((i % 4 != 0).astype(np.uint8) | ((i % 16 == 5).astype(np.uint8) << 1)).astype(np.uint8)
    w12, wnm12 = want.search_by_bow_kf(kk, kd, mp, fk, fd, mp2, np.float32(0.75), True)
    assert n1 == len(w12) and np.array_equal(got12, w12) and nm12 == wnm12 > 30
