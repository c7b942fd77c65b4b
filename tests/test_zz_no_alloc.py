"""borb_undistort_grid_noalloc (include/borb.h): the allocation-free variant of the per-keypoint step.

No counterpart in the reference -- added for callers that cannot call
cudaMalloc.  Its scratch memory is a static __device__ array inside the library; inputs and outputs are the caller's
device buffers.  Checked here THROUGH THE C ABI against the existing entry point (borb_set_camera + extract +
borb_frame_grid) on the same keypoints: bit-exact (the tolerance every parity test of this repository uses), on two
calls in a row with different frames, so that what the first call left in the shared scratch array would show.

Collected last on purpose (file name).
"""
import ctypes as C

import numpy as np
import pytest

from orb_slam2_test_b200 import synth

pytestmark = pytest.mark.gpu

TUM1_K = (517.306408, 516.469215, 318.643040, 255.313989)
TUM1_D = (0.262383, -0.953104, -0.005358, 0.002628, 1.163314)
CELLS = 64 * 48


def _existing(frames, K, dist, nfeatures=1000):
    """The existing entry points: extract with a camera set, then borb_frame_grid per frame."""
    from orb_slam2_test_b200 import ORBextractor
    ext = ORBextractor(nfeatures, 1.2, 8, 20, 7, max_batch=len(frames), want_pyramid=False)
    ext.set_camera(K, dist)
    res = ext.extract_batch(frames)
    stride = ext._lib.borb_max_keypoints(ext._h)
    grids = [ext.frame_grid(f) for f in range(len(frames))]
    return [k for k, _ in res], grids, stride


def _call_noalloc(lib, torch, kps_list, stride, cols, rows, K, dist, stream_ptr=None):
    """Upload the keypoints, call the new entry point on caller-owned device buffers, bring the outputs back."""
    from orb_slam2_test_b200.extractor import KP_DTYPE
    n = len(kps_list)
    host = np.zeros((n, stride), KP_DTYPE)
    counts = np.zeros(n, np.int32)
    for f, k in enumerate(kps_list):
        host[f, :len(k)] = k
        counts[f] = len(k)
    d_kps = torch.from_numpy(host.view(np.uint8).reshape(n, stride * 28)).cuda()
    d_cnt = torch.from_numpy(counts).cuda()
    # outputs pre-filled with a pattern: everything the call should write, it must write
    d_un = torch.full((n, stride, 2), -7.0, dtype=torch.float32, device="cuda")
    d_cs = torch.full((n, CELLS + 1), -7, dtype=torch.int32, device="cuda")
    d_ci = torch.full((n, stride), -7, dtype=torch.int32, device="cuda")
    d_bd = torch.full((n, 4), -7.0, dtype=torch.float32, device="cuda")
    k4 = np.asarray(K, np.float32)
    d = np.asarray(dist, np.float32)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = lib.borb_undistort_grid_noalloc(d_kps.data_ptr(), d_cnt.data_ptr(), n, stride, cols, rows, k4.ctypes.data,
                                         d.ctypes.data if len(d) else None, len(d), d_un.data_ptr(), d_cs.data_ptr(),
                                         d_ci.data_ptr(), d_bd.data_ptr(), stream_ptr)
    assert rc == 0, lib.borb_last_error().decode()
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return d_un.cpu().numpy(), d_cs.cpu().numpy(), d_ci.cpu().numpy(), d_bd.cpu().numpy(), counts, free_before - free_after


def _compare(out, grids, tag):
    un, cs, ci, bd, counts, _ = out
    for f, (wun, wstart, widx, wb) in enumerate(grids):
        n = int(counts[f])
        assert len(wun) == n
        assert un[f, :n].tobytes() == wun.tobytes(), f"{tag} frame {f}: undistorted positions differ"
        assert np.array_equal(cs[f], wstart), f"{tag} frame {f}: cell_start differs"
        m = int(wstart[-1])
        assert np.array_equal(ci[f, :m], widx), f"{tag} frame {f}: cell_idx differs"
        assert bd[f].tobytes() == np.asarray(wb, np.float32).tobytes(), f"{tag} frame {f}: bounds differ"


def test_noalloc_entry_point_equals_the_existing_one_on_two_calls_in_a_row():
    import torch
    from orb_slam2_test_b200 import load_library
    lib = load_library()
    # call 1: three distorted TUM frames; call 2: two OTHER frames with fewer keypoints per frame (so the scratch
    # array still holds call 1's cell numbers beyond what call 2 writes) and then an undistorted KITTI camera
    fa = synth.make_sequence(synth.TUM_SHAPE, seed=201, start=0, count=3)
    fb = synth.make_sequence(synth.TUM_SHAPE, seed=202, start=5, count=2)
    ka, ga, sa = _existing(fa, TUM1_K, TUM1_D, nfeatures=1000)
    kb, gb, sb = _existing(fb, TUM1_K, TUM1_D, nfeatures=500)
    assert all(len(k) >= 1000 for k in ka) and all(500 <= len(k) < 1000 for k in kb)
    out_a = _call_noalloc(lib, torch, ka, sa, 640, 480, TUM1_K, TUM1_D)
    out_b = _call_noalloc(lib, torch, kb, sb, 640, 480, TUM1_K, TUM1_D)
    _compare(out_a, ga, "call 1")
    _compare(out_b, gb, "call 2")
    assert out_b[5] == 0, f"the second call changed the device's free memory by {out_b[5]} bytes"
    # the identity shortcut (k1 == 0, src/Frame.cc:544-548) and a non-default stream
    fc = synth.make_sequence(synth.KITTI_SHAPE, seed=203, start=0, count=2)
    Kk = (718.856, 718.856, 607.1928, 185.2157)
    kc, gc, sc = _existing(fc, Kk, (0.0, 0.0, 0.0, 0.0), nfeatures=2000)
    st = torch.cuda.Stream()
    out_c = _call_noalloc(lib, torch, kc, sc, 1241, 376, Kk, (0.0, 0.0, 0.0, 0.0), stream_ptr=st.cuda_stream)
    _compare(out_c, gc, "call 3 (own stream)")
    assert out_c[5] == 0
    # call 1 again: same bytes as the first time
    out_a2 = _call_noalloc(lib, torch, ka, sa, 640, 480, TUM1_K, TUM1_D)
    for x, y in zip(out_a[:4], out_a2[:4]):
        assert x.tobytes() == y.tobytes()


def test_noalloc_entry_point_rejects_what_does_not_fit():
    import torch
    from orb_slam2_test_b200 import load_library
    lib = load_library()
    buf = torch.zeros(1 << 16, dtype=torch.uint8, device="cuda")
    p = buf.data_ptr()
    k4 = np.asarray(TUM1_K, np.float32)
    args = lambda n, stride: (p, p, n, stride, 640, 480, k4.ctypes.data, None, 0, p, p, p, p, None)  # noqa: E731
    max_slots = 262144  # BORB_NOALLOC_MAX_SLOTS
    assert lib.borb_undistort_grid_noalloc(*args(max_slots // 1024 + 1, 1024)) == -4
    msg = lib.borb_last_error().decode()
    assert "BORB_NOALLOC_MAX_SLOTS" in msg and str(max_slots) in msg
    assert lib.borb_undistort_grid_noalloc(*args(0, 8)) == -1
    assert lib.borb_undistort_grid_noalloc(None, p, 1, 8, 640, 480, k4.ctypes.data, None, 0, p, p, p, p, None) == -1
    torch.cuda.synchronize()
