"""Phase split of k_fast_cells in the single-frame shape (library built with -DBORB_FAST_PROBE, see k_fast.cuh):
BORB_LIB=$PWD/variants/probe.so python tools/probe_fast.py"""
import sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, '/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth

img = synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=0)
dev = torch.from_numpy(img).cuda()
one = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
lib = one._lib
n_cells = 1220
names = ["entry", "tma issued", "window landed", "geometry", "stage 1", "stage 2", "nms", "end"]
for rep in range(6):
    one.extract_device(dev.data_ptr(), 1, 1241, 376); one.sync()
    buf = np.zeros(2048 * 8, np.uint64)
    assert lib.borb_debug_fast_probe(buf.ctypes.data_as(C.c_void_p), buf.size) == 0
    t = buf.reshape(2048, 8)[:n_cells].astype(np.int64)
    t0 = t[:, 0].min()
    if rep < 3: continue
    print(f"rep {rep}: span first entry -> last end {t[:, 7].max() - t0} ns; entry skew p50 {int(np.percentile(t[:, 0] - t0, 50))} "
          f"p90 {int(np.percentile(t[:, 0] - t0, 90))} max {int((t[:, 0] - t0).max())}")
    d = np.diff(t, axis=1)
    for k in range(7):
        print(f"   {names[k]:>14} -> {names[k + 1]:<14} p50 {int(np.percentile(d[:, k], 50)):6d}  p90 {int(np.percentile(d[:, k], 90)):6d}  max {int(d[:, k].max()):6d}")
    tot = t[:, 7] - t[:, 0]
    print(f"   CTA total p50 {int(np.percentile(tot, 50))} p90 {int(np.percentile(tot, 90))} max {int(tot.max())}; slowest CTA {int(tot.argmax())}")
