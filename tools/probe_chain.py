"""Phase split of k_pyr_chain in the single-frame shape (library built with -DBORB_PCH_PROBE, see k_pyramid.cuh):
BORB_LIB=$PWD/variants/probe.so python tools/probe_chain.py"""
import sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, '/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth

img = synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=0)
dev = torch.from_numpy(img).cuda()
one = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
lib = one._lib
n_tiles = 136
for rep in range(5):
    one.extract_device(dev.data_ptr(), 1, 1241, 376); one.sync()
    if rep < 3: continue
    buf = np.zeros(256 * 20, np.uint64)
    assert lib.borb_debug_chain_probe(buf.ctypes.data_as(C.c_void_p), buf.size) == 0
    t = buf.reshape(256, 20)[:n_tiles, :18].astype(np.int64)
    t0 = t[:, 0].min()
    print(f"rep {rep}: first entry -> last end {t[:, 17].max() - t0} ns; entry skew max {int((t[:, 0] - t0).max())}; "
          f"tile total p50 {int(np.percentile(t[:, 17] - t[:, 0], 50))} max {int((t[:, 17] - t[:, 0]).max())}")
    d = np.diff(t, axis=1)
    names = ["level 0 staged"] + [x for l in range(8) for x in (f"level {l} computed", f"level {l} written")]
    for k in range(17):
        print(f"   {names[k]:>18}  p50 {int(np.percentile(d[:, k], 50)):6d}  max {int(d[:, k].max()):6d}")
