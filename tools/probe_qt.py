"""Phase log of k_quadtree in the single-frame shape (library built with -DBORB_QT_PROBE, see k_quadtree.cuh):
BORB_LIB=$PWD/variants/probe.so python tools/probe_qt.py"""
import sys, ctypes as C
import numpy as np, torch
sys.path.insert(0, '/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth

img = synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=0)
dev = torch.from_numpy(img).cuda()
one = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
lib = one._lib
names = {0: "entry", 1: "after wait (n)", 2: "roots counted", 3: "roots listed (S)", 4: "breadth pass (S)", 5: "ordered: ranked (E)",
         6: "ordered: counted", 7: "ordered: prefix (n_proc)", 8: "ordered pass (S)", 9: "end",
         10: "  scan: own sums + warp scan", 11: "  scan: barrier 1", 12: "  scan: warp 0", 13: "  scan: barrier 2", 14: "  scan: emit", 15: "  scan: barrier 3",
         16: "  relabel (+count)", 17: "  ordered: reset issued", 18: "  ordered: reset barrier", 19: "  ordered: ranks computed"}
for rep in range(5):
    one.extract_device(dev.data_ptr(), 1, 1241, 376); one.sync()
    if rep < 4: continue
    buf = np.zeros(8 * 128 * 2, np.uint64); cnt = np.zeros(8, np.int32)
    assert lib.borb_debug_qt_probe(buf.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p)) == 0
    log = buf.reshape(8, 128, 2)
    t_first = min(int(log[l, 0, 1]) for l in range(8))
    for l in range(8):
        print(f"level {l}: entry +{int(log[l, 0, 1]) - t_first} ns, total {int(log[l, cnt[l] - 1, 1]) - int(log[l, 0, 1])} ns")
        if l > 0: continue
        for k in range(1, cnt[l]):
            tag = int(log[l, k, 0])
            print(f"    +{int(log[l, k, 1]) - int(log[l, k - 1, 1]):6d} ns  {names[tag & 255]} {tag >> 8 if tag >> 8 else ''}")
