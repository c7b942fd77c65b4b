"""Cost of a call whose device input buffer is new on every call (graph cache miss): capture vs direct launch."""
import sys, time
import numpy as np, torch
sys.path.insert(0, '/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth
img = synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=0)
ext = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
bufs = [torch.from_numpy(img).cuda() for _ in range(40)]
ext.extract_device(bufs[0].data_ptr(), 1, 1241, 376); ext.sync()
t = []
for b in bufs[1:]:
    a = time.perf_counter(); ext.extract_device(b.data_ptr(), 1, 1241, 376); ext.sync(); t.append((time.perf_counter() - a) * 1e6)
print("misses 1-7 (captured) p50 us", round(float(np.median(t[:7])), 1), "| misses 9+ (launched directly) p50 us", round(float(np.median(t[9:])), 1))
