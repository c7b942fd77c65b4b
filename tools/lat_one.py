"""One single-frame call shape, a few times (for ncu launch lists of the latency shape)."""
import sys, numpy as np, torch
sys.path.insert(0, '/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth
img = synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=0)
dev = torch.from_numpy(img).cuda()
one = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
for _ in range(12):
    one.extract_device(dev.data_ptr(), 1, 1241, 376); one.sync()
