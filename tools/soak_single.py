"""One-off soak of the single-frame shape (programmatic launches, 1024-thread quadtree, level chain, results by kernel)
against the batched shape on many different frames: python tools/soak_single.py [n_frames]"""
import sys, time
import numpy as np
sys.path.insert(0, '/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth

n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 1536
B = 64
for shape, nf, tag in ((synth.KITTI_SHAPE, 2000, "K2k"), (synth.TUM_SHAPE, 1000, "T1k")):
    big = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=B, want_pyramid=False)
    one = ORBextractor(nf, 1.2, 8, 20, 7, max_batch=2, want_pyramid=False)
    differ = total_kp = 0
    t0 = time.time()
    for b0 in range(0, n_frames, B):
        frames = [synth.make_noise_frame(shape, seed=f) if f % 97 == 0 else synth.make_frame(shape, seed=1000 + f // 16, frame=f % 16, eye=f & 1)
                  for f in range(b0, b0 + B)]
        res = big.extract_batch(np.stack(frames))
        for i, (img, (k, d)) in enumerate(zip(frames, res)):
            if i % 5 == 4 and i + 1 < B:   # now and then as a pair (the two-frame chain shape)
                pair = one.extract_batch(np.stack([img, frames[i + 1]]))
                k1, d1 = pair[0]
                k2, d2 = pair[1]
                differ += not (k2.tobytes() == res[i + 1][0].tobytes() and d2.tobytes() == res[i + 1][1].tobytes())
            else:
                k1, d1 = one(img)
            differ += not (k1.tobytes() == k.tobytes() and d1.tobytes() == d.tobytes())
            total_kp += len(k)
    print(f"{tag}: {n_frames} frames, {total_kp} keypoints, {differ} frames differ between the shapes ({time.time() - t0:.0f} s)")
    assert differ == 0
