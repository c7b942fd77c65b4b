import sys, time, numpy as np, torch
sys.path.insert(0,'/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth
img = synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=0)
dev = torch.from_numpy(img).cuda()
one = ORBextractor(2000,1.2,8,20,7,max_batch=1,want_pyramid=False)
for _ in range(30): one.extract_device(dev.data_ptr(),1,1241,376); one.sync()
ks=[]
for _ in range(300):
    one.extract_device(dev.data_ptr(),1,1241,376); one.sync(); ks.append(one.last_device_ms()*1e3)
print('kernels only p50 us', np.percentile(ks,50), 'p10', np.percentile(ks,10))
st = [one.stage_times(dev.data_ptr(), 1, 1241, 376) for _ in range(50)]
print('stages us (median, serial launches):', {k: round(float(np.median([s[k] for s in st])) * 1e3, 1) for k in st[0]})
big = ORBextractor(2000,1.2,8,20,7,max_batch=512,want_pyramid=False)
batch = torch.from_numpy(np.stack([synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=f) for f in range(64)])).cuda().repeat(8,1,1).contiguous()
st = [big.stage_times(batch.data_ptr(), 512, 1241, 376) for _ in range(6)][1:]
print('stages ms per 512 frames:', {k: round(float(np.median([s[k] for s in st])), 3) for k in st[0]})
