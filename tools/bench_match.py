import sys, time
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np
import matchcase
from oracle import pyoracle
from orb_slam2_test_b200 import ORBextractor, synth
cur,last,cam,bounds,scale = matchcase.make_case(seed=1)
bf,mb=float(cam[4]),float(cam[5])
L=ORBextractor(2000,1.2,8,20,7); L.set_camera(matchcase.KITTI_K,(0.,0.,0.,0.)); R=ORBextractor(2000,1.2,8,20,7)
L(synth.make_frame(synth.KITTI_SHAPE,seed=1,frame=1,eye=0)); R(synth.make_frame(synth.KITTI_SHAPE,seed=1,frame=1,eye=1)); L.stereo_match(R,bf,mb)
d=dict(last); d['octave']=last['kps']['octave']; d['angle']=last['kps']['angle']
for th in (7.0,15.0,30.0):
    for _ in range(5): g,nm=L.search_by_projection(d,cur['Tcw'],bf,mb,th,use_stereo=True)
    ts=[]
    for _ in range(50):
        t0=time.perf_counter(); g,nm=L.search_by_projection(d,cur['Tcw'],bf,mb,th,use_stereo=True); ts.append((time.perf_counter()-t0)*1e6)
    tc=[]
    for _ in range(20):
        t0=time.perf_counter(); w,wnm=pyoracle.search_by_projection(cur,last,cam,bounds,scale,th); tc.append((time.perf_counter()-t0)*1e6)
    print('th',th,'gpu call p50 us',np.median(ts),'rounds',L.match_rounds(),'rescans',L.match_rescans(),'nm',nm,'cpu oracle p50 us',np.median(tc), 'equal', np.array_equal(g,w))
