"""Where the host-side part of a single-frame call goes: pageable input, pinned input, device input + fetch."""
import sys, time, ctypes as C
import numpy as np, torch
sys.path.insert(0, '/root/repo')
from orb_slam2_test_b200 import ORBextractor, synth

img = synth.make_frame(synth.KITTI_SHAPE, seed=7, frame=0)
ext = ORBextractor(2000, 1.2, 8, 20, 7, max_batch=1, want_pyramid=False)
lib, h = ext._lib, ext._h
rows, cols = img.shape


def p50(fn, n=400):
    for _ in range(30): fn()
    t = []
    for _ in range(n):
        a = time.perf_counter(); fn(); t.append((time.perf_counter() - a) * 1e6)
    return round(float(np.percentile(t, 50)), 1)


print("pageable input, full Python call (results as numpy):", p50(lambda: ext(img)))
raw = lambda ptr: lib.borb_extract_host(h, ptr, 1, cols, rows, cols, rows * cols, 0)
print("pageable input, C ABI call only:", p50(lambda: raw(img.ctypes.data)))
lib.borb_host_alloc.restype = C.c_void_p
lib.borb_host_alloc.argtypes = [C.c_size_t]
pin = lib.borb_host_alloc(img.size)
C.memmove(pin, img.ctypes.data, img.size)
print("pinned input, C ABI call only:", p50(lambda: raw(pin)))
dev = torch.from_numpy(img).cuda()
def devcall():
    ext.extract_device(dev.data_ptr(), 1, cols, rows); ext.sync()
print("device input, kernels + sync (no result download):", p50(devcall))
def devfetch():
    ext.extract_device(dev.data_ptr(), 1, cols, rows); lib.borb_fetch(h, 0)
print("device input + result download:", p50(devfetch))
