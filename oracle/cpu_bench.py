"""Time the reference's CPU ORB extractor on this host's cores -- TEST/BENCH INFRASTRUCTURE.

    python -m oracle.cpu_bench --procs P --frames-per-proc N [--rows R --cols C --nfeatures F]

One OS process per core (the reference's heavy std::vector / std::list churn serialises badly on a
shared malloc arena when driven from threads of one process, so processes give the CPU its best
number), each owning one extractor instance and looping over seeded synthetic frames: the
"all host cores across frames" mode of BASELINE.md section 3.  Uses oracle/_ref (the reference's
own ORBextractor.cc compiled in place) when present, else the restatement.  Prints one JSON line.
Run as a separate process by bench.py so that it never shares an address space with CUDA.
"""
from __future__ import annotations

import argparse
import json
import multiprocessing as mp
import os
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))


def _pin(idx, per):
    """Pin this worker to `per` cores of its own (no migration, no two workers on one core)."""
    try:
        cpus = sorted(os.sched_getaffinity(0))
        mine = {cpus[(idx * per + k) % len(cpus)] for k in range(per)}
        os.sched_setaffinity(0, mine)
    except Exception:
        pass


def _worker_stereo(idx, args, barrier, out):
    """Kseq arm: the reference's own stereo Frame constructor (src/Frame.cc:79-151: two extraction threads,
    UndistortKeyPoints, ComputeStereoMatches, AssignFeaturesToGrid) on pairs of the synthetic sequence."""
    from oracle import pyoracle
    from orb_slam2_test_b200 import synth
    _pin(idx, 2)
    pairs = [synth.kseq_pair(idx * 4 + i) for i in range(min(4, args.frames_per_proc))]
    K, bf = synth.KSEQ_K, synth.KSEQ_BF
    run = lambda p: pyoracle.RefFrame.stereo(p[0], p[1], K, (0.0, 0.0, 0.0, 0.0), bf, bf / K[0], nfeatures=args.nfeatures,  # noqa: E731
                                             canonical=False)
    run(pairs[0])
    barrier.wait()
    t0 = time.perf_counter()
    for i in range(args.frames_per_proc):
        run(pairs[i % len(pairs)])
    out.put((idx, time.perf_counter() - t0, "reference"))


def _worker(idx, args, barrier, out):
    import numpy as np  # noqa: F401
    from oracle import pyoracle
    from orb_slam2_test_b200 import synth
    _pin(idx, 1)
    frames = synth.make_sequence((args.rows, args.cols), seed=7, start=0, count=min(4, args.frames_per_proc))
    if pyoracle.Reference.available(mono=False):
        kind = "reference"
        inst = pyoracle.Reference(args.nfeatures, 1.2, 8, 20, 7, mono=False)
    else:
        kind = "port"
        inst = pyoracle.Oracle(args.nfeatures, 1.2, 8, 20, 7)
    inst.extract(frames[0])
    barrier.wait()
    t0 = time.perf_counter()
    for i in range(args.frames_per_proc):
        inst.extract(frames[i % len(frames)])
    out.put((idx, time.perf_counter() - t0, kind))


def primitive_ratio(rows: int, cols: int) -> dict:
    """How the scalar-C OpenCV stand-ins the reference arm runs on (oracle/cvprim.c) compare with the SIMD
    OpenCV 4.13 wheel, one thread each, on the work one frame of this shape does: the 7 resizes of the pyramid
    chain, the per-cell FAST calls (30-px cells with the reference's +6 overlap, threshold 20) and the 8 blurs.
    ratio > 1: the stand-in is slower, i.e. the CPU arm is handicapped by that factor in that stage."""
    import ctypes as C

    import numpy as np
    try:
        import cv2
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"cv2: {type(e).__name__}"}
    from oracle import pyoracle
    from orb_slam2_test_b200 import synth
    cv2.setNumThreads(1)
    L = pyoracle.Oracle.load()
    u8p = C.POINTER(C.c_uint8)

    class KP(C.Structure):
        _fields_ = [("x", C.c_int), ("y", C.c_int), ("s", C.c_int)]
    L.cvp_resize_linear_u8.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, u8p, C.c_size_t, C.c_int, C.c_int]
    L.cvp_gaussian7_sigma2.argtypes = [u8p, C.c_size_t, u8p, C.c_size_t, C.c_int, C.c_int]
    L.cvp_fast9_16.argtypes = [u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.POINTER(KP), C.c_int]
    p = lambda a: a.ctypes.data_as(u8p)  # noqa: E731
    img = synth.make_frame((rows, cols), seed=7, frame=0)
    sc, levels = np.float32(1.0), [img]
    for _ in range(7):
        sc = np.float32(float(sc) * float(np.float32(1.2)))
        inv = np.float32(1.0) / sc
        w, h = int(np.rint(np.float32(cols) * inv)), int(np.rint(np.float32(rows) * inv))
        levels.append(cv2.resize(levels[-1], (w, h), interpolation=cv2.INTER_LINEAR))

    def best(fn, reps=5):
        t = []
        for _ in range(reps):
            t0 = time.perf_counter()
            fn()
            t.append(time.perf_counter() - t0)
        return min(t)

    outs = [np.zeros_like(l) for l in levels]

    def ours_resize():
        for a, b in zip(levels[:-1], outs[1:]):
            L.cvp_resize_linear_u8(p(a), a.shape[1], a.shape[1], a.shape[0], p(b), b.shape[1], b.shape[1], b.shape[0])

    def cv_resize():
        for a, b in zip(levels[:-1], outs[1:]):
            cv2.resize(a, (b.shape[1], b.shape[0]), dst=b, interpolation=cv2.INTER_LINEAR)

    def ours_blur():
        for a, b in zip(levels, outs):
            L.cvp_gaussian7_sigma2(p(a), a.shape[1], p(b), b.shape[1], a.shape[1], a.shape[0])

    def cv_blur():
        for a, b in zip(levels, outs):
            cv2.GaussianBlur(a, (7, 7), 2, dst=b, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)

    cells = []
    for l in levels:
        H, W = l.shape
        for y in range(16, H - 16 - 6, 30):
            for x in range(16, W - 16 - 6, 30):
                cells.append(np.ascontiguousarray(l[y:min(y + 36, H - 16), x:min(x + 36, W - 16)]))
    buf = (KP * 4096)()
    det = cv2.FastFeatureDetector_create(threshold=20, nonmaxSuppression=True)

    def ours_fast():
        for c in cells:
            L.cvp_fast9_16(p(c), c.shape[1], c.shape[1], c.shape[0], 20, buf, 4096)

    def cv_fast():
        for c in cells:
            det.detect(c)

    r = {}
    for name, a, b in (("resize", ours_resize, cv_resize), ("fast", ours_fast, cv_fast), ("blur", ours_blur, cv_blur)):
        ta, tb = best(a), best(b)
        r[name] = {"cvprim_ms": ta * 1e3, "cv2_ms": tb * 1e3, "ratio": ta / tb}
    r["what"] = ("per-frame time of oracle/cvprim.c (scalar C, what the reference arm runs on) / the cv2 4.13.0 wheel (SIMD), one "
                 f"thread, {cols}x{rows}: 7 chained resizes, {len(cells)} per-cell FAST calls at threshold 20, 8 blurs; the cv2 FAST "
                 "time includes the Python call and KeyPoint objects per cell")
    return r


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", type=int, default=len(os.sched_getaffinity(0)))
    ap.add_argument("--frames-per-proc", type=int, default=8)
    ap.add_argument("--rows", type=int, default=376)
    ap.add_argument("--cols", type=int, default=1241)
    ap.add_argument("--nfeatures", type=int, default=2000)
    ap.add_argument("--stereo", action="store_true", help="time the reference's stereo Frame constructor on pairs (2 threads per process)")
    ap.add_argument("--primitives", action="store_true", help="cvprim.c vs cv2 on one frame's resize / FAST / blur work, one thread")
    args = ap.parse_args()
    if args.primitives:
        print(json.dumps(primitive_ratio(args.rows, args.cols)))
        return
    ctx = mp.get_context("fork")
    barrier = ctx.Barrier(args.procs + 1)
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker_stereo if args.stereo else _worker, args=(i, args, barrier, out)) for i in range(args.procs)]
    for p in procs:
        p.start()
    barrier.wait()
    t0 = time.perf_counter()
    res = [out.get() for _ in procs]
    wall = time.perf_counter() - t0
    for p in procs:
        p.join()
    n = args.procs * args.frames_per_proc
    print(json.dumps({"frames_per_s": n / wall, "frames": n, "procs": args.procs, "wall_s": wall,
                      "ms_per_frame_per_core": 1e3 * sum(r[1] for r in res) / n, "kind": res[0][2]}))


if __name__ == "__main__":
    main()
