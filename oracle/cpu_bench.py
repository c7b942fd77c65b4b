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


def _worker(idx, args, barrier, out):
    import numpy as np  # noqa: F401
    from oracle import pyoracle
    from orb_slam2_test_b200 import synth
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--procs", type=int, default=len(os.sched_getaffinity(0)))
    ap.add_argument("--frames-per-proc", type=int, default=8)
    ap.add_argument("--rows", type=int, default=376)
    ap.add_argument("--cols", type=int, default=1241)
    ap.add_argument("--nfeatures", type=int, default=2000)
    args = ap.parse_args()
    ctx = mp.get_context("fork")
    barrier = ctx.Barrier(args.procs + 1)
    out = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(i, args, barrier, out)) for i in range(args.procs)]
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
