#!/usr/bin/env python3
"""Generate tests/golden/kseq_digests.json: BASELINE.json configs[4] (the 4541-pair KITTI-00-shaped stereo
sequence) through the REFERENCE's own code -- the stereo Frame constructor of src/Frame.cc (both eyes extracted
by the reference's ORBextractor.cc on two threads, ComputeStereoMatches) compiled in place into
oracle/_ref/libreframe.so, with the monotone node allocator in the extraction threads (canonical quadtree
tie-break).

Run in the build container (needs /root/reference built into oracle/_ref):

    python tests/golden/make_golden_kseq.py [--pairs 4541] [--procs 8]

Per pair the digest is orb_slam2_test_b200.sequence.item_digest's: SHA-256 over mvKeys, mDescriptors,
mvKeysRight, mDescriptorsRight, mvuRight, mvDepth.  The file keeps the per-pair digests of the first 64 pairs, the
first 16 hex digits of every pair's digest (to localise a mismatch), and the digest-of-digests of the prefixes 16, 64, 256, 1024 and 4541, so that a GPU test can check a short
prefix in seconds and bench.py the whole sequence.
"""
import argparse
import hashlib
import json
import multiprocessing as mp
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
HERE = Path(__file__).resolve().parent
PREFIXES = (16, 64, 256, 1024, 4541)


def pair_digest(f: int):
    from oracle.pyoracle import RefFrame
    from orb_slam2_test_b200 import synth
    left, right = synth.kseq_pair(f)
    K, bf = synth.KSEQ_K, synth.KSEQ_BF
    r = RefFrame.stereo(left, right, K, (0.0, 0.0, 0.0, 0.0), bf, bf / K[0], nfeatures=2000, canonical=True)
    h = hashlib.sha256()
    for k in ("kps_l", "desc_l", "kps_r", "desc_r", "u_right", "depth"):
        h.update(np.ascontiguousarray(r[k]).tobytes())
    # canonical only while every quadtree node came from the monotone arena (oracle/ref_driver.cpp)
    lib = RefFrame.load()
    lib.reforb_malloc_fallbacks.restype = __import__("ctypes").c_long
    return f, h.hexdigest(), len(r["kps_l"]), len(r["kps_r"]), int((r["depth"] > 0).sum()), int(lib.reforb_malloc_fallbacks())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pairs", type=int, default=4541)
    ap.add_argument("--procs", type=int, default=4)
    args = ap.parse_args()
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(args.procs) as pool:
        res = pool.map(pair_digest, range(args.pairs), chunksize=16)
    res.sort()
    missed = max(r[5] for r in res)
    if missed != 0:
        raise SystemExit(f"{missed} quadtree nodes were allocated outside the monotone arena: the run is not canonical")
    dig = [r[1] for r in res]
    out = {"what": "reference src/Frame.cc stereo constructor (oracle/_ref/libreframe.so, canonical tie-break) over synth.kseq_pair(f)",
           "seed": 4541, "pairs": args.pairs, "first_64": dig[:64],
           # the first 16 hex digits of every pair's digest: enough to say WHICH pair differs when a prefix digest does
           "all_short": "".join(d[:16] for d in dig),
           "keypoints_left_mean": float(np.mean([r[2] for r in res])), "keypoints_right_mean": float(np.mean([r[3] for r in res])),
           "matched_mean": float(np.mean([r[4] for r in res])), "prefix": {}}
    for p in PREFIXES:
        if p <= args.pairs:
            h = hashlib.sha256()
            for d in dig[:p]:
                h.update(bytes.fromhex(d))
            out["prefix"][str(p)] = h.hexdigest()
    (HERE / "kseq_digests.json").write_text(json.dumps(out, indent=1) + "\n")
    print(f"{args.pairs} pairs in {time.perf_counter() - t0:.1f} s; prefixes {list(out['prefix'])}")


if __name__ == "__main__":
    main()
