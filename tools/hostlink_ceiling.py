#!/usr/bin/env python3
"""Aggregate host<->device copy ceiling of the box at N = 1/2/4/8 GPUs.

Every rank loops plain cudaMemcpyAsync of the host-fed extraction loop's exact sizes (one 128-frame KITTI
batch = 59.7 MB page-locked -> device, its results = 15.6 MB device -> page-locked, two upload and two download streams
(the feeder's two rings), best of three runs, no
kernels; orb_slam2_test_b200/hostlink.py); rank 0 prints one JSON line per N with the aggregate GB/s
(total bytes / slowest rank's time) and what that allows in frames/s.

    torchrun --nproc-per-node N tools/hostlink_ceiling.py            one N, under the launcher bench.py runs under
    python tools/hostlink_ceiling.py --ns 1,2,4,8 [--out FILE]       runs the line above for every N in turn
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))

ROWS, COLS, KP, SUB = 376, 1241, 2032, 128
H2D = SUB * ROWS * COLS                    # one 128-frame batch of KITTI frames
D2H = SUB * KP * 60 + SUB * 4 + 16         # its result slab: 28-B records + 32-B descriptors + counts


def one(args):
    import torch
    import torch.distributed as dist
    from orb_slam2_test_b200 import hostlink
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    if world > 1:
        dist.init_process_group("gloo")
    barrier = dist.barrier if world > 1 else None
    r = hostlink.measure(local, args.h2d, args.d2h, iters=args.iters, barrier=barrier)
    keys = ("both_s", "h2d_alone_s", "d2h_alone_s")
    t = torch.tensor([r.get(k, 0.0) for k in keys], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        both, up, down = (float(v) for v in t)
        it = args.iters
        line = {"n_gpus": world, "h2d_bytes_per_copy": args.h2d, "d2h_bytes_per_copy": args.d2h, "iters": it,
                "aggregate_both_GBps": world * (args.h2d + args.d2h) * it / both / 1e9,
                "aggregate_h2d_GBps_while_both": world * args.h2d * it / both / 1e9,
                "aggregate_d2h_GBps_while_both": world * args.d2h * it / both / 1e9,
                "aggregate_h2d_alone_GBps": world * args.h2d * it / up / 1e9,
                "aggregate_d2h_alone_GBps": world * args.d2h * it / down / 1e9 if down > 0 else None,
                "frames_per_s_ceiling": world * SUB * it / both * (args.h2d / H2D),
                "how": "plain copies, two upload + two download streams per rank (as the two feeder rings), best of 3 runs, no kernels; time = slowest rank (barrier before and after)"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ns", default="")
    ap.add_argument("--iters", type=int, default=60)
    ap.add_argument("--h2d", type=int, default=H2D)
    ap.add_argument("--d2h", type=int, default=D2H)
    ap.add_argument("--out", default="")
    ap.add_argument("--repeat", type=int, default=2)
    args = ap.parse_args()
    if not args.ns:
        one(args)
        return
    lines = []
    for n in (int(v) for v in args.ns.split(",")):
        for rep in range(args.repeat):
            cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1",
                   "--master-port", str(29610 + n), __file__, "--iters", str(args.iters), "--h2d", str(args.h2d), "--d2h", str(args.d2h)]
            p = subprocess.run(cmd, capture_output=True, text=True)
            got = [l for l in p.stdout.splitlines() if l.startswith("{")]
            if p.returncode != 0 or not got:
                print(p.stdout[-2000:], p.stderr[-2000:], file=sys.stderr)
                continue
            d = json.loads(got[-1])
            d["rep"] = rep
            lines.append(d)
            print(json.dumps(d), flush=True)
    if args.out:
        Path(args.out).write_text("\n".join(json.dumps(l) for l in lines) + "\n")


if __name__ == "__main__":
    main()
