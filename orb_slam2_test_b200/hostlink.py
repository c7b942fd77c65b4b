"""Raw host<->device copy ceiling of this box, measured with the sizes the host-fed extraction loop moves.

The host-fed rate of a frame-wise sharded sequence can never exceed what the box's
host<->device path delivers for plain copies.  `measure` runs exactly the traffic pattern of one rank's
feeder -- a page-locked -> device copy of one batch of frames on one CUDA stream, a device -> page-locked
copy of one batch of results on a second stream, both kept busy, NO kernels -- and returns GB/s.  Run
on every rank at once (bench.py does, tools/hostlink_ceiling.py does for N = 1/2/4/8) it gives the
aggregate ceiling the N-GPU host-fed number is to be judged against.

torch is used for the streams, events and page-locked buffers only.
"""
from __future__ import annotations

import time


def measure(device: int, h2d_bytes: int, d2h_bytes: int, iters: int = 40, warmup: int = 5, barrier=None, lanes: int = 2,
            repeats: int = 3) -> dict:
    """-> {"h2d_GBps", "d2h_GBps", "both_GBps": (h2d+d2h bytes)/s with both directions busy,
    "h2d_alone_GBps", "d2h_alone_GBps"} for THIS rank; the caller sums over ranks (after taking the max time).
    lanes = upload (and download) streams kept busy at once: the host-fed loop runs two feeder rings per GPU, each with
    its own upload and download stream, and one stream's copies alone do not fill the link on every box (a single copy
    engine per stream).  Every figure is the best of `repeats` runs: a ceiling is what the link CAN do."""
    import torch
    dev = torch.device("cuda", device)
    torch.cuda.set_device(dev)
    lanes = max(1, int(lanes))
    nbuf = 2 * lanes
    h_in = [torch.empty(h2d_bytes, dtype=torch.uint8).pin_memory() for _ in range(nbuf)]
    d_in = [torch.empty(h2d_bytes, dtype=torch.uint8, device=dev) for _ in range(nbuf)]
    d_out = [torch.empty(max(d2h_bytes, 1), dtype=torch.uint8, device=dev) for _ in range(nbuf)]
    h_out = [torch.empty(max(d2h_bytes, 1), dtype=torch.uint8).pin_memory() for _ in range(nbuf)]
    for t in h_in:
        t.fill_(7)
    ups = [torch.cuda.Stream(dev) for _ in range(lanes)]
    downs = [torch.cuda.Stream(dev) for _ in range(lanes)]

    def run_once(do_up: bool, do_down: bool, n: int) -> float:
        torch.cuda.synchronize(dev)
        if barrier is not None:
            barrier()
        t0 = time.perf_counter()
        for i in range(n):
            if do_up:
                with torch.cuda.stream(ups[i % lanes]):
                    d_in[i % nbuf].copy_(h_in[i % nbuf], non_blocking=True)
            if do_down and d2h_bytes:
                with torch.cuda.stream(downs[i % lanes]):
                    h_out[i % nbuf].copy_(d_out[i % nbuf], non_blocking=True)
        for st in ups + downs:
            st.synchronize()
        dt = time.perf_counter() - t0
        if barrier is not None:
            barrier()
        return dt

    def run(do_up: bool, do_down: bool, n: int) -> float:
        return min(run_once(do_up, do_down, n) for _ in range(max(1, repeats)))

    out = {}
    run_once(True, True, warmup)
    dt = run(True, True, iters)
    out["both_s"] = dt
    out["both_GBps"] = (h2d_bytes + d2h_bytes) * iters / dt / 1e9
    out["h2d_GBps"] = h2d_bytes * iters / dt / 1e9
    out["d2h_GBps"] = d2h_bytes * iters / dt / 1e9
    dt = run(True, False, iters)
    out["h2d_alone_s"] = dt
    out["h2d_alone_GBps"] = h2d_bytes * iters / dt / 1e9
    if d2h_bytes:
        dt = run(False, True, iters)
        out["d2h_alone_s"] = dt
        out["d2h_alone_GBps"] = d2h_bytes * iters / dt / 1e9
    out["iters"] = iters
    out["lanes"] = lanes
    out["repeats"] = repeats
    out["h2d_bytes"] = h2d_bytes
    out["d2h_bytes"] = d2h_bytes
    return out
