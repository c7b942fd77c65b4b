"""N>1 host logic on CPU: two gloo ranks shard a sequence frame-wise (stereo pairs kept together),
run their shards through a stand-in extract function, and rank 0 re-assembles the sequence in
order.  The stand-in is NOT an extractor (no pixels are processed on the CPU anywhere in the
product): it only makes results depend on the frame bytes so that mis-routing is detected."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from orb_slam2_test_b200 import sequence


def test_shard_indices_partition():
    for n in (0, 1, 7, 4541):
        for world in (1, 2, 4, 8):
            parts = [list(sequence.shard_indices(n, r, world)) for r in range(world)]
            assert sorted(sum(parts, [])) == list(range(n))
            assert max(map(len, parts)) - min(map(len, parts)) <= 1
            assert sequence.shard_sizes(n, world) == list(map(len, parts))
    with pytest.raises(ValueError):
        sequence.shard_indices(10, 2, 2)


def _load(f, e):
    rng = np.random.default_rng(f * 2 + e)
    return rng.integers(0, 256, (8, 12), dtype=np.uint8)


def _fake_extract(imgs):
    out = []
    for im in imgs:
        n = int(im[0, 0]) % 5
        k = np.zeros(n, np.dtype([("x", "<f4"), ("y", "<f4")]))
        k["x"] = im.reshape(-1)[:n]
        out.append((k, np.repeat(im[:1, :1], n * 32).reshape(n, 32)))
    return out


def _expected(n_frames, stereo):
    res = sequence.run_shard(range(n_frames), _load, _fake_extract, batch=4, stereo=stereo)
    return sequence.gather_to_rank0(res, 0, 1)


def _worker(rank, world, port, n_frames, stereo, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sequence.shard_indices(n_frames, rank, world)
        local = sequence.run_shard(mine, _load, _fake_extract, batch=6, stereo=stereo)
        assert all(r.frame % world == rank for r in local)
        allr = sequence.gather_to_rank0(local, rank, world)
        if rank == 0:
            sequence.check_complete(allr, n_frames, stereo)
            q.put([(r.frame, r.eye, r.n_keypoints, r.digest) for r in allr])
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("stereo", [False, True])
def test_two_ranks_reassemble_the_sequence(stereo):
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    n_frames, world = 23, 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_frames, stereo, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = [(r.frame, r.eye, r.n_keypoints, r.digest) for r in _expected(n_frames, stereo)]
    assert got == want


def test_incomplete_sequence_is_detected():
    res = _expected(5, True)
    with pytest.raises(RuntimeError):
        sequence.check_complete(res[:-1], 5, True)
