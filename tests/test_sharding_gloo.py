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


# ---- the pipelined runner (sequence.run_stream_shard) over a stand-in for OrbStream -------------------------------

class _FakeStream:
    """The ring bookkeeping of OrbStream (push / pop / release / slots_free) over the stand-in extract function;
    no pixels are processed the way the product does -- only routing and ordering are under test here."""

    def __init__(self, batch, slots, rows=8, cols=12):
        self.batch, self.slots, self.rows, self.cols, self.eyes = batch, slots, rows, cols, 2
        self.q, self.held = [], None

    def slots_free(self):
        return self.slots - len(self.q) - (self.held is not None)

    def in_flight(self):
        return len(self.q)

    def push_ptr(self, lp, rp, n, tag=0):
        import ctypes as C
        assert self.slots_free() > 0
        def arr(p):
            buf = (C.c_uint8 * (n * self.rows * self.cols)).from_address(p)
            return np.frombuffer(buf, np.uint8).reshape(n, self.rows, self.cols).copy()
        self.q.append((tag, n, _fake_extract(arr(lp)), _fake_extract(arr(rp))))

    def pop(self):
        self.held = self.q.pop(0)
        return self.held[0], self.held[1]

    def release(self):
        self.held = None

    def result(self, item, eye=0, copy=True):
        return self.held[2 + eye][item]

    def count(self, item, eye=0):
        return len(self.held[2 + eye][item][0])

    def stereo_result(self, item, copy=True):
        n = self.count(item, 0)
        return np.full(n, 1.5, np.float32), np.full(n, 2.5, np.float32), n // 2


def _stream_worker(rank, world, port, n_frames, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = list(sequence.shard_indices(n_frames, rank, world))
        left = np.stack([_load(f, 0) for f in mine])
        right = np.stack([_load(f, 1) for f in mine])
        res = sequence.run_stream_shard(_FakeStream(batch=3, slots=2), mine, left, right)
        assert [r.frame for r in res] == mine
        merged = sequence.gather_items_to_rank0(res, rank, world)
        if rank == 0:
            q.put(sequence.sequence_digest(merged, n_frames))
        dist.barrier()
    finally:
        dist.destroy_process_group()


def test_two_ranks_stream_runner_gives_the_single_rank_digest():
    n_frames, world = 29, 2
    left = np.stack([_load(f, 0) for f in range(n_frames)])
    right = np.stack([_load(f, 1) for f in range(n_frames)])
    one = sequence.run_stream_shard(_FakeStream(batch=4, slots=3), range(n_frames), left, right)
    want = sequence.sequence_digest(sequence.gather_items_to_rank0(one, 0, 1), n_frames)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_stream_worker, args=(r, world, port, n_frames, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == want
    with pytest.raises(RuntimeError):
        sequence.sequence_digest(one[1:], n_frames)


def test_kseq_golden_file_is_complete():
    import json
    from pathlib import Path
    g = json.loads((Path(__file__).parent / "golden" / "kseq_digests.json").read_text())
    assert g["pairs"] == 4541 and len(g["first_64"]) == 64 and set(g["prefix"]) == {"16", "64", "256", "1024", "4541"}
    import hashlib
    h = hashlib.sha256()
    for d in g["first_64"]:
        h.update(bytes.fromhex(d))
    assert h.hexdigest() == g["prefix"]["64"]
