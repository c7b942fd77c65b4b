"""Frame-wise sharding of an image sequence over the GPUs of one node.

The extractor keeps no cross-frame state (the reference overwrites mvImagePyramid every call,
src/ORBextractor.cc:1417), so frame f of a sequence -- a stereo pair stays together, as
src/Frame.cc:110-113 extracts left and right of one Frame -- simply goes to rank f mod world.
Nothing is exchanged on the data path: no collective, no peer traffic.  Each rank pushes its own
frames through its own extractor handles (pinned host buffers, async copies) and only the small
per-frame result summaries travel to rank 0 for re-ordering, over whatever process group the
launcher set up (NCCL ranks use a gloo side group or torch's object collectives; tests use gloo).

`run_stream_shard` is the pipelined runner: it keeps the ring of an `OrbStream` (stream.py; the feeder
lives in the library, csrc/borb_stream.inl) full, so that uploads, kernels and downloads of successive
batches overlap on the rank's GPU.  `run_shard` is the synchronous helper the small tests use.
"""
from __future__ import annotations

import hashlib
from dataclasses import dataclass
from typing import Callable, Iterable

import numpy as np


def shard_indices(n_items: int, rank: int, world: int) -> range:
    """Round-robin: item f belongs to rank f mod world."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} of {world}")
    return range(rank, n_items, world)


def shard_sizes(n_items: int, world: int) -> list[int]:
    return [len(shard_indices(n_items, r, world)) for r in range(world)]


@dataclass
class FrameResult:
    """What comes back to the host for one image of the sequence."""
    frame: int          # sequence index
    eye: int            # 0 left / mono, 1 right
    n_keypoints: int
    digest: str         # SHA-256 over keypoint records + descriptors (order-sensitive)

    @staticmethod
    def of(frame: int, eye: int, kps: np.ndarray, desc: np.ndarray) -> "FrameResult":
        h = hashlib.sha256()
        h.update(np.ascontiguousarray(kps).view(np.uint8).tobytes())
        h.update(np.ascontiguousarray(desc).tobytes())
        return FrameResult(frame, eye, int(len(kps)), h.hexdigest())


def run_shard(frames: Iterable[int], load: Callable[[int, int], np.ndarray],
              extract_batch: Callable[[np.ndarray], list], batch: int, stereo: bool) -> list[FrameResult]:
    """Push this rank's frames through `extract_batch` ([n, rows, cols] u8 -> list of (kps, desc)),
    `batch` images per call; a stereo pair contributes two consecutive images of one call."""
    eyes = (0, 1) if stereo else (0,)
    todo = [(f, e) for f in frames for e in eyes]
    per_call = max(len(eyes), batch - batch % len(eyes))
    out: list[FrameResult] = []
    for i in range(0, len(todo), per_call):
        chunk = todo[i:i + per_call]
        imgs = np.stack([load(f, e) for f, e in chunk])
        res = extract_batch(imgs)
        assert len(res) == len(chunk)
        out.extend(FrameResult.of(f, e, k, d) for (f, e), (k, d) in zip(chunk, res))
    return out


def gather_to_rank0(local: list[FrameResult], rank: int, world: int, group=None) -> list[FrameResult] | None:
    """Collect every rank's summaries on rank 0, ordered by (frame, eye).  Control plane only."""
    if world == 1:
        return sorted(local, key=lambda r: (r.frame, r.eye))
    import torch.distributed as dist
    bucket = [None] * world if rank == 0 else None
    dist.gather_object(local, bucket, dst=0, group=group)
    if rank != 0:
        return None
    merged = [r for part in bucket for r in part]
    return sorted(merged, key=lambda r: (r.frame, r.eye))


def check_complete(results: list[FrameResult], n_frames: int, stereo: bool) -> None:
    want = [(f, e) for f in range(n_frames) for e in ((0, 1) if stereo else (0,))]
    got = [(r.frame, r.eye) for r in results]
    if got != want:
        missing = sorted(set(want) - set(got))[:5]
        dup = len(got) - len(set(got))
        raise RuntimeError(f"sequence incomplete: {len(got)} of {len(want)} results, missing {missing}, {dup} duplicates")


# ---------------------------------------------------------------- pipelined runner over an OrbStream

@dataclass
class ItemResult:
    """One item (frame, or stereo pair) of a sequence as it came back through a stream."""
    frame: int
    n_left: int
    n_right: int       # -1 for a monocular stream
    n_matched: int     # stereo association: entries with depth > 0; -1 for a monocular stream
    digest: str        # SHA-256 over everything that came back for the item ("" when not asked for)


def item_digest(stream, item: int) -> str:
    """Order-sensitive SHA-256 over an item's keypoint records and descriptors (both eyes) and, for a pair,
    mvuRight / mvDepth -- what the reference's stereo Frame constructor leaves in mvKeys, mDescriptors,
    mvKeysRight, mDescriptorsRight, mvuRight, mvDepth (src/Frame.cc:79-151)."""
    h = hashlib.sha256()
    for e in range(stream.eyes):
        k, d = stream.result(item, e, copy=False)
        h.update(k.view(np.uint8).tobytes() if len(k) else b"")
        h.update(d.tobytes())
    if stream.eyes == 2:
        u, z, _ = stream.stereo_result(item, copy=False)
        h.update(u.tobytes())
        h.update(z.tobytes())
    return h.hexdigest()


def run_stream_shard(stream, frames, left, right=None, fill=None, digest: bool = True) -> list[ItemResult]:
    """Push this rank's items through `stream`, keeping its ring full.

    frames: the sequence indices of this rank's items, in order.  Input, one of
      left / right   [len(frames), rows, cols] uint8 arrays holding the items in that order -- page-locked
                     memory (extractor.pinned_empty) is uploaded from where it lies, with no staging copy;
      fill(f, l, r)  writes item `f` into the page-locked slot views l, r ([rows, cols]; r None for mono).
    Returns one ItemResult per item, in push order."""
    frames = list(frames)
    n, B = len(frames), stream.batch
    out: list[ItemResult] = []
    fsz = stream.rows * stream.cols
    i = 0
    while i < n or stream.in_flight():
        while i < n and stream.slots_free() > 0:
            m = min(B, n - i)
            if fill is not None:
                lb = stream.input_buffer(0)
                rb = stream.input_buffer(1) if stream.eyes == 2 else None
                for j in range(m):
                    fill(frames[i + j], lb[j], rb[j] if rb is not None else None)
                stream.push(None, n=m, tag=i)
            else:
                stream.push_ptr(left.ctypes.data + i * fsz, right.ctypes.data + i * fsz if right is not None else None, m, tag=i)
            i += m
        tag, m = stream.pop()
        for j in range(m):
            if stream.eyes == 2:
                nm = stream.stereo_result(j, copy=False)[2]
                out.append(ItemResult(frames[tag + j], stream.count(j, 0), stream.count(j, 1), nm,
                                      item_digest(stream, j) if digest else ""))
            else:
                out.append(ItemResult(frames[tag + j], stream.count(j, 0), -1, -1, item_digest(stream, j) if digest else ""))
        stream.release()
    return out


def gather_items_to_rank0(local: list[ItemResult], rank: int, world: int, group=None) -> list[ItemResult] | None:
    """Every rank's item summaries on rank 0, ordered by sequence index.  Control plane only (a few dozen
    bytes per item); the data path has no collective."""
    if world == 1:
        return sorted(local, key=lambda r: r.frame)
    import torch.distributed as dist
    bucket = [None] * world if rank == 0 else None
    dist.gather_object(local, bucket, dst=0, group=group)
    if rank != 0:
        return None
    return sorted((r for part in bucket for r in part), key=lambda r: r.frame)


def sequence_digest(results: list[ItemResult], n_frames: int) -> str:
    """Digest of digests over the whole sequence in frame order; raises when an item is missing or doubled."""
    got = [r.frame for r in results]
    if got != list(range(n_frames)):
        missing = sorted(set(range(n_frames)) - set(got))[:5]
        raise RuntimeError(f"sequence incomplete: {len(got)} of {n_frames} items, missing {missing}, "
                           f"{len(got) - len(set(got))} duplicates")
    h = hashlib.sha256()
    for r in results:
        h.update(bytes.fromhex(r.digest))
    return h.hexdigest()
