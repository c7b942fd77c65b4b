"""Frame-wise sharding of an image sequence over the GPUs of one node.

The extractor keeps no cross-frame state (the reference overwrites mvImagePyramid every call,
src/ORBextractor.cc:1417), so frame f of a sequence -- a stereo pair stays together, as
src/Frame.cc:110-113 extracts left and right of one Frame -- simply goes to rank f mod world.
Nothing is exchanged on the data path: no collective, no peer traffic.  Each rank pushes its own
frames through its own extractor handles (pinned host buffers, async copies) and only the small
per-frame result summaries travel to rank 0 for re-ordering, over whatever process group the
launcher set up (NCCL ranks use a gloo side group or torch's object collectives; tests use gloo).
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
