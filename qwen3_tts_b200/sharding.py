"""Utterance / stream sharding across the GPUs of one box.

The decode path has no exchange step: utterances are independent, weights are replicated, so the only
multi-GPU logic is a deterministic host-side partition (longest-processing-time-first) that every
rank computes identically, plus an optional host-side gather of results.  No data-path collective."""
from __future__ import annotations

from typing import List, Sequence


def frames_cost(num_frames: int, chunk: int = 300, left_ctx: int = 25) -> int:
    """Frames actually computed for an utterance, including the recomputed left context of every
    chunk after the first (reference chunked_decode, modeling :885-895)."""
    if num_frames <= 0:
        return 0
    nchunks = (num_frames + chunk - 1) // chunk
    return num_frames + left_ctx * (nchunks - 1)


def shard_utterances(lengths: Sequence[int], world_size: int, rank: int) -> List[int]:
    """Indices of the utterances rank `rank` decodes.  Greedy LPT: sort by cost descending (ties by
    index), give each to the currently least-loaded rank (ties by rank).  Deterministic."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad world_size / rank")
    order = sorted(range(len(lengths)), key=lambda i: (-frames_cost(int(lengths[i])), i))
    load = [0] * world_size
    mine: List[int] = []
    for i in order:
        r = min(range(world_size), key=lambda k: (load[k], k))
        load[r] += frames_cost(int(lengths[i]))
        if r == rank:
            mine.append(i)
    return sorted(mine)


def shard_loads(lengths: Sequence[int], world_size: int) -> List[int]:
    return [sum(frames_cost(int(lengths[i])) for i in shard_utterances(lengths, world_size, r)) for r in range(world_size)]


def bucket_by_length(indices: Sequence[int], lengths: Sequence[int], max_frames: int) -> List[List[int]]:
    """Group a rank's utterances into decode batches: sorted by length so padding is small, each batch
    at most `max_frames` padded frames (batch_size * longest)."""
    idx = sorted(indices, key=lambda i: (-int(lengths[i]), i))
    batches: List[List[int]] = []
    cur: List[int] = []
    longest = 0
    for i in idx:
        L = max(1, int(lengths[i]))
        new_longest = max(longest, L)
        if cur and new_longest * (len(cur) + 1) > max_frames:
            batches.append(cur)
            cur, new_longest = [], L
        cur.append(i)
        longest = new_longest
    if cur:
        batches.append(cur)
    return batches
