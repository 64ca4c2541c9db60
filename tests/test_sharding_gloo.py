"""N > 1 host logic on CPU: world_size-2 gloo processes compute their shards independently and
check (through a host-side gather, the only communication the path has) that the shards are
disjoint, cover every utterance and are balanced."""
import os
import random

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from qwen3_tts_b200.sharding import bucket_by_length, frames_cost, shard_loads, shard_utterances


def _worker(rank, world, port, lengths, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    mine = shard_utterances(lengths, world, rank)
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    load = torch.tensor([float(sum(frames_cost(lengths[i]) for i in mine))])
    mx = load.clone(); dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = load.clone(); dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    if rank == 0:
        q.put((gathered, mx.item(), sm.item()))
    dist.destroy_process_group()


def test_two_rank_shards_cover_and_balance():
    rng = random.Random(0)
    lengths = [rng.randint(20, 750) for _ in range(101)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + rng.randint(0, 2000)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lengths, q)) for r in range(2)]
    for p in procs:
        p.start()
    gathered, mx, sm = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    flat = sorted(i for g in gathered for i in g)
    assert flat == list(range(len(lengths)))          # disjoint cover
    assert mx <= sm / 2 * 1.02                        # LPT keeps the max load within 2 % of the mean


def test_sharding_is_deterministic_and_single_rank_is_identity():
    lengths = [375] * 16 + [10, 900, 301]
    assert shard_utterances(lengths, 1, 0) == list(range(len(lengths)))
    for w in (2, 4, 8):
        shards = [shard_utterances(lengths, w, r) for r in range(w)]
        assert sorted(i for s in shards for i in s) == list(range(len(lengths)))
        assert shards == [shard_utterances(lengths, w, r) for r in range(w)]
        assert max(shard_loads(lengths, w)) - min(shard_loads(lengths, w)) <= frames_cost(900)


def test_frames_cost_counts_left_context_recompute():
    assert frames_cost(125) == 125
    assert frames_cost(375) == 400      # 300 + (25 + 75)
    assert frames_cost(750) == 800      # 300 + 325 + 175 (cfg 4)
    assert frames_cost(0) == 0


def test_buckets_respect_frame_budget():
    lengths = [750, 700, 300, 120, 100, 90, 10]
    batches = bucket_by_length(range(len(lengths)), lengths, max_frames=1500)
    assert sorted(i for b in batches for i in b) == list(range(len(lengths)))
    for b in batches:
        assert len(b) == 1 or len(b) * max(lengths[i] for i in b) <= 1500
