"""CPU test of the multi-GPU dispatcher's host logic (qwen3_tts_b200/sharding.py ShardedDecoder): two worker PROCESSES with an
injected fake decoder (no GPU, no CUDA library call): partition, per-worker batches, gather in input order, error propagation."""
import numpy as np
import pytest
import torch

from qwen3_tts_b200.sharding import ShardedDecoder, batch_cost, decode_shard, global_batches, plan_batches, shard_batches, shard_utterances


def _fake_worker(rank, device, spec, jobs, results):
    results.put(("ready", rank, None))
    while True:
        job = jobs.get()
        if job is None:
            return
        job_id, indices, codes, pcm16 = job
        if spec.get("fail_on") is not None and len(indices) and spec["fail_on"] in indices:
            results.put(("error", rank, "boom"))
            continue
        if pcm16 == "encode":   # "codes" of a waveform: one row per 1920 samples holding the frame's first sample, tagged with the rank
            items = [(i, torch.stack([w[::1920].to(torch.float32) + 1000.0 * rank] * 16, dim=1).to(torch.int64)) for i, w in zip(indices, codes)]
            results.put(("done", rank, (job_id, items)))
            continue
        # "waveform" of an utterance: its first codebook column repeated `up` times, tagged with the rank that decoded it
        items = [(i, torch.repeat_interleave(c[:, 0].to(torch.float32), spec["up"]) + 1000.0 * rank) for i, c in zip(indices, codes)]
        results.put(("done", rank, (job_id, items)))


def test_dispatcher_gathers_in_input_order_and_balances():
    rng = np.random.default_rng(0)
    lengths = [int(v) for v in rng.integers(3, 40, 23)]
    codes = [torch.from_numpy(rng.integers(1, 100, (n, 16))) for n in lengths]
    with ShardedDecoder(["cpu:0", "cpu:1"], {"up": 3}, max_frames=64, worker=_fake_worker) as sd:
        for _ in range(2):                                   # two jobs on the same pool
            wavs, sr = sd.decode_sharded(codes)
            assert sr == 24000 and len(wavs) == len(codes)
            owners = []
            for i, (w, c) in enumerate(zip(wavs, codes)):
                rank = int(round(float(w[0] - c[0, 0]) / 1000.0))
                owners.append(rank)
                np.testing.assert_array_equal(w - 1000.0 * rank, np.repeat(c[:, 0].numpy().astype(np.float32), 3))
            # pull-based: which worker took which batch is not fixed, but every batch was decoded whole by ONE worker
            for b in global_batches(lengths, 2, 64):
                assert len({owners[i] for i in b}) == 1
            assert set(owners) <= {0, 1}
    plans = plan_batches(lengths, 2, 64)
    assert sorted(i for rank in plans for b in rank for i in b) == list(range(len(lengths)))
    assert all(max(lengths[i] for i in b) * len(b) <= 64 or len(b) == 1 for rank in plans for b in rank)


def test_dispatcher_propagates_worker_errors():
    codes = [torch.ones((5, 16), dtype=torch.long) for _ in range(4)]
    with ShardedDecoder(["cpu:0", "cpu:1"], {"up": 2, "fail_on": 3}, worker=_fake_worker) as sd:
        with pytest.raises(RuntimeError, match="boom"):
            sd.decode_sharded(codes)


def test_decode_shard_buckets_through_the_tokenizer_interface():
    class _Tok:
        calls = []

        def decode(self, payload):
            self.calls.append(len(payload))
            return [np.full(int(p["audio_codes"].shape[0]), float(p["audio_codes"][0, 0])) for p in payload], 24000
    tok = _Tok()
    codes = [torch.full((n, 16), n) for n in (10, 3, 7, 10, 2)]
    out = decode_shard(tok, codes, [0, 2, 3, 4], max_frames=20)
    assert sorted(out) == [0, 2, 3, 4] and all(out[i].shape == (codes[i].shape[0],) and out[i][0] == codes[i].shape[0] for i in out)
    assert tok.calls == [2, 2]                               # (10, 10) then (7, 2): each at most 20 padded frames


def test_shard_batches_padding_is_independent_of_world_size():
    """Global buckets first, then deal: the padded frames decoded are (almost) the same for every N and the deal balances."""
    rng = np.random.default_rng(5)
    lengths = [int(v) for v in rng.integers(187, 376, 4096)]
    real = sum(lengths)
    waste = {}
    for world in (1, 2, 4, 8):
        plan = shard_batches(lengths, world, 64 * 375)
        assert sorted(i for r in plan for b in r for i in b) == list(range(4096))
        padded = [sum(len(b) * max(lengths[i] for i in b) for b in r) for r in plan]
        cost = [sum(batch_cost(b, lengths) for b in r) for r in plan]
        waste[world] = sum(padded) / real - 1.0
        assert max(cost) / (sum(cost) / world) < 1.04, (world, cost)
        assert all(len(r) >= min(8, 52 // world) for r in plan)
    assert waste[8] < 0.03 and waste[8] <= waste[1] + 0.005, waste
    # the old scheme (shard first, bucket per rank) for comparison: its padding grows with N
    old8 = plan_batches(lengths, 8, 64 * 375)
    old_waste = sum(len(b) * max(lengths[i] for i in b) for r in old8 for b in r) / real - 1.0
    assert old_waste > waste[8]


def test_dispatcher_encode_side_gathers_in_input_order():
    """encode_sharded: the same pull-based dispatch for waveforms -> codes (bucketed by code frames, results in input order)."""
    rng = np.random.default_rng(5)
    lengths = [int(v) for v in rng.integers(1, 30000, 17)]
    wavs = [np.arange(n, dtype=np.float32) % 97 for n in lengths]
    with ShardedDecoder(["cpu:0", "cpu:1"], {"up": 1}, worker=_fake_worker) as sd:
        codes = sd.encode_sharded(wavs, max_samples=60000)
        assert len(codes) == len(wavs)
        for c, w in zip(codes, wavs):
            assert c.dtype == np.int64 and c.shape == (-(-len(w) // 1920), 16)
            np.testing.assert_array_equal(c[:, 0] % 1000, w[::1920].astype(np.int64))
        wav2, sr = sd.decode_sharded([torch.ones((4, 16), dtype=torch.long)])          # the pool still serves decode jobs
        assert sr == 24000 and wav2[0].shape == (4,)
