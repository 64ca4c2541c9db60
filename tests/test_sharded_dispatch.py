"""CPU test of the multi-GPU dispatcher's host logic (qwen3_tts_b200/sharding.py ShardedDecoder): two worker PROCESSES with an
injected fake decoder (no GPU, no CUDA library call): partition, per-worker batches, gather in input order, error propagation."""
import numpy as np
import pytest
import torch

from qwen3_tts_b200.sharding import ShardedDecoder, decode_shard, plan_batches, shard_utterances


def _fake_worker(rank, device, spec, jobs, results):
    results.put(("ready", rank, None))
    while True:
        job = jobs.get()
        if job is None:
            return
        job_id, indices, codes, max_frames, pcm16 = job
        if spec.get("fail_on") == rank:
            results.put(("error", rank, "boom"))
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
            for r in (0, 1):
                assert sorted(i for i, o in enumerate(owners) if o == r) == shard_utterances(lengths, 2, r)
    plans = plan_batches(lengths, 2, 64)
    assert sorted(i for rank in plans for b in rank for i in b) == list(range(len(lengths)))
    assert all(max(lengths[i] for i in b) * len(b) <= 64 or len(b) == 1 for rank in plans for b in rank)


def test_dispatcher_propagates_worker_errors():
    codes = [torch.ones((5, 16), dtype=torch.long) for _ in range(4)]
    with ShardedDecoder(["cpu:0", "cpu:1"], {"up": 2, "fail_on": 1}, worker=_fake_worker) as sd:
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
