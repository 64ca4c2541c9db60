"""-m gpu test of the multi-GPU dispatcher: ShardedDecoder with one worker PROCESS per visible GPU (two when the box
has them, else one) must return, in input order, exactly what a direct Qwen3TTSTokenizer.decode of each utterance's batch
returns - the workers build the same seeded weights - and the ragged batches must follow the reference length rule."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_sharded_decoder_matches_direct_decode(gpu_model, cfg):
    from qwen3_tts_b200 import Qwen3TTSTokenizer, ShardedDecoder
    from qwen3_tts_b200.sharding import decode_shard, global_batches
    from qwen3_tts_b200.synthetic import make_codes
    ndev = min(2, torch.cuda.device_count())
    devices = [f"cuda:{i}" for i in range(ndev)]
    lengths = [40, 12, 33, 7, 25, 40, 18, 3, 29]
    codes = [make_codes(1, n, cfg, seed=300 + i)[0] for i, n in enumerate(lengths)]
    with ShardedDecoder(devices, {"seed": 0}, max_frames=96) as sd:
        wavs, sr = sd.decode_sharded(codes)
        pcm, _ = sd.decode_sharded(codes[:3], pcm16=True)
    assert sr == 24000 and len(wavs) == len(codes)
    # the same global buckets, dealt the same way, decoded in this process on the session's model (same seeded weights); an item's
    # waveform depends on the padded length of its batch (length rule, modeling :1012-1015), so the batches must be the same ones
    tok = Qwen3TTSTokenizer.from_model(gpu_model)
    batches = global_batches(lengths, ndev, 96)
    assert sorted(i for b in batches for i in b) == list(range(len(codes)))
    want = decode_shard(tok, codes, [], 96, batches=batches)
    for i, w in want.items():
        assert wavs[i].dtype == np.float32 and wavs[i].shape == w.shape
        np.testing.assert_array_equal(wavs[i], w)
    assert pcm[0].dtype == np.int16 and pcm[0].shape[0] > 0


def test_sharded_encode_matches_direct_encode():
    """encode_sharded on one / two GPUs returns, in input order, what model.encode_host returns for the same global buckets."""
    from qwen3_tts_b200 import ShardedDecoder, synthetic_model
    from qwen3_tts_b200.sharding import global_batches
    from qwen3_tts_b200.synthetic import make_audio
    ndev = min(2, torch.cuda.device_count())
    lengths = [30000, 4000, 52000, 24000, 9000, 52000, 1]
    wav = make_audio(len(lengths), max(lengths), seed=21, lengths=lengths)
    wavs = [wav[b, :n].numpy() for b, n in enumerate(lengths)]
    with ShardedDecoder([f"cuda:{i}" for i in range(ndev)], {"seed": 0, "with_encoder": True}) as sd:
        got = sd.encode_sharded(wavs, max_samples=120000)
    model = synthetic_model(seed=0, device="cuda:0", with_encoder=True)
    frames = [-(-n // 1920) for n in lengths]
    batches = global_batches(frames, ndev, 120000 // 1920)
    assert sorted(i for b in batches for i in b) == list(range(len(lengths)))
    for b in batches:
        want = model.encode_host([wavs[i] for i in b])
        for i, w in zip(b, want):
            assert got[i].dtype == np.int64 and got[i].shape == (frames[i], 16)
            np.testing.assert_array_equal(got[i], w)
