"""-m gpu test of the multi-GPU dispatcher: ShardedDecoder with one worker PROCESS per visible GPU (two when the box
has them, else one) must return, in input order, exactly what a direct Qwen3TTSTokenizer.decode of each utterance's batch
returns - the workers build the same seeded weights - and the ragged batches must follow the reference length rule."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_sharded_decoder_matches_direct_decode(gpu_model, cfg):
    from qwen3_tts_b200 import Qwen3TTSTokenizer, ShardedDecoder
    from qwen3_tts_b200.sharding import decode_shard, shard_utterances
    from qwen3_tts_b200.synthetic import make_codes
    ndev = min(2, torch.cuda.device_count())
    devices = [f"cuda:{i}" for i in range(ndev)]
    lengths = [40, 12, 33, 7, 25, 40, 18, 3, 29]
    codes = [make_codes(1, n, cfg, seed=300 + i)[0] for i, n in enumerate(lengths)]
    with ShardedDecoder(devices, {"seed": 0}, max_frames=96) as sd:
        wavs, sr = sd.decode_sharded(codes)
        pcm, _ = sd.decode_sharded(codes[:3], pcm16=True)
    assert sr == 24000 and len(wavs) == len(codes)
    # the same shards, same buckets, decoded in this process on the session's model (same seeded weights)
    tok = Qwen3TTSTokenizer.from_model(gpu_model)
    for r in range(ndev):
        idx = shard_utterances(lengths, ndev, r)
        dense = [codes[i] for i in idx]
        want = decode_shard(tok, dense, list(range(len(dense))), 96)
        for k, i in enumerate(idx):
            assert wavs[i].dtype == np.float32 and wavs[i].shape == want[k].shape
            np.testing.assert_array_equal(wavs[i], want[k])
    assert pcm[0].dtype == np.int16 and pcm[0].shape[0] > 0
