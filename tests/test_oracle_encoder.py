"""CPU: the encode-path oracle (oracle/encoder_oracle.py: SEANet conv stack, 8-layer windowed transformer, replicate-padded
down-sampling conv, RVQ search) against tests/golden/encode_B2.npz, which the reference's own Qwen3TTSTokenizerV2Encoder class wrote
(oracle/make_golden_encoder.py), and against that class run live when /root/reference (or oracle/_ref) is present."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import encoder_oracle as XO
from oracle.make_golden_encoder import LENGTHS, golden_inputs
from qwen3_tts_b200.synthetic import make_audio, make_encoder_state_dict


def test_frame_count_rule():
    # every strided conv pads its input up to a whole frame (modeling_mimi.py:273-285): frames = ceil-chain over 4, 5, 6, 8, 2
    assert [XO.num_frames(n) for n in (1, 1920, 1921, 24000, 264333, 48001)] == [1, 1, 2, 13, 138, 26]


def test_encoder_oracle_matches_reference_golden():
    g = load_golden("encode_B2.npz")
    assert tuple(g["lengths"]) == LENGTHS
    sd = make_encoder_state_dict(seed=0)
    wav, mask = golden_inputs()
    codes, lat, margin = XO.model_encode(sd, wav, mask)
    want_lat = g["latents"]
    assert lat.shape == want_lat.shape == (2, 512, 138)
    scale = float(np.abs(want_lat).max())
    assert np.abs(lat - want_lat).max() <= 2e-5 * scale            # float64 restatement vs the reference's fp32: fp32 rounding only
    with torch.inference_mode():
        x = XO.seanet_encoder(sd, wav.double().unsqueeze(1)).numpy()[:, ::8, ::4]
    assert np.abs(x - g["seanet"]).max() <= 2e-5 * float(np.abs(g["seanet"]).max())
    for b, want in enumerate((g["codes0"], g["codes1"])):
        assert codes[b].shape == want.shape == (-(-LENGTHS[b] // 1920), 16)
        compared, wrong = XO.decidable(codes[b].T, want.astype(np.int64).T, margin[b][:, :want.shape[0]], 1e-5)
        assert wrong == 0 and compared >= 0.97 * want.size, (b, compared, wrong)
        assert (codes[b] == want).mean() > 0.99
    assert len(np.unique(codes[0])) > 300                            # a real search over the codebooks, not a degenerate fixture


def test_sliding_window_and_replicate_padding_matter():
    """The fixture must exercise what it claims: an 11 s item reaches past the 250-frame attention window, and the down-sampling
    conv's replicate padding differs from zero padding in the first frame."""
    sd = make_encoder_state_dict(seed=0)
    wav, _ = golden_inputs()
    with torch.inference_mode():
        x = XO.seanet_encoder(sd, wav[:1].double().unsqueeze(1)).transpose(1, 2)
        assert x.shape[1] > 250
        y = XO.encoder_transformer(sd, x)
        y_nowin = XO.encoder_transformer(sd, x, sliding_window=100000)
        assert torch.equal(y[:, :250], y_nowin[:, :250]) and (y[:, 250:] - y_nowin[:, 250:]).abs().max() > 1e-6
        w = sd["encoder.downsample.conv.weight"].double()
        a = XO.causal_conv1d(y.transpose(1, 2), w, None, stride=2, pad_mode="replicate")
        b = XO.causal_conv1d(y.transpose(1, 2), w, None, stride=2, pad_mode="constant")
        assert (a[..., 0] - b[..., 0]).abs().max() > 1e-3 and torch.equal(a[..., 2:-1], b[..., 2:-1])


def test_encoder_oracle_matches_reference_live():
    from oracle import ref_shim
    if not ref_shim.available():
        pytest.skip("reference sources not present")
    sd = make_encoder_state_dict(seed=5)
    lengths = (24000 + 7, 30000)
    wav = make_audio(2, max(lengths), seed=5, lengths=lengths)
    mask = torch.zeros_like(wav, dtype=torch.long)
    for b, n in enumerate(lengths):
        mask[b, :n] = 1
    want = ref_shim.reference_model_encode(ref_shim.build_reference_encoder(sd), wav, mask)
    codes, _lat, margin = XO.model_encode(sd, wav, mask)
    for b in range(2):
        w = want[b].numpy()
        assert codes[b].shape == w.shape
        compared, wrong = XO.decidable(codes[b].T, w.T, margin[b][:, :w.shape[0]], 1e-5)
        assert wrong == 0 and compared >= 0.97 * w.size
