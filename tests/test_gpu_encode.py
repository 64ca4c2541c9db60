"""-m gpu: the whole 12 Hz encode path through the C-ABI: q3tts_encode / q3tts_encode_host behind
Qwen3TTSTokenizerV2Model.encode and Qwen3TTSTokenizer.encode, against the golden the reference's own encoder class wrote
(tests/golden/encode_B2.npz) and the float64 oracle (oracle/encoder_oracle.py).

Bars: latents within 1e-4 of the largest latent (the reference's own fp32 run sits at 7e-7 from the float64 oracle; this path
multiplies fp16 hi + lo pairs on the tensor cores, ~2e-6); codes identical wherever the oracle's two best codes are further apart
than that error can bridge (relative margin >= 1e-4), and then only up to a frame's first near-tie that went the other way."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import encoder_oracle as XO
from oracle.make_golden_encoder import LENGTHS, golden_inputs
from qwen3_tts_b200.synthetic import make_audio, make_encoder_state_dict

pytestmark = pytest.mark.gpu
LATENT_TOL = 1e-4
MARGIN_TOL = 1e-4


@pytest.fixture(scope="module")
def enc_model(sd):
    from qwen3_tts_b200 import synthetic_model
    return synthetic_model(state_dict=sd, device="cuda:0", with_encoder=True)


def _check_codes(got, want, margin, what):
    """got / want [T, K]; margin [K, T]"""
    assert got.shape == want.shape, (what, got.shape, want.shape)
    compared, wrong = XO.decidable(got.T, want.T, margin[:, :want.shape[0]], MARGIN_TOL)
    same = float((got == want).mean())
    print(f"{what}: {compared} decidable codes, {wrong} wrong, {same:.4f} of all codes identical")
    assert wrong == 0 and compared >= 0.8 * want.size and same >= 0.9, (what, compared, wrong, same)   # few-frame items: one near-tie is a large share


def test_encode_matches_reference_golden(enc_model):
    g = load_golden("encode_B2.npz")
    wav, mask = golden_inputs()
    codes, lat = enc_model.encode(wav, mask, return_latents=True)
    lat = lat.cpu().numpy()
    want_lat = g["latents"]
    assert lat.shape == want_lat.shape
    err = float(np.abs(lat - want_lat).max() / np.abs(want_lat).max())
    print(f"latents vs the reference's fp32 run: max-abs error {err:.2e} of the largest latent")
    assert err <= LATENT_TOL
    sd = make_encoder_state_dict(seed=0)
    _o, _l, margin = XO.model_encode(sd, wav, mask)
    for b, want in enumerate((g["codes0"], g["codes1"])):
        assert codes[b].shape == (-(-LENGTHS[b] // 1920), 16) and codes[b].dtype == torch.int64
        _check_codes(codes[b].cpu().numpy(), want.astype(np.int64), margin[b], f"golden item {b}")


def test_encode_output_forms_and_host_entry(enc_model):
    """Qwen3TTSTokenizer.encode (wrapper :208-257) and the host entry give the same codes as model.encode; return forms as the reference."""
    from qwen3_tts_b200 import Qwen3TTSTokenizer, Qwen3TTSTokenizerV2EncoderOutput
    tok = Qwen3TTSTokenizer.from_model(enc_model)
    lengths = (30000, 24000 + 77, 1)
    wav = make_audio(3, max(lengths), seed=3, lengths=lengths)
    arrays = [wav[b, :n].numpy() for b, n in enumerate(lengths)]
    out = tok.encode(arrays, sr=24000)
    assert isinstance(out, Qwen3TTSTokenizerV2EncoderOutput) and len(out.audio_codes) == 3
    assert [tuple(c.shape) for c in out.audio_codes] == [(16, 16), (13, 16), (1, 16)]
    tup = tok.encode(arrays, sr=24000, return_dict=False)
    assert isinstance(tup, tuple) and all(torch.equal(a, b) for a, b in zip(tup[0], out.audio_codes))
    host = enc_model.encode_host(arrays)
    assert all(np.array_equal(h, c.cpu().numpy()) for h, c in zip(host, out.audio_codes))
    one = tok.encode(arrays[1], sr=24000)                                   # a single ndarray is a batch of one
    # batch of one: the same item without the batch's zero padding behind it.  Everything is causal, so every frame that ends inside
    # the real samples is unchanged; the last, partial frame sees the conv stack's own right padding instead of padded samples pushed
    # through the biases - in the reference too (it encodes the padded batch as it is, modeling :979-980)
    assert one.audio_codes[0].shape == out.audio_codes[1].shape and torch.equal(one.audio_codes[0][:-1], out.audio_codes[1][:-1])
    with pytest.raises(ValueError):
        tok.encode(arrays)                                                  # numpy input needs sr (wrapper :195-196)
    with pytest.raises(TypeError):
        tok.encode([arrays[0], "x.wav"], sr=24000)                          # mixed input types (wrapper :199-200)
    # encode -> decode round trip runs end to end (different random codebooks either side: no identity expected)
    wavs, sr = tok.decode(out)
    assert sr == 24000 and [w.shape[0] for w in wavs] == [min(int((c[:, 0] > 0).sum()) * 1920, tok.num_decoded_samples(16)) for c in out.audio_codes]


def test_encode_matches_oracle_on_a_ragged_batch_at_size(enc_model):
    """12 items of 0.5 .. 14 s in one call (sub-batched by the engine), against the float64 oracle item by item.  The oracle gets
    the item as the batch holds it - zero-padded to the longest item - because the reference encodes the padded batch as it is
    (modeling :979-980): an item's last, partial frame sees the padding."""
    rng = np.random.default_rng(11)
    lengths = [int(x) for x in rng.integers(12000, 336000, size=12)]
    lengths[3] = 24000 * 14                                                  # 350 frames at 25 Hz: past the attention window
    wav = make_audio(12, max(lengths), seed=11, lengths=lengths)
    mask = torch.zeros_like(wav, dtype=torch.long)
    for b, n in enumerate(lengths):
        mask[b, :n] = 1
    codes = enc_model.encode(wav, mask).audio_codes
    sd = make_encoder_state_dict(seed=0)
    for b in (3, 0, 7):
        n = lengths[b]
        want, _lat, margin = XO.model_encode(sd, wav[b:b + 1], mask[b:b + 1])
        _check_codes(codes[b].cpu().numpy(), want[0], margin[0], f"ragged item {b} ({n} samples)")


@pytest.mark.parametrize("samples", [1, 7, 959, 1919, 1920, 1921, 3841, 24000 * 40 + 13])
def test_encode_edge_lengths_match_oracle(enc_model, samples):
    """Lengths around every rounding-up boundary of the conv stack (strides 4, 5, 6, 8, 2: one 12.5 Hz frame = 1920 samples), down to
    a single sample (every stage is then one row and the few-row tile packing of the GEMM kernels is what runs), and a 40 s item
    (1000 frames at 25 Hz: the attention window slides four times).  Batch of two equal items: both rows must give the same codes."""
    wav = make_audio(1, samples, seed=samples % 97).repeat(2, 1)
    mask = torch.ones_like(wav, dtype=torch.long)
    codes, lat = enc_model.encode(wav, mask, return_latents=True)
    assert enc_model.encode_num_frames(samples) == XO.num_frames(samples) == lat.shape[2]
    assert codes[0].shape == (-(-samples // 1920), 16) and torch.equal(codes[0], codes[1]) and torch.equal(lat[0], lat[1])
    sd = make_encoder_state_dict(seed=0)
    want, lat64, margin = XO.model_encode(sd, wav[:1], mask[:1])
    err = float(np.abs(lat[:1].cpu().numpy() - lat64).max() / max(np.abs(lat64).max(), 1e-6))
    print(f"{samples} samples: latents rel err {err:.2e}")
    assert err <= LATENT_TOL
    _check_codes(codes[0].cpu().numpy(), want[0], margin[0], f"{samples} samples")


def test_encoder_config_reaches_the_kernels(sd):
    """encoder_config of the model config (MimiConfig fields) is what the encode path runs with: a 60-frame attention window must
    reproduce the oracle run with that window, and differ from the default 250-frame one on an item longer than the window."""
    from qwen3_tts_b200 import Qwen3TTSTokenizerV2Config, Qwen3TTSTokenizerV2Model
    esd = make_encoder_state_dict(seed=0)
    cfg = Qwen3TTSTokenizerV2Config(encoder_config={"sliding_window": 60})
    model = Qwen3TTSTokenizerV2Model(cfg, device="cuda:0")
    model.decoder.load_state_dict({**sd, **esd})
    wav = make_audio(1, 24000 * 6, seed=9)
    _codes, lat = model.encode(wav, return_latents=True)
    with torch.inference_mode():
        x = XO.seanet_encoder(esd, wav.double().unsqueeze(1)).transpose(1, 2)
        w = esd["encoder.downsample.conv.weight"].double()
        want60 = XO.causal_conv1d(XO.encoder_transformer(esd, x, sliding_window=60).transpose(1, 2), w, None, stride=2, pad_mode="replicate").numpy()
        want250 = XO.causal_conv1d(XO.encoder_transformer(esd, x, sliding_window=250).transpose(1, 2), w, None, stride=2, pad_mode="replicate").numpy()
    got = lat.cpu().numpy()
    scale = np.abs(want60).max()
    assert np.abs(got - want60).max() <= LATENT_TOL * scale and np.abs(got - want250).max() > 1e-3 * scale


def test_encode_is_independent_of_the_decode_compute_mode(sd, enc_model):
    """The encoder always multiplies fp16 hi + lo pairs, whatever operand type the handle's DECODE path was built with: a bf16-mode
    model must return the very same codes and latents."""
    from qwen3_tts_b200 import synthetic_model
    wav = make_audio(2, 30000, seed=4)
    a_codes, a_lat = enc_model.encode(wav, return_latents=True)
    other = synthetic_model(state_dict=sd, device="cuda:0", compute_mode="bf16", with_encoder=True)
    b_codes, b_lat = other.encode(wav, return_latents=True)
    assert torch.equal(a_lat, b_lat) and all(torch.equal(x, y) for x, y in zip(a_codes, b_codes))
    many = make_audio(300, 1921, seed=5)                       # many short items in one call: 2 frames each
    out = enc_model.encode(many).audio_codes
    assert len(out) == 300 and all(c.shape == (2, 16) for c in out)
    assert torch.equal(out[7], enc_model.encode(many[7:8]).audio_codes[0])


def test_encode_long_item_and_sub_batching_are_prefix_consistent(enc_model):
    """Size-independent property at sizes the oracle does not reach: everything in the path is causal, so the codes of a 5-minute
    item (7500 frames at 25 Hz, 30 attention windows; the largest single launch) must START with the codes of its first 20 s, bit
    for bit, except the prefix's last, partial-context frame; and a batch larger than one launch's sample budget (sub-batched by
    the engine) must give every item the codes it gets alone."""
    long_wav = make_audio(1, 24000 * 300, seed=31)
    full = enc_model.encode(long_wav).audio_codes[0]
    assert full.shape == (3750, 16)
    head = enc_model.encode(long_wav[:, :24000 * 20]).audio_codes[0]
    assert head.shape == (250, 16) and torch.equal(head[:-1], full[:249])
    with pytest.raises(ValueError):
        enc_model.encode(make_audio(1, 24000 * 321, seed=1))            # 8025 frames at 25 Hz: past max_position_embeddings
    many = make_audio(40, 24000 * 20, seed=32)                            # 19.2 M samples: two launches of at most 16 M
    out = enc_model.encode(many).audio_codes
    for b in (0, 33, 34, 39):
        assert torch.equal(out[b], enc_model.encode(many[b:b + 1]).audio_codes[0])


def test_encode_needs_encoder_weights(gpu_model):
    with pytest.raises(RuntimeError, match="not loaded"):
        gpu_model.encode(torch.zeros(1, 4800))
