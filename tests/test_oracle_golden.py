"""The oracle (CPU restatement) against golden vectors produced by the UNMODIFIED reference
(oracle/make_golden.py) — this is what pins the oracle.  fp32 restatement vs fp32 reference:
tolerance 2e-4 max-abs on waveforms in [-1, 1] (different op order inside torch kernels only)."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import decoder_oracle as O
from oracle.make_golden import weight_fingerprint

TOL = 2e-4


def test_weights_regenerate_bit_identically(sd):
    g = load_golden("forward_B2_T12.npz")
    np.testing.assert_array_equal(weight_fingerprint(sd), g["weight_fingerprint"])


def test_forward_matches_reference_golden(sd, cfg):
    g = load_golden("forward_B2_T12.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2)
    taps = {}
    with torch.inference_mode():
        wav = O.decoder_forward(sd, cfg, codes, taps=taps)
    assert wav.shape == (2, 1, 12 * 1920 - 555)
    assert np.abs(wav.numpy() - g["wav"]).max() < TOL
    np.testing.assert_allclose(taps["quantizer"].numpy(), g["quantizer"], atol=1e-5)  # dequantised embeddings: fp32 rounding
    np.testing.assert_allclose(taps["pre_conv"].numpy(), g["pre_conv"], atol=1e-4)
    np.testing.assert_allclose(taps["pre_transformer"].numpy(), g["pre_transformer"], atol=2e-4)
    np.testing.assert_allclose(taps["upsample.1.1"].numpy()[:, ::8], g["upsample_1_1"], atol=5e-4)
    np.testing.assert_allclose(taps["decoder.0"].numpy()[:, ::8], g["decoder_0"], atol=5e-4)
    np.testing.assert_allclose(taps["decoder.1"].numpy()[:, ::16, ::4], g["decoder_1"], atol=1e-3)
    np.testing.assert_allclose(taps["decoder.4"].numpy()[:, ::8, ::64], g["decoder_4"], atol=2e-3)


def test_chunked_small_chunks_matches_reference_golden(sd, cfg):
    g = load_golden("chunked_B1_T24_c10_l3.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2)
    with torch.inference_mode():
        wav = O.chunked_decode(sd, cfg, codes, chunk_size=10, left_context_size=3)
    assert wav.shape == g["wav"].shape == (1, 1, 24 * 1920 - 3 * 555)
    assert np.abs(wav.numpy() - g["wav"]).max() < TOL


def test_ragged_decode_matches_reference_golden(sd, cfg):
    g = load_golden("decode_ragged_B3_T14.npz")
    codes = torch.from_numpy(g["codes"])
    with torch.inference_mode():
        wavs = O.model_decode(sd, cfg, codes)
    total = 14 * 1920 - 555
    for i, n in enumerate(g["lengths"]):
        want = g[f"wav{i}"]
        assert wavs[i].shape == want.shape == (min(int(n) * 1920, total),)  # (codes[...,0] > 0).sum()*1920, clipped by the slice
        assert np.abs(wavs[i].numpy() - want).max() < TOL


def test_two_chunk_decode_matches_reference_golden(sd, cfg):
    g = load_golden("chunked_B1_T310.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2)
    with torch.inference_mode():
        wav = O.chunked_decode(sd, cfg, codes)
    assert wav.shape[-1] == int(g["num_samples"]) == 310 * 1920 - 2 * 555
    assert np.abs(wav.numpy()[..., ::8] - g["wav_sub8"]).max() < TOL


def test_chunk_plan_and_lengths():
    assert O.chunk_plan(125) == [(0, 125, 0)]
    assert O.chunk_plan(750) == [(0, 300, 0), (275, 600, 25), (575, 750, 25)]
    assert O.chunk_plan(24, 10, 3) == [(0, 10, 0), (7, 20, 3), (17, 24, 3)]
    # start - left_context > 0 is strict: a chunk starting at frame 25 takes ctx = start = 25
    assert O.chunk_plan(50, 25, 25) == [(0, 25, 0), (0, 50, 25)]
    assert O.expected_num_samples(125) == 239445
    assert O.expected_num_samples(750) == 1438335


def test_wrong_codebook_count_raises(sd, cfg):
    with pytest.raises(ValueError, match="Expected 16 layer of codes, got 15"):
        O.decoder_forward(sd, cfg, torch.zeros(1, 15, 4, dtype=torch.int64))
