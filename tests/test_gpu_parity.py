"""-m gpu parity tests: the CUDA path (through the C-ABI, via the reference-shaped Python classes)
against (a) golden vectors written by the UNMODIFIED reference and (b) the CPU oracle on seeded inputs.

Tolerances (north_star): dequantised embeddings to fp32 rounding; waveform max-abs <= 2e-3 and
SNR >= 40 dB vs the fp32 reference for the parity mode (fp16 tensor-core operands, fp32 accumulate);
the bf16 variant is reported separately with its own documented bound."""
import ctypes as C

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import decoder_oracle as O

pytestmark = pytest.mark.gpu

MAX_ABS = 2e-3
MIN_SNR = 40.0


def _stats(ref, got):
    ref = torch.as_tensor(ref).float().cpu()
    got = torch.as_tensor(got).float().cpu()
    return (ref - got).abs().max().item(), O.snr_db(ref, got)


@pytest.mark.parametrize("case", [
    (128, 64, 64, 1, 1, 0), (300, 1024, 1024, 1, 1, 0), (300, 512, 1024, 3, 1, 0), (1000, 96, 96, 7, 9, 0), (999, 96, 96, 1, 1, 0),
    (700, 192, 192, 7, 3, 0), (300, 1024, 6144, 1, 1, 0), (257, 1536, 6144, 2, 1, 0), (300, 768, 1920, 2, 1, 0), (300, 192, 288, 2, 1, 0),
    (1, 1024, 3072, 1, 1, 0), (129, 4096, 1024, 1, 1, 0), (300, 1024, 1024, 1, 1, 1), (1000, 96, 96, 7, 9, 1)])
def test_tcgen05_gemm_selftest(case):
    """tcgen05 implicit-GEMM (TMA + TMEM) vs the SIMT cross-check kernel on the device."""
    from qwen3_tts_b200 import _capi
    lib = _capi.load_library()
    d, r = C.c_float(-1), C.c_float(-1)
    assert lib.q3tts_selftest_gemm(0, *case, C.byref(d), C.byref(r)) == 0
    assert r.value > 0.1
    assert d.value <= 2e-4 * max(1.0, r.value), (case, d.value, r.value)   # fp32 accumulation-order noise only


def test_rvq_dequant_matches_reference_to_fp32_rounding(gpu_model):
    g = load_golden("forward_B2_T12.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2).contiguous()
    q = gpu_model.decoder.quantizer.decode(codes).cpu().numpy()
    assert q.shape == g["quantizer"].shape == (2, 512, 12)
    assert np.abs(q - g["quantizer"]).max() <= 1e-5 * np.abs(g["quantizer"]).max()


def test_forward_matches_reference_golden(gpu_model):
    g = load_golden("forward_B2_T12.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2).contiguous()
    wav = gpu_model.decoder(codes)
    assert wav.shape == (2, 1, 12 * 1920 - 555) and wav.dtype == torch.float32
    mx, snr = _stats(g["wav"], wav)
    print(f"forward B2 T12: max-abs {mx:.3e} snr {snr:.1f} dB")
    assert mx <= MAX_ABS and snr >= MIN_SNR
    assert wav.abs().max().item() <= 1.0


def test_stage_probes_track_the_oracle(gpu_model, sd, cfg):
    g = load_golden("forward_B2_T12.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2).contiguous()
    taps = {}
    with torch.inference_mode():
        O.decoder_forward(sd, cfg, codes, taps=taps)
    taps["decoder.0.snake"] = O.snake_beta(taps["decoder.0"], sd["decoder.1.block.0.alpha"], sd["decoder.1.block.0.beta"])
    taps["pre_transformer"] = taps["pre_transformer"].transpose(1, 2)
    stages = ["quantizer", "pre_conv", "pre_transformer", "upsample.0.0", "upsample.0.1", "upsample.1.0", "upsample.1.1", "decoder.0.snake"]
    stages += [f"decoder.{i}.block.{j}" for i in range(1, 5) for j in (1, 2, 3, 4)]
    for st in stages:
        got = gpu_model.decoder.probe(codes, st)
        assert got.shape == taps[st].shape, st
        mx, snr = _stats(taps[st], got)
        assert snr >= 48.0, (st, mx, snr)


def test_simt_cross_check_engine_agrees(gpu_model):
    g = load_golden("forward_B2_T12.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2).contiguous()
    dec = gpu_model.decoder
    a = dec(codes)
    dec.set_engine("simt")
    try:
        b = dec(codes)
    finally:
        dec.set_engine("tc")
    mx, snr = _stats(a, b)
    # two fp16 paths that differ only in accumulation order; the stack amplifies that like any other perturbation
    assert mx <= 4e-3 and snr >= 45.0


def test_chunked_decode_small_chunks_golden(gpu_model):
    g = load_golden("chunked_B1_T24_c10_l3.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2).contiguous()
    wav = gpu_model.decoder.chunked_decode(codes, chunk_size=10, left_context_size=3)
    assert tuple(wav.shape) == g["wav"].shape
    mx, snr = _stats(g["wav"], wav)
    assert mx <= MAX_ABS and snr >= MIN_SNR


def test_two_chunk_decode_golden(gpu_model):
    """The real two-chunk path (300 + 10 frames, 25 frames of left context) against the reference: every 8th sample of the whole
    waveform and EVERY sample of the 16384 around the chunk seam (sample 300*1920 - 555, where chunk 2's first emitted sample
    follows chunk 1's last)."""
    g = load_golden("chunked_B1_T310.npz")
    out = gpu_model.decode(torch.from_numpy(g["codes"]).cuda(), return_dict=True)
    wav = out.audio_values[0]
    assert wav.shape == (int(g["num_samples"]),) == (310 * 1920 - 2 * 555,)
    mx, snr = _stats(g["wav_sub8"][0, 0], wav[::8])
    print(f"two-chunk T310: max-abs {mx:.3e} snr {snr:.1f} dB")
    assert mx <= MAX_ABS and snr >= MIN_SNR
    s0 = int(g["seam_start"])
    assert s0 < 300 * 1920 - 555 < s0 + g["wav_seam"].shape[-1]
    mx, snr = _stats(g["wav_seam"][0, 0], wav[s0:s0 + g["wav_seam"].shape[-1]])
    print(f"two-chunk T310 seam: max-abs {mx:.3e} snr {snr:.1f} dB")
    assert mx <= MAX_ABS and snr >= MIN_SNR


def test_wrapper_decode_payload_forms_and_length_rule(gpu_model):
    """Qwen3TTSTokenizer.decode: list[dict] / dict / numpy / tensor payloads, zero padding, a[:l] trim."""
    from qwen3_tts_b200 import Qwen3TTSTokenizer
    g = load_golden("decode_ragged_B3_T14.npz")
    tok = Qwen3TTSTokenizer.from_model(gpu_model)
    codes = torch.from_numpy(g["codes"])
    lengths = [int(x) for x in g["lengths"]]
    total = 14 * 1920 - 555
    want = [g[f"wav{i}"] for i in range(3)]
    # list of dicts with unpadded per-sample codes (numpy and tensor mixed) -> pad_sequence path, host e2e entry
    payload = [{"audio_codes": codes[0, :lengths[0]].numpy()}, {"audio_codes": codes[1, :lengths[1]]}, {"audio_codes": codes[2, :lengths[2]]}]
    wavs, sr = tok.decode(payload)
    assert sr == 24000 and len(wavs) == 3
    for w, ref, n in zip(wavs, want, lengths):
        assert isinstance(w, np.ndarray) and w.dtype == np.float32 and w.ndim == 1
        assert w.shape == ref.shape == (min(n * 1920, total),)
        mx, snr = _stats(ref, w)
        assert mx <= MAX_ABS and snr >= MIN_SNR
    # dict with a padded batch tensor on the device -> model.decode path
    wavs2, _ = tok.decode({"audio_codes": codes.cuda()})
    for a, b in zip(wavs, wavs2):
        np.testing.assert_array_equal(a, b)
    # single (T, Q) tensor -> batch of one
    w1, _ = tok.decode({"audio_codes": codes[1, :lengths[1]]})
    # decoded alone the 9-frame item is NOT padded to 14, so it ends 555 samples early (look-ahead of the transposed convs)
    assert len(w1) == 1 and w1[0].shape == (lengths[1] * 1920 - 555,)

    class _Out:
        audio_codes = [codes[0]]
    w0, _ = tok.decode(_Out())
    assert w0[0].shape == want[0].shape
    assert tok.get_model_type() == "qwen3_tts_tokenizer_12hz"
    assert (tok.get_input_sample_rate(), tok.get_output_sample_rate(), tok.get_encode_downsample_rate(), tok.get_decode_upsample_rate()) == (
        24000, 24000, 1920, 1920)


def test_decode_host_pageable_and_pinned_buffers_agree(gpu_model, cfg):
    """q3tts_decode_host with a pinned result buffer stores the waveform straight into it from the output conv; a pageable buffer
    takes the staged device-to-host copy.  Same kernels, same bits; a row stride wider than the output keeps both paths honest."""
    from qwen3_tts_b200 import _capi
    from qwen3_tts_b200.modeling import _stream_ptr
    from qwen3_tts_b200.synthetic import make_codes
    d = gpu_model.decoder
    B, T = 3, 11
    codes = make_codes(B, T, cfg, seed=9).contiguous()
    codes[2, 7:] = 0
    total = T * 1920 - 555
    stride = total + 37
    outs = []
    for pinned in (True, False):
        wav = torch.full((B, stride), 7.0, dtype=torch.float32, pin_memory=pinned)
        lengths = (C.c_int64 * B)()
        rc = d._lib.q3tts_decode_host(d._h, codes.data_ptr(), B, T, wav.data_ptr(), stride, lengths, None, _stream_ptr(d.device))
        _capi.check(d._h, rc, "decode_host")
        assert [int(x) for x in lengths] == [total, total, 7 * 1920]
        outs.append((wav, [int(x) for x in lengths]))
    (a, la), (b, lb) = outs
    for i in range(B):
        assert torch.equal(a[i, :la[i]], b[i, :lb[i]])
        assert a[i, :la[i]].abs().max().item() <= 1.0
    assert torch.equal(a[:, total:], torch.full((B, stride - total), 7.0)) and torch.equal(b[:, total:], a[:, total:])  # padding columns untouched


def test_pcm16_and_int16_codes_wire_formats(gpu_model, cfg):
    """int16 code frames in, PCM16 out.  The int16 path must decode exactly what the int64 path decodes, and the
    PCM16 samples must be rint(wav * 32767) of the float result (what sf.write stores), ragged lengths included."""
    from qwen3_tts_b200 import Qwen3TTSTokenizer
    from qwen3_tts_b200.synthetic import make_codes
    tok = Qwen3TTSTokenizer.from_model(gpu_model)
    items = [{"audio_codes": make_codes(1, T, cfg, seed=20 + i)[0]} for i, T in enumerate((40, 17, 33))]
    wavs, sr = tok.decode(items)
    pcm, sr2 = tok.decode_pcm16([{"audio_codes": it["audio_codes"].to(torch.int16).numpy()} for it in items])
    assert sr == sr2 == 24000 and len(pcm) == 3
    for w, q in zip(wavs, pcm):
        assert q.dtype == np.int16 and q.shape == w.shape
        assert np.array_equal(q, np.rint(np.clip(w, -1.0, 1.0) * np.float32(32767.0)).astype(np.int16))


def test_error_behaviour_matches_reference(gpu_model):
    with pytest.raises(ValueError, match="Expected 16 layer of codes, got 15"):
        gpu_model.decoder(torch.zeros(1, 15, 4, dtype=torch.int64))
    from qwen3_tts_b200 import Qwen3TTSTokenizer
    with pytest.raises(TypeError):
        Qwen3TTSTokenizer.from_model(gpu_model).decode(3.14)
    with pytest.raises(KeyError):
        gpu_model.decoder.probe(torch.ones(1, 16, 4, dtype=torch.int64), "no.such.stage")


def test_full_size_batch_properties(gpu_model, sd, cfg):
    """BASELINE config 2 shape (64 x 10 s): size-independent properties + two items against the oracle."""
    from qwen3_tts_b200.synthetic import make_codes
    codes = make_codes(64, 125, cfg, seed=1)
    dev = codes.cuda()
    out = gpu_model.decode(dev, return_dict=True).audio_values
    assert len(out) == 64 and all(w.shape == (239445,) for w in out)
    assert all(bool(torch.isfinite(w).all()) and w.abs().max().item() <= 1.0 for w in out)
    # determinism and batch independence: an item decoded alone is bit-identical to its slot in the batch
    again = gpu_model.decode(dev, return_dict=True).audio_values
    assert all(torch.equal(a, b) for a, b in zip(out, again))
    solo = gpu_model.decode(dev[5:6], return_dict=True).audio_values[0]
    assert torch.equal(solo, out[5])
    # frame-causality with the 555-sample look-ahead
    pref = gpu_model.decoder(dev[:2, :60].transpose(1, 2).contiguous())
    full = gpu_model.decoder(dev[:2, :110].transpose(1, 2).contiguous())
    assert (pref - full[..., : 60 * 1920 - 555]).abs().max().item() <= 1e-4
    for b in (0, 63):
        with torch.inference_mode():
            ref = O.model_decode(sd, cfg, codes[b:b + 1])[0]
        mx, snr = _stats(ref, out[b])
        print(f"full-size item {b}: max-abs {mx:.3e} snr {snr:.1f} dB")
        assert mx <= MAX_ABS and snr >= MIN_SNR


@pytest.mark.parametrize("T", [1, 2, 17, 73, 130])
def test_edge_lengths_match_oracle(gpu_model, sd, cfg, T):
    """Shortest inputs and the attention corner cases against the CPU oracle: a single frame (every conv sees only its causal
    padding; the transposed convs emit (L-1)*s rows), T = 73 (first query whose window of 72 drops key 0), T = 130 (queries in
    three 64-query tiles, keys staged across tile boundaries), batch 2 with a zero-padded second item (length rule, modeling :1012-1015)."""
    from qwen3_tts_b200.synthetic import make_codes
    codes = make_codes(2, T, cfg, seed=100 + T)
    if T > 2:
        codes[1, T - T // 3:] = 0                                           # padded tail: decoded like any other codes, then cut by the length rule
    got = gpu_model.decode(codes.cuda(), return_dict=True).audio_values
    with torch.inference_mode():
        ref = O.model_decode(sd, cfg, codes)
    assert len(got) == len(ref) == 2
    for b in range(2):
        assert got[b].shape == ref[b].shape, (T, b, got[b].shape, ref[b].shape)
        if ref[b].numel() == 0:
            continue
        mx, snr = _stats(ref[b], got[b])
        print(f"T={T} item {b}: n={ref[b].numel()} max-abs {mx:.3e} snr {snr:.1f} dB")
        assert mx <= MAX_ABS and snr >= MIN_SNR, (T, b, mx, snr)


def test_long_form_and_sharded_shapes(gpu_model, cfg):
    """BASELINE config 4 (32 x 60 s: chunks 300 / 25+300 / 25+150) and the config 5 utterance shape (T = 375: 300 / 25+75),
    through size-independent properties: sample counts, batch independence across chunk boundaries, and the
    chunk seam rule wav(chunked)[start*1920 - k*555 ...] == forward(codes[start-25:end])[25*1920:]."""
    from qwen3_tts_b200.synthetic import make_codes
    dec = gpu_model.decoder
    codes = make_codes(32, 750, cfg, seed=9).cuda()
    out = gpu_model.decode(codes, return_dict=True).audio_values
    assert len(out) == 32 and all(w.shape == (750 * 1920 - 3 * 555,) for w in out)            # 1 438 335
    assert all(bool(torch.isfinite(w).all()) and w.abs().max().item() <= 1.0 for w in out)
    solo = gpu_model.decode(codes[7:8], return_dict=True).audio_values[0]
    assert torch.equal(solo, out[7])
    # second chunk of item 3 re-decoded by hand, the way chunked_decode defines it (modeling :889-893)
    c = codes[3:4].transpose(1, 2).contiguous()
    seam = dec(c[..., 275:600])[0, 0, 25 * 1920:]
    lo = 300 * 1920 - 555
    assert torch.equal(seam, out[3][lo:lo + seam.numel()])
    # config 5 shape, a few utterances of different true lengths (padding rule: codes[..., 0] > 0)
    c5 = make_codes(6, 375, cfg, seed=10)
    c5[4, 200:] = 0
    got = gpu_model.decode(c5.cuda(), return_dict=True).audio_values
    assert [w.numel() for w in got] == [375 * 1920 - 2 * 555] * 4 + [200 * 1920] + [375 * 1920 - 2 * 555]
    assert torch.equal(got[4], gpu_model.decode(c5[4:5].cuda(), return_dict=True).audio_values[0])


def test_streaming_equals_whole_sequence_forward(gpu_model, cfg):
    """Stateful per-packet decode (q3tts_stream_*): concatenated stream output == Decoder.forward(whole sequence) on
    [0, T*1920-555), for packet sizes 1, 3 and 5, streams that start at different steps, and a reset."""
    from qwen3_tts_b200.synthetic import make_codes
    dec = gpu_model.decoder
    T = 30
    codes = make_codes(3, T, cfg, seed=11).cuda().transpose(1, 2).contiguous()    # [3, Q, T]
    whole = dec(codes)[:, 0]                                                          # [3, T*1920-555]
    dec.reserve_streams(4, 5)
    for P in (1, 3, 5):
        ids = [dec.open_stream() for _ in range(3)]
        got = [[] for _ in range(3)]
        # stream 2 joins one step late: a step then mixes first-packet and continuing streams
        t = [0, 0, -P]
        while min(t) < T:
            rows = [i for i in range(3) if 0 <= t[i] < T]
            if rows:
                pk = torch.stack([codes[i, :, t[i]:t[i] + P] for i in rows])
                wav, nv = dec.decode_step([ids[i] for i in rows], pk)
                for k, i in enumerate(rows):
                    got[i].append(wav[k, :nv[k]].clone())
            t = [x + P for x in t]
        for i in range(3):
            w = torch.cat(got[i])
            assert w.shape == (T * 1920 - 555,), (P, i, w.shape)
            mx, snr = _stats(whole[i], w)
            assert mx <= 1e-5, (P, i, mx, snr)
            assert dec.stream_frames(ids[i]) == T
        # reset: the same stream decodes the same codes to the same samples again
        dec.reset_stream(ids[0])
        wav, nv = dec.decode_step([ids[0]], codes[0:1, :, :P])
        assert nv[0] == P * 1920 - 555 and torch.equal(wav[0, :nv[0]], got[0][0])
        for i in ids:
            dec.close_stream(i)
    with pytest.raises(KeyError):
        dec.decode_step([0], codes[0:1, :, :1])      # closed stream


def test_talker_stream_sink_equals_whole_sequence_forward(gpu_model, cfg):
    """SURVEY 8f-1 hand-off: the talker's per-step codec_ids [B, 16], pushed on the device one frame at a time, decode to the same
    samples as Decoder.forward over the stacked sequence (bit for bit); a stream that finishes early stops emitting."""
    from qwen3_tts_b200 import TalkerStreamSink
    from qwen3_tts_b200.synthetic import make_codes
    dec = gpu_model.decoder
    B, T, P = 3, 10, 2
    codes = make_codes(B, T, cfg, seed=21).cuda()                               # [B, T, Q] as the talker stacks them (talker :2280)
    whole = dec(codes.transpose(1, 2).contiguous())[:, 0]                       # [B, T*1920-555]
    short = dec(codes[2:3, :6].transpose(1, 2).contiguous())[0, 0]              # row 2 stops after 6 frames
    sink = TalkerStreamSink(dec, batch=B, packet=P)
    got = [[] for _ in range(B)]
    for t in range(T):
        if t == 6:
            sink.finish(2)
        out = sink.push(codes[:, t])
        assert (out is None) == (t % P == 0)
        if out is not None:
            rows, wav, nv = out
            assert rows == ([0, 1, 2] if t < 6 else [0, 1])
            for k, r in enumerate(rows):
                got[r].append(wav[k, :nv[k]].clone())
    sink.close()
    for r in (0, 1):
        assert torch.equal(torch.cat(got[r]), whole[r])
    assert torch.equal(torch.cat(got[2]), short)
    bad = TalkerStreamSink(dec, batch=1, packet=1, reserve=False)
    with pytest.raises(ValueError):
        bad.push(codes[:, 0])                                                   # [3, 16] for a sink of one stream
    bad.close()


def test_streaming_matches_reference_golden(gpu_model):
    """Streamed in packets of 2 frames, the golden batch must meet the same tolerance as the bulk decode vs the fp32 reference."""
    g = load_golden("forward_B2_T12.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2).contiguous().cuda()
    dec = gpu_model.decoder
    dec.reserve_streams(2, 2)
    ids = [dec.open_stream(), dec.open_stream()]
    parts = []
    for t in range(0, 12, 2):
        wav, nv = dec.decode_step(ids, codes[:, :, t:t + 2])
        assert nv[0] == nv[1]
        parts.append(wav[:, :nv[0]].clone())
    w = torch.cat(parts, dim=1)
    assert w.shape == (2, 12 * 1920 - 555)
    mx, snr = _stats(g["wav"][:, 0], w)
    print(f"streamed B2 T12 (P=2): max-abs {mx:.3e} snr {snr:.1f} dB")
    assert mx <= MAX_ABS and snr >= MIN_SNR


def test_bulk_decode_is_bitwise_reproducible(gpu_model, cfg):
    """Regression for a shared-memory slot hand-back race (DESIGN.md "Residual ring"): every run of the same batch must be
    bit-identical.  The race showed up in ~1 of 10 runs as a one-chunk glitch in the fp32 residual stream."""
    from qwen3_tts_b200.synthetic import make_codes
    codes = make_codes(64, 125, cfg, seed=3).cuda().transpose(1, 2).contiguous()
    ref = gpu_model.decoder(codes).clone()
    for _ in range(25):
        assert torch.equal(gpu_model.decoder(codes), ref)


def test_bf16_variant_reported_separately(sd, cfg):
    """bf16 tensor-core operands: same kernels, 8-bit mantissa.  Bound documented in DESIGN.md (SNR >= 30 dB)."""
    from qwen3_tts_b200 import synthetic_model
    g = load_golden("forward_B2_T12.npz")
    codes = torch.from_numpy(g["codes"]).transpose(1, 2).contiguous()
    m = synthetic_model(state_dict=sd, compute_mode="bf16")
    wav = m.decoder(codes)
    mx, snr = _stats(g["wav"], wav)
    print(f"bf16 variant B2 T12: max-abs {mx:.3e} snr {snr:.1f} dB")
    assert snr >= 30.0 and mx <= 5e-2
    del m
