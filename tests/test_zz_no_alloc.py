"""-m gpu test of q3tts_rvq_decode_i16_static (include/q3tts_decode.h): the int16-codes RVQ gather-sum whose staging lives in a
static __device__ array, so the call allocates nothing.  Collected last on purpose (test_zz_*).

Checked through the C-ABI against the existing entry point q3tts_rvq_decode on the same codes (widened to int64), on two calls in
a row - the second one smaller and with different codes, so stale contents of the shared scratch would show - plus the size
limit.  Tolerance: the one tests/test_gpu_parity.py uses for dequantised embeddings (<= 1e-5 relative to max |x|); the two entry
points run the same gather kernel on the same tables, so in practice they are bit-identical."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _static(dec, codes16_dev, out):
    from qwen3_tts_b200.modeling import _stream_ptr
    B, _, T = codes16_dev.shape
    return dec._lib.q3tts_rvq_decode_i16_static(dec._h, codes16_dev.data_ptr(), B, T, out.data_ptr(), _stream_ptr(dec.device))


def test_static_scratch_entry_matches_rvq_decode_twice(gpu_model, cfg):
    from qwen3_tts_b200 import _capi
    from qwen3_tts_b200.synthetic import make_codes
    dec = gpu_model.decoder
    assert int(dec._lib.q3tts_static_max_codes()) == 512 * 1024
    for B, T, seed in ((4, 300, 81), (2, 37, 82)):                       # two calls in a row; the second reuses a prefix of the scratch
        codes = make_codes(B, T, cfg, seed=seed).transpose(1, 2).contiguous()   # [B, Q, T] int64
        want = dec.quantizer.decode(codes)                                # existing entry point: q3tts_rvq_decode
        c16 = codes.to(torch.int16).cuda().contiguous()
        got = torch.full((B, cfg.codebook_dim, T), float("nan"), dtype=torch.float32, device=dec.device)
        before = torch.cuda.memory_allocated(dec.device)
        rc = _static(dec, c16, got)
        _capi.check(dec._h, rc, "rvq_decode_i16_static")
        torch.cuda.synchronize(dec.device)
        assert torch.cuda.memory_allocated(dec.device) == before          # nothing went through torch's allocator either
        assert got.shape == want.shape and bool(torch.isfinite(got).all())
        err = (got - want).abs().max().item()
        assert err <= 1e-5 * want.abs().max().item(), (B, T, err)
        assert torch.equal(got, want)                                     # same kernel, same tables, same order of additions


def test_static_scratch_entry_rejects_oversize(gpu_model, cfg):
    from qwen3_tts_b200 import _capi
    dec = gpu_model.decoder
    Q = cfg.num_quantizers
    T = 512 * 1024 // Q + 1                                               # one frame more than the scratch holds
    c16 = torch.ones((1, Q, T), dtype=torch.int16, device=dec.device)
    out = torch.zeros((1, cfg.codebook_dim, 8), dtype=torch.float32, device=dec.device)   # never written: the call must refuse first
    rc = _static(dec, c16, out)
    assert rc == _capi.BAD_SHAPE
    assert b"exceed the static scratch" in dec._lib.q3tts_last_error(dec._h)
    assert float(out.abs().max()) == 0.0
    # exactly at the limit is accepted
    T = 512 * 1024 // Q
    c16 = torch.ones((1, Q, T), dtype=torch.int16, device=dec.device)
    out = torch.empty((1, cfg.codebook_dim, T), dtype=torch.float32, device=dec.device)
    _capi.check(dec._h, _static(dec, c16, out), "rvq_decode_i16_static at the limit")
    want = dec.quantizer.decode(c16.to(torch.int64))
    assert torch.equal(out, want)
