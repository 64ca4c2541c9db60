"""Live check of the oracle restatement against the reference module itself; only possible in the
build container (skipped where /root/reference does not exist, e.g. on the GPU box)."""
import pytest
import torch

from oracle import decoder_oracle as O
from oracle import ref_shim

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="reference tree not mounted")


def test_restatement_equals_reference_forward(sd, cfg):
    from qwen3_tts_b200.synthetic import make_codes
    codes = make_codes(1, 6, cfg, seed=11).transpose(1, 2)
    dec = ref_shim.build_reference_decoder(sd, codebook_dim=cfg.codebook_dim, attn_implementation="eager")
    with torch.inference_mode():
        ref = dec(codes)
        got = O.decoder_forward(sd, cfg, codes)
        q_ref = dec.quantizer.decode(codes)
        q_got = O.rvq_decode(sd, cfg, codes)
    assert ref.shape == got.shape
    assert (ref - got).abs().max().item() < 2e-4
    assert (q_ref - q_got).abs().max().item() < 1e-5
