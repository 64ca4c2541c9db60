"""CPU: the encode-side quantiser oracle (oracle/encode_oracle.py) against codes written by the `transformers` package's own
MimiSplitResidualVectorQuantizer (tests/golden/rvq_encode_B2_T40.npz, oracle/make_golden_encode.py), and against HF run live."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import encode_oracle as EO


def test_encode_oracle_matches_hf_golden():
    g = load_golden("rvq_encode_B2_T40.npz")
    sd = EO.make_encoder_quantizer_state_dict(seed=0)
    codes, margin = EO.rvq_encode(sd, EO.make_latents(2, 40, seed=0))
    assert codes.shape == g["codes"].shape == (2, 16, 40) and codes.dtype == np.int64
    assert np.array_equal(codes, g["codes"])                       # integer work: identical indices
    assert margin.min() > 0 and len(np.unique(codes)) > 500       # a real search, not a degenerate fixture


def test_encode_oracle_matches_hf_live():
    pytest.importorskip("transformers.models.mimi.modeling_mimi")
    from oracle.make_golden_encode import hf_quantizer
    sd = EO.make_encoder_quantizer_state_dict(seed=3)
    x = EO.make_latents(3, 25, seed=3)
    with torch.inference_mode():
        want = hf_quantizer(sd).encode(x, num_quantizers=16).permute(1, 0, 2).numpy()
    codes, margin = EO.rvq_encode(sd, x)
    # HF scores in fp32 (torch.cdist); a frame may differ only where the two best codes tie to fp32 rounding, and a differing
    # index changes the residual, so everything after it in that frame is excluded as well
    for b in range(3):
        for t in range(25):
            k_bad = next((k for k in range(16) if codes[b, k, t] != want[b, k, t]), None)
            assert k_bad is None or margin[b, k_bad, t] < 1e-5, (b, t, k_bad)
    assert (codes == want).mean() > 0.99
