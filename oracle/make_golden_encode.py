"""ORACLE - writes tests/golden/rvq_encode_B2_T40.npz by running the `transformers` package's own
MimiSplitResidualVectorQuantizer.encode (the third-party code the reference's encode path ends in, modeling :960-990) on seeded
synthetic weights and latents.     python -m oracle.make_golden_encode"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import encode_oracle as EO  # noqa: E402


def hf_quantizer(sd, num_quantizers_total: int = 32):
    """HF quantiser with the reference's Mimi shapes (32 codebooks of 2048 x 256, hidden 512) loaded with `sd` (first 16 used)."""
    from transformers import MimiConfig
    from transformers.models.mimi.modeling_mimi import MimiSplitResidualVectorQuantizer
    cfg = MimiConfig(num_quantizers=num_quantizers_total, codebook_size=2048, codebook_dim=256, hidden_size=512, vector_quantization_hidden_dimension=256)
    q = MimiSplitResidualVectorQuantizer(cfg).eval()
    own = q.state_dict()
    for k, v in sd.items():
        kk = k[len("encoder.quantizer."):]
        assert kk in own and own[kk].shape == v.shape, (kk, tuple(v.shape))
        own[kk] = v.clone()
    q.load_state_dict(own)
    return q


def main():
    sd = EO.make_encoder_quantizer_state_dict(seed=0)
    x = EO.make_latents(2, 40, seed=0)
    q = hf_quantizer(sd)
    with torch.inference_mode():
        codes = q.encode(x, num_quantizers=16)          # [K, B, T]
    codes = codes.permute(1, 0, 2).contiguous().numpy()
    out = os.path.join(ROOT, "tests", "golden", "rvq_encode_B2_T40.npz")
    np.savez_compressed(out, codes=codes, transformers_version=np.array(__import__("transformers").__version__))
    print(out, codes.shape, os.path.getsize(out))


if __name__ == "__main__":
    main()
