"""ORACLE - test infrastructure, not product code.

CPU restatement of the quantiser stage at the end of the reference's encode path.  Qwen3TTSTokenizerV2Model.encode
(/root/reference/qwen_tts/core/tokenizer_12hz/modeling_qwen3_tts_tokenizer_v2.py:960-990) calls MimiModel.encode of the
third-party `transformers` package (pinned ==4.57.3 in the reference's pyproject.toml:23; 5.5.0 in this image; NOT vendored in
/root/reference), whose last stage is MimiSplitResidualVectorQuantizer.encode (transformers/models/mimi/modeling_mimi.py):

    MimiEuclideanCodebook.embed      embed_sum / cluster_usage.clamp(min=1e-5)[:, None]
    MimiEuclideanCodebook.quantize   torch.cdist(x, embed, p=2).argmin(-1)                 (first minimum wins)
    MimiResidualVectorQuantizer.encode   r = input_proj(x); per layer: idx = quantize(r); r = r - embed[idx]
    MimiSplitResidualVectorQuantizer.encode   semantic RVQ (1 codebook) and acoustic RVQ both start from the SAME latents

Pinning: tests/golden/rvq_encode_B2_T40.npz holds codes written by HF's own MimiSplitResidualVectorQuantizer on seeded weights
(oracle/make_golden_encode.py; tests/test_oracle_encode.py re-runs HF live when transformers has the Mimi model).
Distances are evaluated in float64 here, so `margin` tells how decisive every argmin was: an fp32 implementation may only differ
on frames whose relative margin is at fp32 rounding level."""
from __future__ import annotations

from typing import Dict, Tuple

import numpy as np

SEM = "encoder.quantizer.semantic_residual_vector_quantizer."
ACO = "encoder.quantizer.acoustic_residual_vector_quantizer."


def _np(t) -> np.ndarray:
    return t.detach().cpu().numpy() if hasattr(t, "detach") else np.asarray(t)


def rvq_encode(sd: Dict[str, object], latents, num_quantizers: int = 16, num_semantic: int = 1) -> Tuple[np.ndarray, np.ndarray]:
    """latents [B, D, T] -> (codes [B, K, T] int64, margin [B, K, T] float64 = (second best - best) / best squared distance)."""
    x = _np(latents).astype(np.float64)
    B, D, T = x.shape
    codes = np.zeros((B, num_quantizers, T), dtype=np.int64)
    margin = np.zeros((B, num_quantizers, T), dtype=np.float64)
    flat = x.transpose(0, 2, 1).reshape(B * T, D)
    k_out = 0
    for prefix, n in ((SEM, num_semantic), (ACO, num_quantizers - num_semantic)):
        w = _np(sd[prefix + "input_proj.weight"]).astype(np.float64)[:, :, 0]        # Conv1d(D -> dq, k = 1, bias = False)
        r = flat @ w.T
        for layer in range(n):
            es = _np(sd[f"{prefix}layers.{layer}.codebook.embed_sum"]).astype(np.float64)
            cu = _np(sd[f"{prefix}layers.{layer}.codebook.cluster_usage"]).astype(np.float64)
            embed = es / np.maximum(cu, 1e-5)[:, None]
            d2 = (r * r).sum(1)[:, None] - 2.0 * (r @ embed.T) + (embed * embed).sum(1)[None, :]
            idx = d2.argmin(1)                                                        # first minimum, like torch.argmin
            part = np.partition(d2, 1, axis=1)
            best = np.maximum(part[:, 0], 1e-300)
            margin[:, k_out, :] = ((part[:, 1] - part[:, 0]) / best).reshape(B, T)
            codes[:, k_out, :] = idx.reshape(B, T)
            r = r - embed[idx]
            k_out += 1
    return codes, margin


def make_encoder_quantizer_state_dict(codebook_size: int = 2048, codebook_dim: int = 512, num_quantizers: int = 16, seed: int = 0):
    """Seeded synthetic `encoder.quantizer.*` tensors with the reference's key names and shapes (no checkpoint is mounted)."""
    import torch
    g = torch.Generator().manual_seed(1234 + seed)
    dq = codebook_dim // 2
    sd = {}
    for prefix, n in ((SEM, 1), (ACO, num_quantizers - 1)):
        sd[prefix + "input_proj.weight"] = torch.randn((dq, codebook_dim, 1), generator=g) / codebook_dim ** 0.5
        for layer in range(n):
            # later layers quantise ever smaller residuals: scale their codebooks down like a trained RVQ's
            scale = 0.7 ** layer
            sd[f"{prefix}layers.{layer}.codebook.embed_sum"] = torch.randn((codebook_size, dq), generator=g) * scale
            sd[f"{prefix}layers.{layer}.codebook.cluster_usage"] = torch.rand((codebook_size,), generator=g) * 2.0 + 0.5
    return sd


def make_latents(batch: int, frames: int, codebook_dim: int = 512, seed: int = 0):
    import torch
    g = torch.Generator().manual_seed(4321 + seed)
    return torch.randn((batch, codebook_dim, frames), generator=g)
