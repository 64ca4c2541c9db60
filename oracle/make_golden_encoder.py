"""ORACLE - writes tests/golden/encode_B2.npz by running the reference's own encoder class (Qwen3TTSTokenizerV2Encoder, imported
unmodified through oracle/ref_shim.py; its arithmetic is the `transformers` package's MimiModel) and the reference's
Qwen3TTSTokenizerV2Model.encode post-processing (modeling :979-985) on seeded weights and seeded ragged audio.
    python -m oracle.make_golden_encoder"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from qwen3_tts_b200.synthetic import make_audio, make_encoder_state_dict  # noqa: E402

LENGTHS = (24000 * 11 + 333, 24000 * 2 + 1)      # 11.01 s (276 frames at 25 Hz: past the 250-frame attention window) and 2.00004 s


def golden_inputs():
    wav = make_audio(len(LENGTHS), max(LENGTHS), seed=0, lengths=LENGTHS)
    mask = torch.zeros_like(wav, dtype=torch.long)
    for b, n in enumerate(LENGTHS):
        mask[b, :n] = 1
    return wav, mask


def main():
    sd = make_encoder_state_dict(seed=0)
    wav, mask = golden_inputs()
    enc = ref_shim.build_reference_encoder(sd)
    codes = ref_shim.reference_model_encode(enc, wav, mask)
    with torch.inference_mode():
        x = enc.encoder(wav.unsqueeze(1))
        y = enc.encoder_transformer(x.transpose(1, 2))[0].transpose(1, 2)
        lat = enc.downsample(y)
    out = os.path.join(ROOT, "tests", "golden", "encode_B2.npz")
    np.savez_compressed(out, lengths=np.array(LENGTHS), codes0=codes[0].numpy().astype(np.int16), codes1=codes[1].numpy().astype(np.int16),
                        seanet=x.numpy()[:, ::8, ::4].astype(np.float32), latents=lat.numpy().astype(np.float32),
                        transformers_version=np.array(__import__("transformers").__version__))
    print(out, [tuple(c.shape) for c in codes], tuple(lat.shape), os.path.getsize(out))


if __name__ == "__main__":
    main()
