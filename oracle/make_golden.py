"""ORACLE — writes tests/golden/*.npz by running the UNMODIFIED reference decoder
(/root/reference, through oracle/ref_shim.py) on seeded synthetic weights and codes.

Run in the build container only:  python -m oracle.make_golden
The reference cannot travel to the GPU box, these small fixtures can.  Each file also stores a
fingerprint of the seeded weights so a test can prove it regenerated the same parameters."""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim  # noqa: E402
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig  # noqa: E402
from qwen3_tts_b200.synthetic import make_codes, make_decoder_state_dict  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")
FP_KEYS = ["pre_conv.conv.weight", "decoder.4.block.3.conv1.conv.weight", "quantizer.rvq_rest.vq.layers.7._codebook.embedding_sum",
           "decoder.2.block.0.beta", "pre_transformer.layers.5.mlp.down_proj.weight"]


def weight_fingerprint(sd) -> np.ndarray:
    return np.array([[float(sd[k].double().sum()), float(sd[k].double().abs().sum())] for k in FP_KEYS], dtype=np.float64)


def main():
    torch.set_num_threads(os.cpu_count() or 1)
    os.makedirs(OUT, exist_ok=True)
    cfg = Qwen3TTSTokenizerV2DecoderConfig()
    sd = make_decoder_state_dict(cfg, seed=0)
    fp = weight_fingerprint(sd)
    dec = ref_shim.build_reference_decoder(sd, codebook_dim=cfg.codebook_dim, attn_implementation="eager")
    taps = {}
    hooks = []

    def hook(name):
        def f(_m, _i, o):
            taps[name] = (o.last_hidden_state if hasattr(o, "last_hidden_state") else o).detach().clone()
        return f

    for name, mod in (("pre_conv", dec.pre_conv), ("pre_transformer", dec.pre_transformer), ("upsample.1.1", dec.upsample[1][1]),
                      ("decoder.0", dec.decoder[0]), ("decoder.1", dec.decoder[1]), ("decoder.2", dec.decoder[2]),
                      ("decoder.3", dec.decoder[3]), ("decoder.4", dec.decoder[4])):
        hooks.append(mod.register_forward_hook(hook(name)))

    with torch.inference_mode():
        # g1: forward, B=2, T=12, with per-stage taps
        codes = make_codes(2, 12, cfg, seed=0)
        wav = dec(codes.transpose(1, 2))
        quant = dec.quantizer.decode(codes.transpose(1, 2))
        np.savez_compressed(os.path.join(OUT, "forward_B2_T12.npz"), codes=codes.numpy(), wav=wav.numpy(), quantizer=quant.numpy(),
                            pre_conv=taps["pre_conv"].numpy(), pre_transformer=taps["pre_transformer"].numpy(),
                            upsample_1_1=taps["upsample.1.1"].numpy()[:, ::8], decoder_0=taps["decoder.0"].numpy()[:, ::8],
                            decoder_1=taps["decoder.1"].numpy()[:, ::16, ::4], decoder_2=taps["decoder.2"].numpy()[:, ::16, ::16],
                            decoder_3=taps["decoder.3"].numpy()[:, ::16, ::32], decoder_4=taps["decoder.4"].numpy()[:, ::8, ::64],
                            weight_fingerprint=fp)
        for h in hooks:
            h.remove()
        # g2: chunked_decode with small chunks (3 chunks, left context 3), B=1, T=24
        codes = make_codes(1, 24, cfg, seed=1)
        wav = dec.chunked_decode(codes.transpose(1, 2), chunk_size=10, left_context_size=3)
        np.savez_compressed(os.path.join(OUT, "chunked_B1_T24_c10_l3.npz"), codes=codes.numpy(), wav=wav.numpy(), weight_fingerprint=fp)
        # g3: model.decode with ragged (zero-padded) batch, B=3, T=14, lengths 14/9/5
        lengths = [14, 9, 5]
        codes = make_codes(3, 14, cfg, seed=2, lengths=lengths)
        wavs = ref_shim.reference_model_decode(dec, codes)
        np.savez_compressed(os.path.join(OUT, "decode_ragged_B3_T14.npz"), codes=codes.numpy(), lengths=np.array(lengths),
                            **{f"wav{i}": w.numpy() for i, w in enumerate(wavs)}, weight_fingerprint=fp)
        # g4: the real two-chunk path (chunk 300, left context 25), B=1, T=310; waveform stored every 8th sample
        codes = make_codes(1, 310, cfg, seed=3)
        wav = dec.chunked_decode(codes.transpose(1, 2))
        seam = 300 * 1920 - 555                      # first sample emitted by the second chunk
        s0 = seam - 8192
        np.savez_compressed(os.path.join(OUT, "chunked_B1_T310.npz"), codes=codes.numpy(), wav_sub8=wav.numpy()[..., ::8],
                            wav_seam=wav.numpy()[..., s0:s0 + 16384], seam_start=np.array(s0),
                            num_samples=np.array(wav.shape[-1]), weight_fingerprint=fp)
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
