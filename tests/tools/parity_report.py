"""Parity numbers per BASELINE.json config and compute mode -> JSON (copied to profiles/r2_parity.json).

    python tests/tools/parity_report.py [--out gpurun_out/r2_parity.json] [--modes fp16,bf16]

For every config the CUDA path (through the reference-shaped Python classes over the C-ABI) is compared with the CPU oracle
(oracle/decoder_oracle.py, fp32, pinned to the unmodified reference by tests/golden): max-abs error and SNR =
10 log10(sum ref^2 / sum (ref - out)^2), worst over the compared items.  Tolerance (north_star): max-abs <= 2e-3, SNR >= 40 dB for
the parity mode; the bf16 variant is reported separately.  The oracle is the checker here, never the thing measured."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from oracle import decoder_oracle as O  # noqa: E402
from qwen3_tts_b200 import synthetic_model  # noqa: E402
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig  # noqa: E402
from qwen3_tts_b200.synthetic import make_codes, make_decoder_state_dict  # noqa: E402


def stats(refs, gots):
    mx, snr = 0.0, 1e9
    for r, g in zip(refs, gots):
        r = r.float().cpu(); g = g.float().cpu()
        assert r.shape == g.shape, (r.shape, g.shape)
        mx = max(mx, (r - g).abs().max().item())
        snr = min(snr, O.snr_db(r, g))
    return {"max_abs": mx, "snr_db": snr, "items": len(refs), "within_tolerance": bool(mx <= 2e-3 and snr >= 40.0)}


def stream_decode(dec, codes_dev, P):
    """codes_dev [S, Q, T] -> list of S concatenated stream outputs."""
    S, _, T = codes_dev.shape
    dec.reserve_streams(S, max(P, 1))
    ids = [dec.open_stream() for _ in range(S)]
    parts = []
    for t in range(0, T, P):
        wav, nv = dec.decode_step(ids, codes_dev[:, :, t:t + P].contiguous())
        parts.append(wav[:, :nv[0]].clone())
    for i in ids:
        dec.close_stream(i)
    return list(torch.cat(parts, dim=1))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "r2_parity.json"))
    ap.add_argument("--modes", default="fp16,bf16")
    args = ap.parse_args()
    torch.set_num_threads(os.cpu_count() or 1)
    cfg = Qwen3TTSTokenizerV2DecoderConfig()
    sd = make_decoder_state_dict(cfg, seed=0)
    sds = make_decoder_state_dict(cfg, seed=0, stress=True)
    t0 = time.time()
    # ---- inputs and oracle references (computed once, shared by every mode)
    c1 = make_codes(1, 125, cfg, seed=0)
    c2 = make_codes(64, 125, cfg, seed=1); pick2 = [0, 9, 18, 27, 36, 45, 54, 63]
    c3 = make_codes(256, 96, cfg, seed=32).transpose(1, 2).contiguous(); pick3 = [0, 31, 128, 255]
    c4 = make_codes(32, 750, cfg, seed=9); pick4 = [11]
    c5 = make_codes(6, 375, cfg, seed=10); c5[4, 200:] = 0; pick5 = [2, 4]
    cs = make_codes(2, 12, cfg, seed=0)
    with torch.inference_mode():
        r1 = O.model_decode(sd, cfg, c1)
        r2 = O.model_decode(sd, cfg, c2[pick2])
        r3 = list(O.decoder_forward(sd, cfg, c3[pick3])[:, 0])
        r4 = O.model_decode(sd, cfg, c4[pick4])
        r5 = O.model_decode(sd, cfg, c5[pick5])
        rs = O.model_decode(sds, cfg, cs)
        rq = O.rvq_decode(sd, cfg, c1.transpose(1, 2))
    oracle_s = time.time() - t0
    report = {"tolerance": {"max_abs": 2e-3, "snr_db": 40.0, "against": "fp32 CPU oracle (oracle/decoder_oracle.py), pinned to the unmodified reference by tests/golden"},
              "weights": "seeded synthetic initialiser (qwen3_tts_b200/synthetic.py): no checkpoint is mounted; 'stress' = SURVEY 8(c) Snake beta ~ N(0, 0.3)",
              "oracle_seconds": round(oracle_s, 1), "modes": {}}
    for mode in args.modes.split(","):
        m = synthetic_model(state_dict=sd, compute_mode=mode)
        dec = m.decoder
        out = {}
        q = dec.quantizer.decode(c1.transpose(1, 2).contiguous()).cpu()
        out["dequantised_embeddings"] = {"max_abs": (q - rq).abs().max().item(), "rel_to_max": ((q - rq).abs().max() / rq.abs().max()).item(),
                                         "note": "fp32 gather-sum in every mode; bar = fp32 rounding (<= 1e-5 relative)"}
        g = m.decode(c1.cuda(), return_dict=True).audio_values
        out["config1_B1_T125"] = stats(r1, g)
        g = m.decode(c2.cuda(), return_dict=True).audio_values
        out["config2_B64_T125"] = dict(stats(r2, [g[b] for b in pick2]), compared=f"items {pick2} of 64")
        for P in (1, 4):
            g = stream_decode(dec, c3.cuda(), P)
            out[f"config3_S256_P{P}_T96"] = dict(stats(r3, [g[i] for i in pick3]), compared=f"streams {pick3} of 256, whole sequences")
        g = m.decode(c4.cuda(), return_dict=True).audio_values
        out["config4_B32_T750"] = dict(stats(r4, [g[b] for b in pick4]), compared=f"item {pick4} of 32, all 1438335 samples")
        g = m.decode(c5.cuda(), return_dict=True).audio_values
        out["config5_T375"] = dict(stats(r5, [g[b] for b in pick5]), compared=f"items {pick5} of a ragged batch of 6")
        del m, dec
        torch.cuda.empty_cache()
        ms = synthetic_model(state_dict=sds, compute_mode=mode)
        g = ms.decode(cs.cuda(), return_dict=True).audio_values
        out["stress_weights_B2_T12"] = stats(rs, g)
        del ms
        torch.cuda.empty_cache()
        report["modes"][mode] = out
        print(mode, json.dumps(out), flush=True)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(report, open(args.out, "w"), indent=1)
    print("wrote", args.out)


if __name__ == "__main__":
    main()
