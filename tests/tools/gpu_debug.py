"""GPU bring-up script: tcgen05 GEMM self-tests, then stage-by-stage parity against the oracle.
Run on the GPU box:  python tests/tools/gpu_debug.py [T] [B]   (writes gpurun_out/debug.log too)."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch

from oracle import decoder_oracle as O
from qwen3_tts_b200 import _capi, synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes, make_decoder_state_dict

T = int(sys.argv[1]) if len(sys.argv) > 1 else 12
B = int(sys.argv[2]) if len(sys.argv) > 2 else 2
os.makedirs("gpurun_out", exist_ok=True)
log = open("gpurun_out/debug.log", "a")


def P(*a):
    s = " ".join(str(x) for x in a)
    print(s, flush=True)
    log.write(s + "\n"); log.flush()


lib = _capi.load_library()
P("== GEMM self-tests (tcgen05 vs SIMT), (rows, cin, n, taps, dil, bf16) -> max|diff| / max|ref|")
for case in [(128, 64, 64, 1, 1, 0), (300, 1024, 1024, 1, 1, 0), (300, 512, 1024, 3, 1, 0), (1000, 96, 96, 7, 9, 0), (1000, 96, 96, 1, 1, 0),
             (700, 192, 192, 7, 3, 0), (300, 1024, 6144, 1, 1, 0), (257, 1536, 6144, 2, 1, 0), (300, 768, 1920, 2, 1, 0), (300, 192, 288, 2, 1, 0),
             (300, 1024, 1024, 1, 1, 1)]:
    d, r = C.c_float(-1), C.c_float(-1)
    t0 = time.time()
    rc = lib.q3tts_selftest_gemm(0, *case, C.byref(d), C.byref(r))
    P(f"  {case}: rc={rc} diff={d.value:.3e} ref={r.value:.3e} ({time.time()-t0:.2f}s)")

cfg = Qwen3TTSTokenizerV2DecoderConfig()
sd = make_decoder_state_dict(cfg, 0)
codes = make_codes(B, T, cfg, 0).transpose(1, 2).contiguous()  # [B,Q,T]
taps = {}
with torch.inference_mode():
    ref = O.decoder_forward(sd, cfg, codes, taps=taps)
# oracle tap for our fused "decoder.0.snake" probe
taps["decoder.0.snake"] = O.snake_beta(taps["decoder.0"], sd["decoder.1.block.0.alpha"], sd["decoder.1.block.0.beta"])
model = synthetic_model(state_dict=sd)
dec = model.decoder
stages = ["quantizer", "pre_conv", "pre_transformer", "upsample.0.0", "upsample.0.1", "upsample.1.0", "upsample.1.1", "decoder.0.snake"]
for i in range(1, 5):
    stages += [f"decoder.{i}.block.{j}" for j in (1, 2, 3, 4)]
for engine in ("simt", "tc"):
    dec.set_engine(engine)
    P(f"== stage parity, engine={engine}, B={B} T={T}")
    q = dec.quantizer.decode(codes).cpu()
    P(f"  rvq_decode fp32: maxabs diff {(q - taps['quantizer']).abs().max().item():.3e} (ref absmax {taps['quantizer'].abs().max().item():.2f})")
    for st in stages:
        try:
            got = dec.probe(codes, st).cpu()
        except Exception as e:  # noqa
            P(f"  {st:24s} FAILED: {e}")
            break
        want = taps[st] if st != "pre_transformer" else taps[st].transpose(1, 2)
        if got.shape != want.shape:
            P(f"  {st:24s} SHAPE got {tuple(got.shape)} want {tuple(want.shape)}")
            continue
        P(f"  {st:24s} {tuple(got.shape)} maxabs {(got - want).abs().max().item():.3e} snr {O.snr_db(want, got):.1f} dB (ref absmax {want.abs().max().item():.2f})")
    try:
        wav = dec(codes).cpu()
        P(f"  wav {tuple(wav.shape)} maxabs {(wav - ref).abs().max().item():.3e} snr {O.snr_db(ref, wav):.1f} dB")
    except Exception as e:  # noqa
        P(f"  forward FAILED: {e}")
P("done")
