"""Same-box, same-process A/B of programmatic dependent launch: alternating off / on, bulk step (64 x 10 s) and streaming step.
    python tools/ab_pdl.py [streams]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from qwen3_tts_b200 import _capi, synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

S = int(sys.argv[1]) if len(sys.argv) > 1 else 256
lib = _capi.load_library()
cfg = Qwen3TTSTokenizerV2DecoderConfig()
model = synthetic_model()
dec = model.decoder
codes = make_codes(64, 125, cfg, seed=1).cuda().transpose(1, 2).contiguous()
dec.reserve(64 * 125)
ref = dec(codes).clone()


def bulk(n=6):
    ts = []
    for _ in range(n):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); w = dec(codes); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    assert torch.equal(w, ref)
    return sorted(ts)[len(ts) // 2]


dec.reserve_streams(S, 4)
sc = make_codes(S, 400, cfg, seed=2).cuda().transpose(1, 2).contiguous()
ids = [dec.open_stream() for _ in range(S)]
out = torch.empty((S, 1920), dtype=torch.float32, device="cuda")
pos = [0]


def stream(n=150):
    ts = []
    for _ in range(n):
        t = pos[0] % 400
        pk = sc[:, :, t:t + 1].contiguous()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); dec.decode_step(ids, pk, out=out); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b)); pos[0] += 1
    return sorted(ts)[len(ts) // 2]


for on in (1, 0, 1, 0, 1, 0):
    lib.q3tts_debug_set_pdl(on)
    bulk(2); stream(20)
    print(f"pdl={on}  bulk step {bulk():7.3f} ms   stream step ({S} x 1) {stream():6.3f} ms", flush=True)
