"""Per-op device time of one steady-state streaming step (S streams x P frames).  usage: profile_stream.py [S] [P]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

S = int(sys.argv[1]) if len(sys.argv) > 1 else 256
P = int(sys.argv[2]) if len(sys.argv) > 2 else 1
cfg = Qwen3TTSTokenizerV2DecoderConfig()
dec = synthetic_model(compute_mode=os.environ.get("MODE", "fp16")).decoder
dec.reserve_streams(S, max(P, 4))
ids = [dec.open_stream() for _ in range(S)]
codes = make_codes(S, 40 * P, cfg, seed=5).cuda().transpose(1, 2).contiguous()
out = torch.empty((S, P * 1920), dtype=torch.float32, device="cuda")
for i in range(20):
    dec.decode_step(ids, codes[:, :, i * P:(i + 1) * P], out=out)
torch.cuda.synchronize()
names = {0: "rvq", 1: "gemm", 2: "rmsnorm", 3: "attn", 4: "dwln", 5: "tail", 6: "unit", 7: "s_attn", 8: "st_load", 9: "st_save", 10: "pos"}
acc = None
N = 10
for i in range(20, 20 + N):
    dec.set_profile(True)
    dec.decode_step(ids, codes[:, :, i * P:(i + 1) * P], out=out)
    torch.cuda.synchronize()
    prof = dec.get_profile()
    dec.set_profile(False)
    if acc is None:
        acc = [[p[0], 0.0, p[2], p[3], p[4]] for p in prof]
    for a, p in zip(acc, prof):
        a[1] += p[1] / N
tot = sum(a[1] for a in acc)
print(f"S={S} P={P}: {len(acc)} ops, sum of op times {tot:.3f} ms")
for i, (ty, ms, fl, by, dm) in enumerate(acc):
    print(f"{i:3d} {names[ty]:8s} {ms*1e3:8.1f} us  rows {dm[0]:7d} n {dm[1]:5d} cin {dm[2]:5d} taps {dm[3]}  {fl/ms/1e9 if ms else 0:7.1f} TF/s {by/ms/1e6 if ms else 0:7.0f} GB/s")
bt = {}
for ty, ms, *_ in acc:
    bt[names[ty]] = bt.get(names[ty], 0) + ms
print({k: round(v, 3) for k, v in bt.items()})
