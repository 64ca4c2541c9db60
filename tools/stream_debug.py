import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes
cfg = Qwen3TTSTokenizerV2DecoderConfig()
dec = synthetic_model().decoder
T = int(os.environ.get("T", 30))
codes = make_codes(3, T, cfg, seed=11).cuda().transpose(1, 2).contiguous()
whole = dec(codes)[:, 0]
torch.cuda.synchronize()
dec.reserve_streams(4, 5)
for P in (1, 3, 5):
    ids = [dec.open_stream() for _ in range(3)]
    got = [[] for _ in range(3)]
    t = [0, 0, -P]
    while min(t) < T:
        rows = [i for i in range(3) if 0 <= t[i] < T]
        if rows:
            pk = torch.stack([codes[i, :, t[i]:t[i] + P] for i in rows])
            wav, nv = dec.decode_step([ids[i] for i in rows], pk)
            torch.cuda.synchronize()
            for k, i in enumerate(rows):
                got[i].append(wav[k, :nv[k]].clone())
        t = [x + P for x in t]
    for i in range(3):
        w = torch.cat(got[i])
        print(f"P={P} stream {i}: {tuple(w.shape)} max diff {(w - whole[i]).abs().max().item():.3e}", flush=True)
    dec.reset_stream(ids[0])
    print("reset", flush=True)
    wav, nv = dec.decode_step([ids[0]], codes[0:1, :, :P])
    torch.cuda.synchronize()
    print("after reset", nv, torch.equal(wav[0, :nv[0]], got[0][0]), flush=True)
    for i in ids: dec.close_stream(i)
