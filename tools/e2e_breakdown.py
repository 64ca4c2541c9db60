"""Where the end-to-end time of Qwen3TTSTokenizer.decode goes (64 x 10 s): device step, C-ABI decode_host, Python wrapper."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from qwen3_tts_b200 import Qwen3TTSTokenizer, synthetic_model
from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
from qwen3_tts_b200.synthetic import make_codes

cfg = Qwen3TTSTokenizerV2DecoderConfig()
model = synthetic_model()
dec = model.decoder
tok = Qwen3TTSTokenizer.from_model(model)
B, T = 64, 125
codes_host = make_codes(B, T, cfg, seed=1)
codes_dev = codes_host.cuda().transpose(1, 2).contiguous()
items = [{"audio_codes": codes_host[b].clone().pin_memory()} for b in range(B)]
padded = codes_host.clone().pin_memory()

def timeit(fn, n=8):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3

print(f"device chunked_decode      {timeit(lambda: dec.chunked_decode(codes_dev)):.2f} ms")
print(f"model.decode_host (padded) {timeit(lambda: model.decode_host(padded)):.2f} ms")
print(f"tok.decode(list of dicts)  {timeit(lambda: tok.decode(items)):.2f} ms")
wav_d = dec.chunked_decode(codes_dev)
wav_h = torch.empty(wav_d.shape, dtype=torch.float32).pin_memory()
print(f"D2H {wav_d.numel()*4/1e6:.1f} MB pinned        {timeit(lambda: wav_h.copy_(wav_d, non_blocking=True)):.2f} ms")
print(f"pinned alloc 61 MB         {timeit(lambda: torch.empty((64, 239445), dtype=torch.float32, pin_memory=True)):.2f} ms")
