"""GPU debug tool: the encode path stage by stage against the float64 oracle, on the TC engine and on the SIMT cross-check engine.
    python tests/tools/encode_stages.py [samples]
Prints, per engine: latent error, codes agreement; then (TC vs SIMT) the latent difference, which isolates tensor-core kernel bugs
from plan / packing bugs (both engines share the plan and the packed weights)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from oracle import encoder_oracle as XO
from qwen3_tts_b200 import synthetic_model
from qwen3_tts_b200.synthetic import make_audio, make_encoder_state_dict

n = int(sys.argv[1]) if len(sys.argv) > 1 else 24000 * 3 + 123
sd = make_encoder_state_dict(0)
lengths = (n, n // 2 + 5)
wav = make_audio(2, n, seed=1, lengths=lengths)
mask = torch.zeros_like(wav, dtype=torch.long)
for b, l in enumerate(lengths): mask[b, :l] = 1
want, lat64, margin = XO.model_encode(sd, wav, mask)
m = synthetic_model(device="cuda:0", with_encoder=True)
res = {}
for eng in ("simt", "tc"):
    m.decoder.set_engine(eng)
    codes, lat = m.encode(wav, mask, return_latents=True)
    torch.cuda.synchronize()
    lat = lat.cpu().numpy(); res[eng] = lat
    err = np.abs(lat - lat64).max() / np.abs(lat64).max()
    print(f"[{eng}] latents: rel max err {err:.3e}  (nan: {np.isnan(lat).any()})  per-frame err head: {np.abs(lat - lat64).max(axis=(0,1))[:6]}")
    for b in range(2):
        g = codes[b].cpu().numpy()
        print(f"[{eng}] item {b}: identical {float((g == want[b]).mean()):.4f}, decidable/wrong at 1e-4: {XO.decidable(g.T, want[b].T, margin[b][:, :g.shape[0]], 1e-4)}")
print("tc vs simt latents:", float(np.abs(res['tc'] - res['simt']).max()))
