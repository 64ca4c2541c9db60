"""GPU debug tool: run-to-run determinism of the encode path on the golden input (latents must be bit-identical between runs)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np, torch
from oracle.make_golden_encoder import golden_inputs
from qwen3_tts_b200 import synthetic_model

g = np.load(os.path.join(os.path.dirname(__file__), "..", "golden", "encode_B2.npz"))
wav, mask = golden_inputs()
m = synthetic_model(device="cuda:0", with_encoder=True)
eng = sys.argv[1] if len(sys.argv) > 1 else "tc"
m.decoder.set_engine(eng)
first = None
for i in range(int(sys.argv[2]) if len(sys.argv) > 2 else 6):
    codes, lat = m.encode(wav, mask, return_latents=True)
    torch.cuda.synchronize()
    lat = lat.cpu().numpy()
    err = np.abs(lat - g["latents"])
    w = np.unravel_index(err.argmax(), err.shape)
    msg = f"run {i}: rel err {err.max() / np.abs(g['latents']).max():.3e} at (b, c, t) = {w}"
    if first is None:
        first = lat
    else:
        d = np.abs(lat - first)
        bad = np.argwhere(d.max(axis=1) > 0)
        msg += f"; vs run 0: max diff {d.max():.3e}, frames that differ: {len(bad)} {bad[:8].tolist()}"
    print(msg, flush=True)
