"""ORACLE - test / baseline infrastructure.  Stages the UNMODIFIED reference decoder under oracle/_ref/ so that it can travel to
the GPU box (oracle/_ref/ is git-ignored: it never enters the history, like the built .so files; it is not gpurun-ignored).

    python -m oracle.stage_ref            (run in the build container, where /root/reference exists; __graft_entry__.build() does)

The reference is a pure-Python package with no native sources, so there is nothing to compile: its Tokenizer-12Hz decode path is
two files (V2 configuration + V2 modeling).  They are copied byte for byte - MANIFEST.json records their sha256 -
and imported through oracle/ref_shim.py exactly as from /root/reference.  Used only by bench.py's baseline legs
(--impl reference, cpu_baseline, torch_gpu context) and the tests; the product path never imports anything from here."""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(os.environ.get("Q3TTS_REFERENCE_ROOT", "/root/reference"), "qwen_tts", "core", "tokenizer_12hz")
DST = os.path.join(ROOT, "oracle", "_ref", "qwen_tts", "core", "tokenizer_12hz")
FILES = ("configuration_qwen3_tts_tokenizer_v2.py", "modeling_qwen3_tts_tokenizer_v2.py")


def staged() -> bool:
    return all(os.path.isfile(os.path.join(DST, f)) for f in FILES)


def stage() -> str:
    """Returns "staged" (copied now), "present" (already there, reference tree absent) or "unavailable"."""
    if not all(os.path.isfile(os.path.join(SRC, f)) for f in FILES):
        return "present" if staged() else "unavailable"
    os.makedirs(DST, exist_ok=True)
    manifest = {}
    for f in FILES:
        shutil.copyfile(os.path.join(SRC, f), os.path.join(DST, f))
        manifest[f] = hashlib.sha256(open(os.path.join(DST, f), "rb").read()).hexdigest()
    json.dump({"source": SRC, "sha256": manifest, "note": "byte-for-byte copies of the reference files; not product source"},
              open(os.path.join(ROOT, "oracle", "_ref", "MANIFEST.json"), "w"), indent=1)
    return "staged"


if __name__ == "__main__":
    print(stage(), DST)
    sys.exit(0)
