"""Builds libq3tts_decode.so (hand-written sm_100a CUDA + C-ABI) in-tree with nvcc.

In-tree so the .so travels to the GPU box with the repo snapshot; nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "libq3tts_decode.so")
LIB_PATH_DEBUG = os.path.join(LIB_DIR, "libq3tts_decode_dbg.so")  # -DQ3TTS_DEBUG: env switches + dbg experiment bits (tools/ only)
OBJ_DIR = os.path.join(HERE, "build")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC"]


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _newest_src_mtime() -> float:
    m = 0.0
    for root in (CSRC, os.path.join(HERE, "..", "include")):
        for f in os.listdir(root):
            m = max(m, os.path.getmtime(os.path.join(root, f)))
    return m


LAST_BUILD = {"action": None, "path": None, "seconds": 0.0}  # what the last build_library() call did: "compiled" or "reused"


def build_library(force: bool = False, verbose: bool = False, debug: bool = False) -> str:
    """Compiles every csrc/*.cu for sm_100a and links the shared library.  An up-to-date library (newer than every source and
    header) is reused unless `force`; LAST_BUILD says which happened so callers can report it."""
    import time
    lib_path = LIB_PATH_DEBUG if debug else LIB_PATH
    obj_dir = OBJ_DIR + ("_dbg" if debug else "")
    if not force and os.path.isfile(lib_path) and os.path.getmtime(lib_path) >= _newest_src_mtime():
        LAST_BUILD.update(action="reused", path=lib_path, seconds=0.0)
        return lib_path
    t0 = time.time()
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(obj_dir, exist_ok=True)

    def compile_one(src: str) -> str:
        obj = os.path.join(obj_dir, src[:-3] + ".o")
        cmd = [NVCC, *FLAGS, *(["-DQ3TTS_DEBUG"] if debug else []), "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
        if verbose:
            sys.stderr.write(r.stderr)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        objs = list(ex.map(compile_one, _sources()))
    r = subprocess.run([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib_path, *objs], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    LAST_BUILD.update(action="compiled", path=lib_path, seconds=time.time() - t0)
    return lib_path


if __name__ == "__main__":
    p = build_library(force="--force" in sys.argv, verbose="-v" in sys.argv, debug="--debug" in sys.argv)
    print(f"{LAST_BUILD['action']} {p}" + (f" in {LAST_BUILD['seconds']:.0f} s" if LAST_BUILD["action"] == "compiled" else ""))
