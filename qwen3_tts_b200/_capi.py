"""ctypes binding of libq3tts_decode.so (include/q3tts_decode.h).  No torch types cross this boundary:
only raw pointers, sizes and a CUDA stream handle.  There is NO CPU fallback: if the library or a
CUDA device is missing, calls raise."""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

from .configuration import Qwen3TTSTokenizerV2DecoderConfig

_LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libq3tts_decode.so")
_lib: Optional[C.CDLL] = None


def set_library_path(path: str) -> None:
    """tools/ only (same-box A/B of two builds, the -DQ3TTS_DEBUG build): load another build of the SAME CUDA library.
    An explicit call, made before the first load; the product never reads an environment variable to find its library."""
    global _LIB_PATH
    if _lib is not None:
        raise RuntimeError("the library is already loaded")
    _LIB_PATH = os.path.abspath(path)


OK, BAD_SHAPE, BAD_CONFIG, CODE_OUT_OF_RANGE, CUDA_ERROR, OOM, BAD_STREAM, BAD_KEY, NOT_READY = 0, -1, -2, -3, -4, -5, -6, -7, -8
MODE_FP16, MODE_BF16 = 0, 1
# bit 0 = operand type, upper bits = how many operands are hi + lo pairs (include/q3tts_decode.h)
MODES = {"fp16": 0, "bf16": 1, "fp16-plain": 2, "bf16-plain": 3, "fp16-split-all": 4, "bf16-split-all": 5}
ENGINE_TC, ENGINE_SIMT = 0, 1

# error code -> Python exception, mirroring what the reference raises on the same condition
_EXC = {
    BAD_SHAPE: ValueError,  # reference: ValueError on wrong codebook count (modeling :869-870)
    BAD_CONFIG: ValueError,
    CODE_OUT_OF_RANGE: IndexError,
    CUDA_ERROR: RuntimeError,
    OOM: MemoryError,
    BAD_STREAM: KeyError,
    BAD_KEY: KeyError,
    NOT_READY: RuntimeError,
}


class q3tts_config(C.Structure):
    _fields_ = [
        ("codebook_size", C.c_int32), ("codebook_dim", C.c_int32), ("hidden_size", C.c_int32), ("latent_dim", C.c_int32),
        ("num_attention_heads", C.c_int32), ("num_key_value_heads", C.c_int32), ("head_dim", C.c_int32),
        ("sliding_window", C.c_int32), ("intermediate_size", C.c_int32), ("num_hidden_layers", C.c_int32),
        ("num_quantizers", C.c_int32), ("decoder_dim", C.c_int32),
        ("num_upsample_rates", C.c_int32), ("upsample_rates", C.c_int32 * 8),
        ("num_upsampling_ratios", C.c_int32), ("upsampling_ratios", C.c_int32 * 4),
        ("max_position_embeddings", C.c_int32), ("rms_norm_eps", C.c_float), ("rope_theta", C.c_float),
    ]


class q3tts_encoder_config(C.Structure):
    _fields_ = [("num_attention_heads", C.c_int32), ("sliding_window", C.c_int32), ("max_position_embeddings", C.c_int32),
                ("rope_theta", C.c_float), ("norm_eps", C.c_float)]


EXPORTS = {
    # name: (restype, argtypes)
    "q3tts_create": (C.c_int, [C.POINTER(q3tts_config), C.c_int, C.POINTER(C.c_void_p)]),
    "q3tts_destroy": (None, [C.c_void_p]),
    "q3tts_last_error": (C.c_char_p, [C.c_void_p]),
    "q3tts_load_tensor": (C.c_int, [C.c_void_p, C.c_char_p, C.c_void_p, C.c_int, C.POINTER(C.c_int64), C.c_int]),
    "q3tts_finalize_weights": (C.c_int, [C.c_void_p, C.c_int]),
    "q3tts_reserve": (C.c_int, [C.c_void_p, C.c_int]),
    "q3tts_set_engine": (C.c_int, [C.c_void_p, C.c_int]),
    "q3tts_forward_num_samples": (C.c_int64, [C.c_void_p, C.c_int]),
    "q3tts_chunked_num_samples": (C.c_int64, [C.c_void_p, C.c_int, C.c_int, C.c_int]),
    "q3tts_rvq_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "q3tts_rvq_encode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "q3tts_set_encoder_config": (C.c_int, [C.c_void_p, C.POINTER(q3tts_encoder_config)]),
    "q3tts_encode_num_frames": (C.c_int64, [C.c_void_p, C.c_int64]),
    "q3tts_encode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "q3tts_encode_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "q3tts_static_max_codes": (C.c_int64, []),
    "q3tts_rvq_decode_i16_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "q3tts_forward": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "q3tts_chunked_decode": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]),
    "q3tts_decode_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "q3tts_decode_host_pcm16": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "q3tts_set_check_codes": (C.c_int, [C.c_void_p, C.c_int]),
    "q3tts_forward_probe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_void_p, C.c_int64,
                                      C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p]),
    "q3tts_stream_reserve": (C.c_int, [C.c_void_p, C.c_int, C.c_int]),
    "q3tts_stream_open": (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.c_void_p]),
    "q3tts_stream_reset": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    "q3tts_stream_close": (C.c_int, [C.c_void_p, C.c_int]),
    "q3tts_stream_frames": (C.c_int64, [C.c_void_p, C.c_int]),
    "q3tts_stream_state_bytes": (C.c_int64, [C.c_void_p]),
    "q3tts_stream_step": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.c_void_p]),
    "q3tts_launch_count": (C.c_int64, [C.c_void_p]),
    "q3tts_set_profile": (C.c_int, [C.c_void_p, C.c_int]),
    "q3tts_get_profile": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_float), C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_int32)]),
}

# include/q3tts_debug.h: self-tests, microbenchmarks, timeline (tests/ and tools/ only)
DEBUG_EXPORTS = {
    "q3tts_debug_set_pdl": (C.c_int, [C.c_int]),
    "q3tts_selftest_snake": (C.c_int, [C.c_int, C.c_void_p, C.c_int, C.c_float, C.c_float, C.c_void_p]),
    "q3tts_unit_trace": (C.c_int, [C.POINTER(C.c_longlong)]),
    "q3tts_bench_mma": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "q3tts_bench_tma": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "q3tts_selftest_pair": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "q3tts_selftest_unit": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float)]),
    "q3tts_selftest_gemm": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
}


def lib_path() -> str:
    return _LIB_PATH


def load_library() -> C.CDLL:
    """Load the in-tree CUDA extension.  Fails loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(_LIB_PATH):
        raise RuntimeError(
            f"{_LIB_PATH} is missing: build it with `python -m qwen3_tts_b200.build` (or __graft_entry__.build()). "
            "There is no CPU fallback for the decode path.")
    lib = C.CDLL(_LIB_PATH)
    for name, (res, args) in list(EXPORTS.items()) + list(DEBUG_EXPORTS.items()):
        fn = getattr(lib, name)  # AttributeError if a declared symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def make_config_struct(cfg: Qwen3TTSTokenizerV2DecoderConfig) -> q3tts_config:
    s = q3tts_config()
    for k in ("codebook_size", "codebook_dim", "hidden_size", "latent_dim", "num_attention_heads", "num_key_value_heads", "head_dim",
              "sliding_window", "intermediate_size", "num_hidden_layers", "num_quantizers", "decoder_dim", "max_position_embeddings"):
        setattr(s, k, int(getattr(cfg, k)))
    if len(cfg.upsample_rates) > 8 or len(cfg.upsampling_ratios) > 4:
        raise ValueError("too many upsample stages")
    s.num_upsample_rates = len(cfg.upsample_rates)
    for i, r in enumerate(cfg.upsample_rates):
        s.upsample_rates[i] = int(r)
    s.num_upsampling_ratios = len(cfg.upsampling_ratios)
    for i, r in enumerate(cfg.upsampling_ratios):
        s.upsampling_ratios[i] = int(r)
    s.rms_norm_eps = float(cfg.rms_norm_eps)
    s.rope_theta = float(cfg.rope_theta)
    return s


def check(handle, rc: int, what: str = "") -> None:
    if rc == OK:
        return
    msg = ""
    if handle:
        raw = load_library().q3tts_last_error(handle)
        msg = raw.decode("utf-8", "replace") if raw else ""
    raise _EXC.get(rc, RuntimeError)(f"q3tts {what} failed ({rc}): {msg}")
