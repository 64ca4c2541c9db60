#!/usr/bin/env python
"""Benchmark of the Tokenizer-12Hz decode path (codes -> 24 kHz waveform).

    python bench.py --gpus N --steps K --warmup W [--impl reference]

A "step" is one bulk decode of BASELINE.json config 2: 64 utterances x 10 s (125 code frames, 16
codebooks) of seeded synthetic codes through Qwen3TTSTokenizerV2Model.decode semantics
(chunked_decode(300, 25) + the length rule) on seeded random-init weights of the named
architecture (no checkpoint is mounted).  With N > 1 every rank decodes its own 64 utterances
(weak scaling, no data-path collective; NCCL is used only for the timing barrier / max).

  value  : audio-seconds decoded per second, codes already resident in HBM (CUDA events)
  e2e    : same metric through the public API Qwen3TTSTokenizer.decode with HOST codes in pinned
           memory: pad, H2D, decode, length rule, D2H into host float32 arrays, every step
  roofline : per-kernel CUDA-event times recorded INSIDE the timed steps (the engine's accumulating profile), dominant kernel =
           the one with the largest share of the step; traffic is null here (DRAM bytes need ncu: profiles/r2_ncu_*.txt)
  cpu_baseline / --impl reference : the UNMODIFIED reference decoder (oracle/_ref staged by oracle/stage_ref.py, imported through
           oracle/ref_shim.py) on the host cores, 4 utterances of the workload in one batch per step
  torch_gpu : context only - the same unmodified reference in eager PyTorch on this B200 (bulk decode and the reference-style
           streaming chunk call), never the product path
  clocks / gpu_launches : see DESIGN.md "Measurement"
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

B_UTT, T_FRAMES = 64, 125
N_SAMPLES = 239445  # T*1920 - 555 per 300-frame chunk (modeling :885-895)
WORKLOADS = {  # name -> (utterances per GPU per step, frames per utterance, BASELINE.json config it stands for)
    "bulk10s": (64, 125, "Tokenizer-12Hz non-streaming bulk decode, batch 64 x 10 s synthetic 16-codebook codes per GPU (BASELINE config 2)"),
    "longform60s": (32, 750, "Long-form decode, batch 32 x 60 s codes per GPU through chunked_decode (chunks 300 / 25+300 / 25+150; BASELINE config 4)"),
    "sharded30s": (64, 375, "strong-scaling form of BASELINE config 5: see run_sharded"),
    "bulk30s": (64, 375, "Offline bulk decode of 30 s utterances, one batch of 64 per GPU per step (4096 utterances = 64 such steps at N=1, "
                         "utterance-sharded over N GPUs; chunks 300 / 25+75; BASELINE config 5)"),
}
WORKLOADS_NAME = WORKLOADS["bulk10s"][2]
FRAME_SEC = 0.08
GFLOP_PER_FRAME = 5.1226  # 2 x 2561.30 MMAC per emitted code frame


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        d = json.load(open(p))
        return {"hbm_gbs": d["hbm_gbs"], "tflops_burst": d["bf16_tflops"], "tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                "source": "measured"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """Samples SM clock / throttle reasons while the timed region runs: NVML in-process (a few microseconds per
    query, every 50 ms); a polling nvidia-smi process is the fallback (its queries stall host-side CUDA calls,
    which showed up as +8 ms per synchronous e2e step)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index: int):
        self.index, self.rows, self.proc, self.nv, self.run = index, [], None, None, False
        self.sm, self.mx, self.reasons, self.source = [], None, set(), None

    def start(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[self.index]) if vis and all(v.strip().isdigit() for v in vis.split(",")) else self.index
            self.h = nv.nvmlDeviceGetHandleByIndex(idx)
            self.mx = float(nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM))
            self.nv, self.run, self.source = nv, True, "nvml"
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
            return
        except Exception:
            self.nv = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.source = "nvidia-smi"
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _poll(self):
        nv = self.nv
        names = (("hw_slowdown", nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else nv.nvmlClocksThrottleReasonHwSlowdown),
                 ("hw_thermal_slowdown", nv.nvmlClocksEventReasonHwThermalSlowdown if hasattr(nv, "nvmlClocksEventReasonHwThermalSlowdown") else nv.nvmlClocksThrottleReasonHwThermalSlowdown),
                 ("sw_thermal_slowdown", nv.nvmlClocksEventReasonSwThermalSlowdown if hasattr(nv, "nvmlClocksEventReasonSwThermalSlowdown") else nv.nvmlClocksThrottleReasonSwThermalSlowdown),
                 ("sw_power_cap", nv.nvmlClocksEventReasonSwPowerCap if hasattr(nv, "nvmlClocksEventReasonSwPowerCap") else nv.nvmlClocksThrottleReasonSwPowerCap))
        while self.run:
            try:
                self.sm.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                get = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
                mask = get(self.h)
                for name, bit in names:
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.05)

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.nv:
            self.run = False
            self.thread.join(timeout=2)
            return {"sm_mhz": statistics.median(self.sm) if self.sm else None, "sm_max_mhz": self.mx, "reasons": sorted(self.reasons),
                    "samples": len(self.sm), "source": self.source}
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm),
                "source": self.source}


# ------------------------------------------------------------------------------------------------ reference (baseline legs only)
REF_UTTERANCES = 4  # bounded sample of the workload the CPU legs decode per step, in ONE batch


def load_reference(device: str = "cpu"):
    """The reference's own decoder for the baseline legs: the UNMODIFIED reference module (oracle/_ref, staged byte for byte by
    oracle/stage_ref.py; /root/reference in the build container) through oracle/ref_shim.py -> kind "reference"; the oracle port
    (pinned to it by tests/golden) only if the staged files are missing -> kind "port".  Returns (decode_fn(codes [B,T,Q]) ->
    list of waveforms, kind).  Never used by the product path."""
    import torch

    from oracle import decoder_oracle as O
    from oracle import ref_shim
    from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
    from qwen3_tts_b200.synthetic import make_decoder_state_dict
    import contextlib
    cfg = Qwen3TTSTokenizerV2DecoderConfig()
    sd = make_decoder_state_dict(cfg, seed=0)
    if ref_shim.available():
        with contextlib.redirect_stdout(sys.stderr):  # transformers prints docstring complaints about the reference on stdout: one JSON line only
            dec = ref_shim.build_reference_decoder(sd, codebook_dim=cfg.codebook_dim, attn_implementation="sdpa").to(device)

        def decode(codes):
            with contextlib.redirect_stdout(sys.stderr):
                return ref_shim.reference_model_decode(dec, codes.to(device))
        return decode, "reference", dec
    sd = {k: v.to(device) for k, v in sd.items()}

    def decode_port(codes):
        with torch.inference_mode():
            return O.model_decode(sd, cfg, codes.to(device))
    return decode_port, "port", None


def cpu_reference_run(steps: int, warmup: int, utterances: int = REF_UTTERANCES):
    """The reference's CPU decode of `utterances` items of the workload in one batch per step, on all host cores."""
    import torch

    from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
    from qwen3_tts_b200.synthetic import make_codes
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cfg = Qwen3TTSTokenizerV2DecoderConfig()
    decode, kind, _ = load_reference("cpu")
    codes = make_codes(B_UTT, T_FRAMES, cfg, seed=1)[:utterances]
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        out = decode(codes)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    assert len(out) == utterances and out[0].shape[0] == N_SAMPLES
    sec = sum(times) / len(times)
    return {"value": utterances * T_FRAMES * FRAME_SEC / sec, "sec_per_step": sec, "cores": torch.get_num_threads(), "kind": kind,
            "sample": f"{utterances} of the {B_UTT} utterances x {T_FRAMES * FRAME_SEC:g} s in one batch per step, mean of {steps} steps after {warmup} warm-up"}


def run_reference(args, rank: int):
    """--impl reference: the reference's own CPU implementation of the path on the host cores, same metric / config / unit as
    our arm; each step is a bounded sample (REF_UTTERANCES utterances in one batch) of the workload."""
    if rank != 0:
        return
    r = cpu_reference_run(steps=args.steps, warmup=args.warmup)
    line = {"impl": "reference", "metric": "audio-sec decoded per sec (codes->24kHz)", "value": r["value"], "unit": "audio-s/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["sec_per_step"] * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
            "config": bench_config(args.gpus),
            "cpu_baseline": {"value": r["value"], "unit": "audio-s/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
            "e2e": {"value": r["value"], "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def bench_config(world: int):
    """`config` of the JSON line: identical for our arm and the reference arm (the driver compares them)."""
    return {"workload": WORKLOADS_NAME, "frames_per_step_per_gpu": B_UTT * T_FRAMES,
            "weights": "seeded random-init, reference architecture (195.08 M params)",
            "l2": "flushed between timed steps (256 MiB write, untimed); activations per step >> 126 MB L2",
            "parallelism": f"utterance-sharded x{world}, no collective"}


def torch_gpu_context(dev, streams: int):
    """Context numbers SURVEY 8(d) / BASELINE.md 4 ask for: the unmodified reference in eager PyTorch (cuDNN / cuBLAS, fp32) on the
    SAME B200 - (a) bulk decode of a sub-batch of config 2, (b) the reference-style streaming chunk call
    forward(codes[..., s-25:e])[..., 25*1920:] (modeling :889-893) for `streams` streams with one new frame.  Baseline only."""
    import torch

    from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
    from qwen3_tts_b200.synthetic import make_codes
    cfg = Qwen3TTSTokenizerV2DecoderConfig()
    decode, kind, ref_dec = load_reference(str(dev))
    nb = 16
    codes = make_codes(B_UTT, T_FRAMES, cfg, seed=1)[:nb].to(dev)

    def timed(fn, reps, warm):
        for _ in range(warm):
            fn()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize(dev)
            ts.append(a.elapsed_time(b))
        return sorted(ts)

    bulk = timed(lambda: decode(codes), 5, 2)
    out = {"impl": f"{kind} (unmodified reference module, eager PyTorch fp32, library defaults: cuDNN / cuBLAS)" if kind == "reference" else "oracle port, eager PyTorch fp32",
           "bulk": {"batch": nb, "frames": T_FRAMES, "ms_p50": bulk[len(bulk) // 2], "audio_s_per_s": nb * T_FRAMES * FRAME_SEC / (bulk[len(bulk) // 2] * 1e-3)}}
    if ref_dec is not None:
        ch = make_codes(streams, 26, cfg, seed=2).transpose(1, 2).contiguous().to(dev)   # 25 frames of left context + 1 new frame

        def chunk_call():
            import contextlib
            with torch.inference_mode(), contextlib.redirect_stdout(sys.stderr):
                return ref_dec(ch)[..., 25 * 1920:]
        c = timed(chunk_call, 20, 3)
        out["stream_chunk_call"] = {"streams": streams, "frames_per_call": 26, "new_frames": 1, "ms_p50": c[len(c) // 2], "ms_p99": c[-1],
                                    "note": "reference-style chunk: forward(codes[..., s-25:e])[..., 25*1920:], 25 frames recomputed per packet"}
    del decode, ref_dec
    torch.cuda.empty_cache()
    return out


def stream_latency(dec, cfg, dev, streams: int, packet: int, steps: int, warmup: int):
    """BASELINE config 3: `streams` concurrent streams, one packet of `packet` frames each per step, left-context state on the device.

    Per step: device time of q3tts_stream_step (CUDA events) and host wall time of the whole chunk (pinned codes -> device, step, waveform
    -> pinned host, synchronize).  The timed steps are steady-state (continuing) steps that cross the K/V ring wrap."""
    import torch

    from qwen3_tts_b200.synthetic import make_codes
    total = warmup + steps
    frames = 1000                                                                                   # SURVEY cfg 3: (256, 16, 1000) seeded codes, cycled
    codes_h = make_codes(streams, frames, cfg, seed=77).transpose(1, 2).contiguous()               # [S,Q,1000]
    ring = [codes_h[:, :, t:t + packet].contiguous().pin_memory() for t in range(0, frames - packet + 1, packet)]
    pk_h = [ring[i % len(ring)] for i in range(total)]
    pk_d = torch.empty_like(pk_h[0], device=dev)
    wav_d = torch.empty((streams, packet * 1920), dtype=torch.float32, device=dev)
    wav_h = torch.empty((streams, packet * 1920), dtype=torch.float32).pin_memory()
    ids = [dec.open_stream() for _ in range(streams)]
    dev_ms, host_ms = [], []
    l0 = dec.launch_count()
    for i in range(total):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        pk_d.copy_(pk_h[i], non_blocking=True)
        a.record()
        dec.decode_step(ids, pk_d, out=wav_d)
        b.record()
        wav_h.copy_(wav_d, non_blocking=True)
        torch.cuda.synchronize(dev)
        t1 = time.perf_counter()
        if i >= warmup:
            dev_ms.append(a.elapsed_time(b)); host_ms.append((t1 - t0) * 1e3)
    launches = (dec.launch_count() - l0) // total
    for i in ids:
        dec.close_stream(i)
    q = lambda v, f: float(sorted(v)[min(len(v) - 1, int(f * len(v)))])
    chunk_audio_s = packet * FRAME_SEC
    return {"streams": streams, "frames_per_packet": packet, "chunk_audio_ms": chunk_audio_s * 1e3, "steps": steps, "warmup": warmup,
            "device_ms": {"p50": q(dev_ms, 0.5), "p99": q(dev_ms, 0.99), "max": max(dev_ms)},
            "e2e_ms": {"p50": q(host_ms, 0.5), "p99": q(host_ms, 0.99), "max": max(host_ms),
                       "includes": "pinned codes h2d, step, waveform d2h to pinned host, synchronize"},
            "realtime_factor_p50": streams * chunk_audio_s / (q(host_ms, 0.5) * 1e-3), "launches_per_step": launches,
            "h2d_bytes_per_step": streams * 16 * packet * 8, "d2h_bytes_per_step": streams * packet * 1920 * 4}


def run_sharded(args, rank, local_rank, world):
    """--workload sharded30s: BASELINE config 5 as STRONG scaling - a FIXED set of `--utterances` ragged utterances (true lengths
    U[0.5, 1] x 375 frames, seeded) is bucketed by length GLOBALLY and the batches are dealt to the ranks longest-first
    queue order (qwen3_tts_b200.sharding.global_batches); the ranks PULL batches through an atomic counter in the launcher's store;
    every rank decodes its shard end to end (host codes -> host waveforms through Qwen3TTSTokenizer.decode); no data-path collective.
    A step = the whole set once; value = audio-seconds of the whole set / max-over-ranks wall time of the step."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from qwen3_tts_b200 import Qwen3TTSTokenizer, synthetic_model
    from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
    from qwen3_tts_b200.sharding import batch_cost, decode_shard, global_batches
    from qwen3_tts_b200.synthetic import make_codes
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)   # timing barrier / max only (the launch contract); the decode path uses no collective
    cfg = Qwen3TTSTokenizerV2DecoderConfig()
    n_utt, Tmax = args.utterances, 375
    lengths = [int(v) for v in np.random.default_rng(5).integers(Tmax // 2, Tmax + 1, n_utt)]
    max_frames = 64 * Tmax
    batches = global_batches(lengths, world, max_frames)            # global length buckets, most expensive first: the shared work queue
    model = synthetic_model(seed=0, device=dev, compute_mode=args.mode)
    model.decoder.reserve(64 * 325)
    tok = Qwen3TTSTokenizer.from_model(model)
    # every rank may pull any batch, so every rank holds all codes (147 MB of int64 for 4096 utterances) in pinned memory
    dense = [make_codes(1, lengths[i], cfg, seed=1000 + i)[0].pin_memory() for i in range(n_utt)]
    for bt in batches:                                                   # warm-up: one untimed pass over every batch shape (launch plans,
        decode_shard(tok, dense, [], max_frames, batches=[bt])           # workspace, pinned result buffers of the host allocator)
    store = dist.distributed_c10d._get_default_store() if world > 1 else None   # the launcher's key-value store: host-side, no collective
    times = []
    l0 = model.decoder.launch_count()
    # breakdown: seconds inside the C decode call (H2D, kernels, direct D2H, synchronise) vs the whole Python step of this rank
    in_call = [0.0]
    _orig_decode_host = model.decode_host

    def _timed_decode_host(*a, **k):
        t = time.perf_counter()
        r = _orig_decode_host(*a, **k)
        in_call[0] += time.perf_counter() - t
        return r
    model.decode_host = _timed_decode_host
    my_cost, n_out = 0.0, 0
    for step in range(args.steps):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        taken = 0
        while True:
            # pull the next batch: an atomic counter per step (rank order does not matter, a faster GPU simply takes more batches)
            k = (store.add(f"q3tts_step{step}", 1) - 1) if store is not None else taken
            if k >= len(batches):
                break
            out = decode_shard(tok, dense, [], max_frames, batches=[batches[k]])
            n_out += sum(int(w.shape[0]) for w in out.values())
            my_cost += batch_cost(batches[k], lengths)
            taken += 1
        torch.cuda.synchronize(dev)
        times.append(time.perf_counter() - t0)
    launches = model.decoder.launch_count() - l0
    tt = torch.tensor(times, dtype=torch.float64, device=dev)
    load = torch.tensor([my_cost / max(1, args.steps), in_call[0], float(sum(times))], dtype=torch.float64, device=dev)
    loads = [load.clone() for _ in range(world)]
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dist.all_gather(loads, load)
    if rank == 0:
        sec = float(tt.sum().item())
        audio_s = sum(lengths) * FRAME_SEC * args.steps
        ld = [float(x[0].item()) for x in loads]
        call_frac = [float(x[1].item()) / max(float(x[2].item()), 1e-9) for x in loads]
        h2d = sum(lengths) * 16 * 8
        line = {"metric": "audio-sec decoded per sec (codes->24kHz)", "value": audio_s / sec, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps,
                "warmup": 1, "ms_per_step": sec / args.steps * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": f"{args.mode} tensor-core operands, fp32 accumulate", "data": "synthetic",
                "config": {"workload": f"Offline bulk decode of a FIXED set of {n_utt} ragged utterances (true lengths U[0.5,1] x 30 s, 16-codebook codes) "
                                       f"sharded by utterance over {world} GPU(s): global length buckets pulled longest-first from one queue (BASELINE config 5)",
                           "audio_seconds_per_step": sum(lengths) * FRAME_SEC, "batches": len(batches),
                           "frames_cost_max_over_mean": max(ld) / (sum(ld) / len(ld)), "parallelism": f"utterance-sharded x{world}, no collective",
                           "l2": "inputs and activations per step >> 126 MB L2"},
                "e2e": {"value": audio_s / sec, "unit": "audio-s/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(sum(min(n * 1920, 375 * 1920 - 1110) for n in lengths)) * 4,
                        "api": "qwen3_tts_b200.sharding.decode_shard -> Qwen3TTSTokenizer.decode (pinned host codes -> host float32 arrays), the whole step"},
                "gpu_launches": launches,
                "per_rank": {"seconds": [float(x[2].item()) for x in loads],
                             "fraction_inside_decode_call": [round(v, 4) for v in call_frac],
                             "note": "decode call = q3tts_decode_host (H2D, kernels, direct D2H, synchronise); the rest is Python host work between batches"}}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


KERNEL_NAMES = {0: "k_rvq", 1: "conv_gemm_tc_kernel", 2: "k_rmsnorm", 3: "k_attention", 4: "k_dwconv_ln", 5: "tail_conv_kernel", 6: "unit_kernel",
                7: "ws_conv_kernel", 8: "slab_conv_kernel", 9: "k_stream_attention", 10: "k_state_copy", 11: "k_split_f32",
                12: "k_enc_conv0", 13: "k_act_split", 14: "k_layernorm_split", 15: "k_enc_attention", 16: "k_enc_to_cf", 17: "k_enc_split_cols",
                18: "k_enc_argmin_update"}
TENSOR_KERNELS = ("unit_kernel", "slab_conv_kernel", "conv_gemm_tc_kernel", "ws_conv_kernel")



# ------------------------------------------------------------------------------------------------ encode path
ENC_UTT, ENC_SEC = 64, 10.0
ENC_NAME = "Tokenizer-12Hz encode, batch 64 x 10 s synthetic 24 kHz audio per GPU -> 16-codebook codes (the mirror of BASELINE config 2)"
ENC_METRIC = "audio-sec encoded per sec (24kHz->codes)"


def encode_reference_cpu(steps: int, warmup: int, utterances: int = REF_UTTERANCES):
    """The reference's own encoder class (Qwen3TTSTokenizerV2Encoder = HF MimiModel, through oracle/ref_shim.py) + the reference's
    model.encode post-processing on the host cores: `utterances` items of the workload in one batch per step."""
    import torch

    from oracle import ref_shim
    from qwen3_tts_b200.synthetic import make_audio, make_encoder_state_dict
    torch.set_num_threads(os.cpu_count() or 1)
    with contextlib.redirect_stdout(sys.stderr):  # transformers prints docstring complaints about the reference on stdout: one JSON line only
        enc = ref_shim.build_reference_encoder(make_encoder_state_dict(seed=0), attn_implementation="sdpa")
    T = int(ENC_SEC * 24000)
    wav = make_audio(utterances, T, seed=1)
    mask = torch.ones_like(wav, dtype=torch.long)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(sys.stderr):
            out = ref_shim.reference_model_encode(enc, wav, mask)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    assert len(out) == utterances and tuple(out[0].shape) == (int(ENC_SEC * 12.5), 16)
    sec = sum(times) / len(times)
    return {"value": utterances * ENC_SEC / sec, "sec_per_step": sec, "cores": torch.get_num_threads(), "kind": "reference",
            "sample": f"{utterances} of the {ENC_UTT} utterances x {ENC_SEC:g} s in one batch per step, mean of {steps} steps after {warmup} warm-up"}


def encode_config(world: int):
    return {"workload": ENC_NAME, "samples_per_step_per_gpu": int(ENC_UTT * ENC_SEC * 24000),
            "weights": "seeded random-init, Mimi encoder architecture (MimiConfig defaults, 16 of 32 codebooks)",
            "l2": "flushed between timed steps (256 MiB write, untimed); activations per step >> 126 MB L2",
            "parallelism": f"utterance-sharded x{world}, no collective"}


def encode_section(model, dev, steps: int, warmup: int, world: int = 1, rank: int = 0):
    """Times encode on `model` (needs the encoder weights): device-resident value with the per-kernel profile recorded inside the
    timed steps, and e2e through the host call (pinned host waveforms in, host codes out)."""
    import torch

    from qwen3_tts_b200.synthetic import make_audio
    T = int(ENC_SEC * 24000)
    wav_host = make_audio(ENC_UTT, T, seed=1 + rank)
    wav = wav_host.to(dev)
    arrays = [wav_host[b].numpy() for b in range(ENC_UTT)]
    dec = model.decoder
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(max(3, warmup)):
        model.encode(wav)
    torch.cuda.synchronize(dev)
    dec.set_profile(True)
    dec.get_profile()
    l0 = dec.launch_count()
    step_ms, acc = [], {}
    for _ in range(steps):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        model.encode(wav)
        b.record()
        prof = dec.get_profile()
        step_ms.append(a.elapsed_time(b))
        for code, t_ms, fl, by, dims in prof:
            k = acc.setdefault(KERNEL_NAMES.get(code, str(code)), {"ms": 0.0, "launches": 0, "flops": 0.0, "min_bytes": 0.0})
            k["ms"] += t_ms; k["launches"] += 1; k["flops"] += fl; k["min_bytes"] += by
    dec.set_profile(False)
    launches = dec.launch_count() - l0
    for _ in range(3):
        codes = model.encode_host(arrays)
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(steps):
        codes = model.encode_host(arrays)
    e2e_s = time.perf_counter() - t0
    assert len(codes) == ENC_UTT and codes[0].shape == (int(ENC_SEC * 12.5), 16)
    return {"step_ms": step_ms, "acc": acc, "launches": launches, "e2e_s": e2e_s, "frames": int(ENC_SEC * 12.5)}


def kernel_table(acc, steps, peaks):
    ksum = sum(k["ms"] for k in acc.values())
    out = {}
    for name, k in sorted(acc.items(), key=lambda kv: -kv[1]["ms"]):
        sec = k["ms"] * 1e-3
        tfl = k["flops"] / sec / 1e12 if sec > 0 else 0.0
        gbs = k["min_bytes"] / sec / 1e9 if sec > 0 else 0.0
        ft, fh = tfl / peaks["tflops_sustained"], gbs / peaks["hbm_gbs"]
        tensor = name in TENSOR_KERNELS and ft >= fh
        out[name] = {"ms_per_step": round(k["ms"] / steps, 4), "launches_per_step": k["launches"] // steps, "share_of_step": round(k["ms"] / ksum, 4),
                     "bound": "tensor" if tensor else "hbm", "achieved": round(tfl if tensor else gbs, 1), "unit": "TFLOP/s" if tensor else "GB/s",
                     "frac": round(ft if tensor else fh, 4)}
    return out, ksum


def run_encode(args, rank, local_rank, world):
    """--workload encode10s: one JSON line for the ENCODE path (same contract as the decode line; its own metric)."""
    if args.impl == "reference":
        if rank != 0:
            return
        r = encode_reference_cpu(steps=args.steps, warmup=args.warmup)
        print(json.dumps({"impl": "reference", "metric": ENC_METRIC, "value": r["value"], "unit": "audio-s/s", "n_gpus": args.gpus, "steps": args.steps,
                          "warmup": args.warmup, "ms_per_step": r["sec_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                          "dtype": "fp32", "data": "synthetic", "config": encode_config(args.gpus),
                          "cpu_baseline": {"value": r["value"], "unit": "audio-s/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]},
                          "e2e": {"value": r["value"], "unit": "audio-s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}), flush=True)
        return
    import torch
    import torch.distributed as dist

    from qwen3_tts_b200 import synthetic_model
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the encode path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)
    args.warmup = max(args.warmup, 3)
    model = synthetic_model(device=dev, with_encoder=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    sampler = ClockSampler(local_rank)
    sampler.start()
    r = encode_section(model, dev, args.steps, args.warmup, world, rank)
    clocks = sampler.stop()
    ms, e2e_s = sum(r["step_ms"]), r["e2e_s"]
    if world > 1:
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s = t[0].item(), t[1].item()
    peaks = load_peaks()
    kernels, ksum = kernel_table(r["acc"], args.steps, peaks)
    if rank == 0:
        audio_s = world * ENC_UTT * ENC_SEC * args.steps
        dom = next(iter(kernels))
        d, kd = r["acc"][dom], kernels[dom]
        tensor = kd["bound"] == "tensor"
        alg = d["flops"] if tensor else d["min_bytes"]
        roof = {"bound": kd["bound"], "kernel": dom, "achieved": alg / (d["ms"] * 1e-3) / (1e12 if tensor else 1e9),
                "peak": peaks["tflops_sustained"] if tensor else peaks["hbm_gbs"], "unit": kd["unit"], "frac": kd["frac"], "traffic": None,
                "peak_source": f"{peaks['source']} " + ("sustained bf16 cuBLAS" if tensor else "copy bandwidth"),
                "share_of_step": d["ms"] / ksum, "launches_per_step": d["launches"] // args.steps,
                "algorithmic_per_launch": alg / d["launches"], "ms_per_launch_avg": d["ms"] / d["launches"],
                "measured": f"CUDA-event pairs around every launch inside the {args.steps} timed steps",
                "sum_kernel_ms_per_step": ksum / args.steps, "sum_kernel_over_step": ksum / sum(r["step_ms"])}
        total_bytes = sum(k["min_bytes"] for k in r["acc"].values()) / args.steps
        line = {"metric": ENC_METRIC, "value": audio_s / (ms * 1e-3), "unit": "audio-s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "fp16 hi + lo operand pairs on the tensor cores (3 MMAs per product), fp32 accumulate, fp32 activations", "data": "synthetic",
                "config": encode_config(world),
                "e2e": {"value": audio_s / e2e_s, "unit": "audio-s/s", "ms_per_step": e2e_s / args.steps * 1e3,
                        "h2d_bytes_per_step": world * ENC_UTT * int(ENC_SEC * 24000) * 4, "d2h_bytes_per_step": world * ENC_UTT * r["frames"] * 16 * 8,
                        "api": "Qwen3TTSTokenizerV2Model.encode_host(list of float32 waveforms) -> list of int64 [frames, 16] arrays (q3tts_encode_host)"},
                "gpu_launches": r["launches"], "clocks": clocks, "roofline": roof,
                "path_frac_of_hbm_roofline": total_bytes / (ms / args.steps * 1e-3) / 1e9 / peaks["hbm_gbs"], "min_hbm_gb_per_step": total_bytes / 1e9,
                "kernels": kernels}
        if world == 1:
            # latency of ONE 10 s item (the voice-clone reference clip, model wrapper :427): launch-bound, not throughput-bound
            from qwen3_tts_b200.synthetic import make_audio
            one = make_audio(1, int(ENC_SEC * 24000), seed=5).to(dev)
            for _ in range(5):
                model.encode(one)
            lat = []
            for _ in range(30):
                a, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record(); model.encode(one); b2.record(); torch.cuda.synchronize(dev)
                lat.append(a.elapsed_time(b2))
            lat.sort()
            line["single_item"] = {"seconds": ENC_SEC, "device_ms_p50": lat[len(lat) // 2], "device_ms_max": lat[-1]}
        if not args.no_torch_gpu and world == 1:
            try:   # context: the reference's own encoder class in eager PyTorch (fp32, library defaults) on this GPU
                from oracle import ref_shim
                from qwen3_tts_b200.synthetic import make_encoder_state_dict
                del model
                torch.cuda.empty_cache()
                with contextlib.redirect_stdout(sys.stderr):
                    enc = ref_shim.build_reference_encoder(make_encoder_state_dict(seed=0), attn_implementation="sdpa").to(dev)
                nb = 16
                w = make_audio(nb, int(ENC_SEC * 24000), seed=1).to(dev)
                msk = torch.ones_like(w, dtype=torch.long)
                ts = []
                for i in range(7):
                    a, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record()
                    with contextlib.redirect_stdout(sys.stderr):
                        ref_shim.reference_model_encode(enc, w, msk)
                    b2.record(); torch.cuda.synchronize(dev)
                    if i >= 2:
                        ts.append(a.elapsed_time(b2))
                ts.sort()
                line["torch_gpu"] = {"impl": "reference encoder class (unmodified, eager PyTorch fp32, library defaults: cuDNN / cuBLAS)", "batch": nb,
                                     "ms_p50": ts[len(ts) // 2], "audio_s_per_s": nb * ENC_SEC / (ts[len(ts) // 2] * 1e-3)}
                line["torch_gpu"]["ours_over_eager"] = line["value"] / line["torch_gpu"]["audio_s_per_s"]
            except Exception as ex:  # context only
                line["torch_gpu"] = {"unavailable": f"{type(ex).__name__}: {ex}"[:300]}
        if not args.no_cpu_baseline and world == 1:
            c = encode_reference_cpu(steps=3, warmup=1)
            line["cpu_baseline"] = {"value": c["value"], "unit": "audio-s/s", "cores": c["cores"], "kind": c["kind"], "sample": c["sample"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="fp16", choices=["fp16", "bf16", "fp16-plain", "bf16-plain", "fp16-split-all", "bf16-split-all"],
                    help="tensor-core operand type (fp32 accumulate) and how many early-stage operands are hi + lo pairs (DESIGN.md Precision)")
    ap.add_argument("--workload", default="bulk10s", choices=sorted(WORKLOADS) + ["encode10s"],
                    help="default = the configuration BASELINE.json's metric is quoted on; encode10s = the encode path (SURVEY 8(f)-2), its own metric")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-streaming", action="store_true", help="skip the streaming-chunk latency section (BASELINE config 3)")
    ap.add_argument("--no-torch-gpu", action="store_true", help="skip the eager-PyTorch-on-this-GPU context section")
    ap.add_argument("--no-encode", action="store_true", help="skip the encode-path section of the default line (its full line: --workload encode10s)")
    ap.add_argument("--stream-only", action="store_true", help="print only the streaming-chunk latency section")
    ap.add_argument("--streams", type=int, default=256)
    ap.add_argument("--utterances", type=int, default=4096, help="size of the fixed utterance set of --workload sharded30s")
    args = ap.parse_args()
    if args.workload == "encode10s":
        run_encode(args, int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return
    if args.workload == "sharded30s":
        if args.impl == "reference":
            raise SystemExit("--workload sharded30s has no reference arm: use the default workload")
        run_sharded(args, int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")))
        return
    global B_UTT, T_FRAMES, N_SAMPLES, WORKLOADS_NAME
    B_UTT, T_FRAMES, WORKLOADS_NAME = WORKLOADS[args.workload]
    N_SAMPLES = T_FRAMES * 1920 - 555 * ((T_FRAMES + 299) // 300)
    if args.workload != "bulk10s":
        args.no_streaming = True  # the streaming and context sections ride on the default line only
        args.no_torch_gpu = True
        args.no_encode = True
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    args.warmup = max(args.warmup, 3)

    import torch
    import torch.distributed as dist

    from qwen3_tts_b200 import Qwen3TTSTokenizer, synthetic_model
    from qwen3_tts_b200.configuration import Qwen3TTSTokenizerV2DecoderConfig
    from qwen3_tts_b200.synthetic import make_codes, make_decoder_state_dict

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the decode path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        if os.environ.get("NCCL_DEBUG", "VERSION").upper() == "VERSION":
            os.environ["NCCL_DEBUG"] = "WARN"                                   # keep NCCL's version banner off stdout: one JSON line only
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    cfg = Qwen3TTSTokenizerV2DecoderConfig()
    sd = make_decoder_state_dict(cfg, seed=0)
    model = synthetic_model(state_dict=sd, device=dev, compute_mode=args.mode, with_encoder=(world == 1 and not args.no_encode))
    del sd
    dec = model.decoder
    if args.stream_only:
        dec.reserve_streams(args.streams, 4)
        out = {"streaming": [stream_latency(dec, cfg, dev, args.streams, P, steps=2000 // P, warmup=200 // P) for P in (1, 2, 4)]}
        print(json.dumps(out), flush=True)
        return
    dec.reserve(B_UTT * min(T_FRAMES, 325))                                  # largest single forward: 300-frame chunk + 25 of left context
    tok = Qwen3TTSTokenizer.from_model(model)
    codes_host = make_codes(B_UTT, T_FRAMES, cfg, seed=1 + rank)            # [B,T,Q] int64
    codes_dev = codes_host.to(dev).transpose(1, 2).contiguous()              # [B,Q,T] resident in HBM
    host_items = [{"audio_codes": codes_host[b].clone().pin_memory()} for b in range(B_UTT)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)            # > 126 MB L2

    def step_device():
        return dec.chunked_decode(codes_dev)

    # ---------------- device-resident throughput.  The per-kernel profile is recorded INSIDE these timed steps: the engine brackets every
    # launch with a CUDA-event pair on the launch stream (two event records per kernel; their cost is inside `value`, not hidden).
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    dec.set_profile(True)
    l0 = dec.launch_count()
    step_ms, acc = [], {}
    for _ in range(args.steps):
        flush.zero_()                                                       # L2 flush between timed iterations (untimed)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        wav = step_device()
        b.record()
        prof = dec.get_profile()                                            # synchronises; read-and-clear
        step_ms.append(a.elapsed_time(b))
        for code, t_ms, fl, by, dims in prof:
            k = acc.setdefault(KERNEL_NAMES.get(code, str(code)), {"ms": 0.0, "launches": 0, "flops": 0.0, "min_bytes": 0.0})
            k["ms"] += t_ms; k["launches"] += 1; k["flops"] += fl; k["min_bytes"] += by
    dec.set_profile(False)
    barrier()
    launches = dec.launch_count() - l0
    ms = sum(step_ms)
    # ---------------- end to end through the public API (host codes -> host float32 waveforms)
    for _ in range(max(3, args.warmup)):                                     # results held as in the timed loop, so the two pinned
        wavs, sr = tok.decode(host_items)                                    # result buffers that alternate there exist before timing starts
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        wavs, sr = tok.decode(host_items)
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    clocks = sampler.stop()
    assert sr == 24000 and len(wavs) == B_UTT and wavs[0].shape == (N_SAMPLES,) and tuple(wav.shape) == (B_UTT, 1, N_SAMPLES)
    if world > 1:
        t = torch.tensor([ms, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)                             # max over ranks
        ms, e2e_s = t[0].item(), t[1].item()
    audio_s = world * B_UTT * T_FRAMES * FRAME_SEC * args.steps
    value = audio_s / (ms / 1e3)
    e2e_value = audio_s / e2e_s
    # ---------------- roofline: per-kernel numbers averaged over the timed steps themselves
    peaks = load_peaks()
    ksum_ms = sum(k["ms"] for k in acc.values())
    kernels = {}
    for name, k in sorted(acc.items(), key=lambda kv: -kv[1]["ms"]):
        per_step = k["ms"] / args.steps
        tfl = k["flops"] / (k["ms"] * 1e-3) / 1e12 if k["ms"] > 0 else 0.0
        gbs = k["min_bytes"] / (k["ms"] * 1e-3) / 1e9 if k["ms"] > 0 else 0.0
        tensor = name in TENSOR_KERNELS
        kernels[name] = {"ms_per_step": round(per_step, 4), "launches_per_step": k["launches"] // args.steps, "share_of_step": round(k["ms"] / ksum_ms, 4),
                         "bound": "tensor" if tensor else "hbm",
                         "achieved": round(tfl if tensor else gbs, 1), "unit": "TFLOP/s" if tensor else "GB/s",
                         "frac": round((tfl / peaks["tflops_sustained"]) if tensor else (gbs / peaks["hbm_gbs"]), 4)}
        if tensor:
            kernels[name]["min_hbm_gbs"] = round(gbs, 1)
    dom = max(acc, key=lambda n: acc[n]["ms"])                               # dominant kernel = largest share of the timed steps
    d = acc[dom]
    dom_tensor = dom in TENSOR_KERNELS
    achieved = (d["flops"] if dom_tensor else d["min_bytes"]) / (d["ms"] * 1e-3) / (1e12 if dom_tensor else 1e9)
    peak = peaks["tflops_sustained"] if dom_tensor else peaks["hbm_gbs"]
    roof = {"bound": "tensor" if dom_tensor else "hbm", "kernel": dom, "achieved": achieved, "peak": peak,
            "unit": "TFLOP/s" if dom_tensor else "GB/s", "frac": achieved / peak,
            "traffic": None,  # DRAM bytes per launch need ncu (dram__bytes_read.sum + dram__bytes_write.sum): profiles/r2_ncu_*.txt of this commit
            "peak_source": f"{peaks['source']} " + ("sustained bf16 cuBLAS (kernel timed inside a long step; fp16 has the same pipe rate)" if dom_tensor else "copy bandwidth"),
            "share_of_step": d["ms"] / ksum_ms, "launches_per_step": d["launches"] // args.steps,
            "algorithmic_per_launch": (d["flops"] if dom_tensor else d["min_bytes"]) / d["launches"], "ms_per_launch_avg": d["ms"] / d["launches"],
            "measured": f"CUDA-event pairs around every launch inside the {args.steps} timed steps",
            "sum_kernel_ms_per_step": ksum_ms / args.steps, "sum_kernel_over_step": ksum_ms / sum(step_ms)}
    path_frac = (B_UTT * T_FRAMES * GFLOP_PER_FRAME * 1e9) / (ms / args.steps * 1e-3) / 1e12 / peaks["tflops_sustained"]
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {"metric": "audio-sec decoded per sec (codes->24kHz)", "value": value, "unit": "audio-s/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": f"{args.mode} tensor-core operands, fp32 accumulate", "data": "synthetic",
            "config": bench_config(world),
            "e2e": {"value": e2e_value, "unit": "audio-s/s", "h2d_bytes_per_step": world * B_UTT * T_FRAMES * 16 * 8,
                    "d2h_bytes_per_step": world * B_UTT * N_SAMPLES * 4, "ms_per_step": e2e_s / args.steps * 1e3,
                    "api": "Qwen3TTSTokenizer.decode(list of {'audio_codes': pinned host int64 [T,16]}) -> host float32 arrays"},
            "gpu_launches": launches, "clocks": clocks, "roofline": roof,
            "path_frac_of_tensor_roofline": path_frac, "kernels": kernels}
    if world == 1:
        # SURVEY 8f-3 wire formats (additive API, informational): int16 code frames in, PCM16 out, conversion fused into the output conv
        items16 = [{"audio_codes": codes_host[b].to(torch.int16)} for b in range(B_UTT)]
        for _ in range(3):
            pcm, _sr = tok.decode_pcm16(items16)
        torch.cuda.synchronize(dev)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            pcm, _sr = tok.decode_pcm16(items16)
        torch.cuda.synchronize(dev)
        t16 = time.perf_counter() - t0
        assert len(pcm) == B_UTT and pcm[0].shape == (N_SAMPLES,) and str(pcm[0].dtype) == "int16"
        line["e2e_pcm16"] = {"value": B_UTT * T_FRAMES * FRAME_SEC * args.steps / t16, "unit": "audio-s/s", "ms_per_step": t16 / args.steps * 1e3,
                             "h2d_bytes_per_step": B_UTT * T_FRAMES * 16 * 2, "d2h_bytes_per_step": B_UTT * N_SAMPLES * 2,
                             "api": "Qwen3TTSTokenizer.decode_pcm16(list of {'audio_codes': int16 [T,16]}) -> host int16 PCM arrays"}
    if not args.no_streaming and world == 1:
        # BASELINE config 3 rides along as its own section: it is a latency, not the throughput metric of this line
        dec.reserve_streams(args.streams, 4)
        line["streaming"] = [stream_latency(dec, cfg, dev, args.streams, 1, steps=2000, warmup=200),
                             stream_latency(dec, cfg, dev, args.streams, 4, steps=500, warmup=50)]
    if not args.no_encode and world == 1:
        # SURVEY 8(f)-2 (informational on this line; `--workload encode10s` prints the encode path's own full line)
        try:
            r = encode_section(model, dev, steps=5, warmup=3)
            ek, eks = kernel_table(r["acc"], 5, peaks)
            top = next(iter(ek))
            line["encode"] = {"metric": ENC_METRIC, "workload": ENC_NAME, "value": ENC_UTT * ENC_SEC * 5 / (sum(r["step_ms"]) * 1e-3), "unit": "audio-s/s",
                              "ms_per_step": sum(r["step_ms"]) / 5, "e2e": {"value": ENC_UTT * ENC_SEC * 5 / r["e2e_s"], "ms_per_step": r["e2e_s"] / 5 * 1e3},
                              "gpu_launches_per_step": r["launches"] // 5, "top_kernel": {"name": top, **ek[top]}}
        except Exception as ex:  # informational: never fail the bench line over it
            line["encode"] = {"unavailable": f"{type(ex).__name__}: {ex}"[:300]}
    if not args.no_torch_gpu and world == 1:
        del model, dec, tok
        torch.cuda.empty_cache()
        try:
            line["torch_gpu"] = torch_gpu_context(dev, args.streams)
            line["torch_gpu"]["ours_over_eager_bulk"] = value / line["torch_gpu"]["bulk"]["audio_s_per_s"]
            if "stream_chunk_call" in line["torch_gpu"] and "streaming" in line:
                line["torch_gpu"]["eager_chunk_call_over_ours_p50"] = line["torch_gpu"]["stream_chunk_call"]["ms_p50"] / line["streaming"][0]["device_ms"]["p50"]
        except Exception as ex:  # context only: never fail the bench line over it
            line["torch_gpu"] = {"unavailable": f"{type(ex).__name__}: {ex}"[:300]}
    if not args.no_cpu_baseline and world == 1:
        r = cpu_reference_run(steps=3, warmup=1)
        line["cpu_baseline"] = {"value": r["value"], "unit": "audio-s/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"]}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
