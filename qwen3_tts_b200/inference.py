"""`Qwen3TTSModel` — the decode hand-off of qwen_tts/inference/qwen3_tts_model.py ("model wrapper").

The reference wrapper does three things: text templating / validation, `self.model.generate(...)` (talker LM +
code predictor, talker :2022-2292) and `self.model.speech_tokenizer.decode([...])` with the voice-clone prefix
handling (model wrapper :612-631, :727, :838).  Only the last one is on the accelerated hot path.  This class keeps
the three public `generate_*` names and their `(wavs: List[np.ndarray], sample_rate)` return, takes the talker as an
injected object (the unmodified reference `Qwen3TTSForConditionalGeneration`, or anything with the same
`generate(...) -> (List[LongTensor (T_i,16)], _)`), and owns the hand-off to the B200 tokenizer.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch


class Qwen3TTSModel:
    def __init__(self, model, processor=None, generate_defaults: Optional[Dict[str, Any]] = None, speech_tokenizer=None,
                 tokenize: Optional[Callable[[List[str]], List[torch.Tensor]]] = None):
        """`model`: talker with `.generate(**kw) -> (codes_list, _)`; its `.speech_tokenizer` is replaced by
        `speech_tokenizer` (a qwen3_tts_b200.Qwen3TTSTokenizer) when one is given.  `tokenize`: texts -> input ids
        (the reference builds them with its processor, model wrapper :278-285); optional for pre-tokenised calls."""
        self.model = model
        self.processor = processor
        self.generate_defaults = generate_defaults or {}
        if speech_tokenizer is not None:
            self.model.speech_tokenizer = speech_tokenizer
        self._tokenize = tokenize
        self.device = getattr(model, "device", None)

    @classmethod
    def wrap(cls, reference_wrapper, speech_tokenizer) -> "Qwen3TTSModel":
        """Take a loaded reference `qwen_tts.Qwen3TTSModel` and route its decode through the B200 tokenizer."""
        tok = getattr(reference_wrapper, "_tokenize_texts", None)
        inst = cls(reference_wrapper.model, getattr(reference_wrapper, "processor", None),
                   getattr(reference_wrapper, "generate_defaults", None), speech_tokenizer, tok)
        inst._ref = reference_wrapper
        return inst

    # ------------------------------------------------------------------ the hot-path hand-off
    def decode_talker_codes(self, talker_codes_list: Sequence[torch.Tensor],
                            ref_code_list: Optional[Sequence[Optional[torch.Tensor]]] = None) -> Tuple[List[np.ndarray], int]:
        """The decode hand-off of the three generate_* calls.  Without reference codes it is one
        `speech_tokenizer.decode([{"audio_codes": c}, ...])` (model wrapper :727 / :838).  With them (voice clone, model wrapper
        :612-631) every item is decoded as reference-prefix + generated codes, and the prefix audio - the first
        int(ref_len / total_len * n_samples) samples - is cut INSIDE the decode call (`skip_samples`): the output conv never stores
        it, chunks that lie inside every item's prefix are not computed, nothing is sliced on the host."""
        tok = self.model.speech_tokenizer
        n = len(talker_codes_list)
        refs = list(ref_code_list) if ref_code_list is not None else [None] * n
        items = [c if r is None else torch.cat([r.to(c.device), c], dim=0) for c, r in zip(talker_codes_list, refs)]
        payload = [{"audio_codes": c} for c in items]
        if all(r is None for r in refs):
            return tok.decode(payload)
        # n_samples of an item = what decode() returns for it: min(valid frames * upsample, chunked length of the padded batch)
        padded_T = max(int(c.shape[0]) for c in items)
        whole = tok.num_decoded_samples(padded_T)
        up = tok.get_decode_upsample_rate()
        skips = []
        for c, r in zip(items, refs):
            if r is None:
                skips.append(0)
                continue
            n_samples = min(int((c[:, 0] > 0).sum()) * up, whole)
            skips.append(int(int(r.shape[0]) / max(int(c.shape[0]), 1) * n_samples))   # model wrapper :627-629
        return tok.decode(payload, skip_samples=skips)

    # ------------------------------------------------------------------ generate_* (talker side injected)
    def _ids(self, texts: Union[str, List[str]]):
        texts = texts if isinstance(texts, list) else [texts]
        if self._tokenize is None:
            raise RuntimeError("no text front-end: construct with tokenize=... or Qwen3TTSModel.wrap(reference_wrapper, tokenizer)")
        return self._tokenize(texts)

    def _merged(self, kwargs):
        out = dict(self.generate_defaults)
        out.update({k: v for k, v in kwargs.items() if v is not None})
        return out

    @torch.no_grad()
    def generate_custom_voice(self, text, speaker=None, language=None, instruct=None, non_streaming_mode: bool = True, **kwargs):
        """model wrapper :732-839."""
        if getattr(self, "_ref", None) is not None:
            return self._ref.generate_custom_voice(text=text, speaker=speaker, language=language, instruct=instruct,
                                                   non_streaming_mode=non_streaming_mode, **kwargs)
        codes, _ = self.model.generate(input_ids=self._ids(text), speakers=speaker, languages=language, instruct_ids=instruct,
                                       non_streaming_mode=non_streaming_mode, **self._merged(kwargs))
        return self.decode_talker_codes(codes)

    @torch.no_grad()
    def generate_voice_design(self, text, instruct, language=None, non_streaming_mode: bool = True, **kwargs):
        """model wrapper :637-728."""
        if getattr(self, "_ref", None) is not None:
            return self._ref.generate_voice_design(text=text, instruct=instruct, language=language, non_streaming_mode=non_streaming_mode, **kwargs)
        codes, _ = self.model.generate(input_ids=self._ids(text), instruct_ids=instruct, languages=language,
                                       non_streaming_mode=non_streaming_mode, **self._merged(kwargs))
        return self.decode_talker_codes(codes)

    @torch.no_grad()
    def generate_voice_clone(self, text, language=None, voice_clone_prompt: Optional[Dict[str, Any]] = None,
                             non_streaming_mode: bool = False, **kwargs):
        """model wrapper :470-633.  `voice_clone_prompt["ref_code"]` (list of LongTensor (T_ref,16) or None) drives the prefix cut."""
        if getattr(self, "_ref", None) is not None:
            return self._ref.generate_voice_clone(text=text, language=language, voice_clone_prompt=voice_clone_prompt,
                                                  non_streaming_mode=non_streaming_mode, **kwargs)
        prompt = voice_clone_prompt or {}
        codes, _ = self.model.generate(input_ids=self._ids(text), voice_clone_prompt=prompt, languages=language,
                                       non_streaming_mode=non_streaming_mode, **self._merged(kwargs))
        return self.decode_talker_codes(codes, prompt.get("ref_code", None))


class TalkerStreamSink:
    """On-device hand-off from the talker's generation loop to the stateful streaming decoder.

    The reference collects every step's `codec_ids` [B, 16] (talker :1681) into Python lists, stacks them after generation
    (talker :2280) and only then decodes (model wrapper :838).  This sink takes the same per-step tensor while it is still on
    the device, buffers `packet` frames per stream and runs one `decode_step` for all live streams when a packet is full, so
    audio leaves `packet * 80 ms` after its codes exist.  The concatenated chunks of a stream equal
    `Decoder.forward(all its codes)` on `[0, T*1920 - 555)` (the streaming contract), for any utterance length:
    `finish(row)` / `close()` decode the frames still buffered as one short packet.  `decode_step` does not synchronise with the
    host (the stream ids travel through a ring of pinned staging blocks), so the sink can be driven from the talker's stream.
    """

    def __init__(self, decoder, batch: int, packet: int = 1, reserve: bool = True):
        """`decoder`: a qwen3_tts_b200 Qwen3TTSTokenizerV2Decoder.  `reserve=False` when the caller has already sized the stream
        pool with `decoder.reserve_streams(...)` (reserving again would drop the state of streams that are open)."""
        if batch < 1 or packet < 1:
            raise ValueError("batch and packet must be positive")
        self.decoder = decoder
        self.packet = int(packet)
        self.Q = int(decoder.config.num_quantizers)
        if reserve:
            decoder.reserve_streams(batch, self.packet)
        self.ids = [decoder.open_stream() for _ in range(batch)]
        self.live = [True] * batch
        self._buf = torch.zeros((batch, self.Q, self.packet), dtype=torch.int64, device=decoder.device)
        self._fill = 0   # frames buffered since the last full packet (the same for every live row: one frame per push)

    def push(self, codec_ids: torch.Tensor):
        """codec_ids: [B, 16] integer tensor on the decoder's device (one generated frame per stream, finished streams included:
        their rows are ignored once `finish(i)` was called).  Returns None until a packet is full, then
        `(rows, wav [len(rows), packet*1920] float32 on the device, n_valid)` for the live streams `rows`."""
        if codec_ids.dim() != 2 or codec_ids.shape != (len(self.ids), self.Q):
            raise ValueError(f"Expected codec_ids of shape ({len(self.ids)}, {self.Q}), got {tuple(codec_ids.shape)}")
        self._buf[:, :, self._fill] = codec_ids.to(device=self._buf.device, dtype=torch.int64)
        self._fill += 1
        if self._fill < self.packet:
            return None
        self._fill = 0
        rows = [i for i, ok in enumerate(self.live) if ok]
        if not rows:
            return None
        codes = self._buf if len(rows) == len(self.ids) else self._buf[rows]
        wav, n_valid = self.decoder.decode_step([self.ids[i] for i in rows], codes.contiguous())
        return rows, wav, n_valid

    def _flush(self, rows: List[int]):
        """Decode the frames rows `rows` have buffered since the last full packet (a short packet of `_fill` frames)."""
        if not rows or self._fill == 0:
            return None
        codes = self._buf[rows][:, :, : self._fill].contiguous()
        wav, n_valid = self.decoder.decode_step([self.ids[i] for i in rows], codes)
        return wav, n_valid

    def finish(self, row: int) -> Optional[torch.Tensor]:
        """Stream `row` has produced its stop token.  The frames it still has buffered (the utterance length need not be a multiple
        of `packet`) are decoded as one short packet and returned as a 1-D device tensor (None when nothing was pending); then the
        stream's state is released and later frames of that row are ignored."""
        if not self.live[row]:
            return None
        out = self._flush([row])
        self.live[row] = False
        self.decoder.close_stream(self.ids[row])
        if out is None:
            return None
        wav, n_valid = out
        return wav[0, : n_valid[0]]

    def close(self) -> Dict[int, torch.Tensor]:
        """Flush and release every stream that is still live: {row: last samples} for the rows that had frames pending."""
        rows = [i for i, ok in enumerate(self.live) if ok]
        out = self._flush(rows)
        tails: Dict[int, torch.Tensor] = {}
        if out is not None:
            wav, n_valid = out
            tails = {r: wav[k, : n_valid[k]] for k, r in enumerate(rows)}
        for i in rows:
            self.live[i] = False
            self.decoder.close_stream(self.ids[i])
        self._fill = 0
        return tails
