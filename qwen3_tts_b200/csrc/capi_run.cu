// extern "C" run-time entry points: rvq_decode / forward / chunked_decode / decode_host / probe / selftest.
#include <cstring>
#include <vector>

#include "engine.h"
#include "kernels.h"

using namespace q3;

namespace {

int ready(Engine& e, const void* codes, int B, int T) {
  if (!e.finalized) return e.fail(Q3TTS_NOT_READY, "weights not finalized");
  if (!codes || B < 1 || T < 1) return e.fail(Q3TTS_BAD_SHAPE, "codes must be [B>=1, Q, T>=1]");
  return 0;
}

// chunk loop of Qwen3TTSTokenizerV2Decoder.chunked_decode (reference modeling :887-894); codes addressed by strides
// item_skip (device, optional): per-item count of leading samples that are never stored (voice-clone prefix, model wrapper :627-630);
// min_skip = the smallest of them: a chunk that ends at or before it is not computed at all.
int chunked(Engine& e, const long long* codes, long long sb, long long sk, long long st, int B, int T, int chunk, int left_ctx, float* wav,
            long long wav_stride, cudaStream_t s, const long long* item_skip = nullptr, long long min_skip = 0) {
  int start = 0;
  long long out_off = 0;
  while (start < T) {
    const int end = (start + chunk < T) ? start + chunk : T;
    const int ctx = (start - left_ctx > 0) ? left_ctx : start;
    const int lo = start - ctx;
    const int Tc = end - lo;
    const long long emitted = (long long)(end - start) * e.total_up - e.lookahead;
    if (emitted > 0) {
      if (!(item_skip && out_off + emitted <= min_skip)) {
        int rc = engine_forward_batched(&e, codes + lo * st, sb, sk, st, B, Tc, wav, wav_stride, ctx * e.total_up, s, out_off, item_skip);
        if (rc) return rc;
      }
      out_off += emitted;
    }
    start = end;
  }
  return Q3TTS_OK;
}

}  // namespace

extern "C" {

int q3tts_rvq_decode(q3tts_handle* h, const int64_t* codes, int B, int T, float* out, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (int rc = ready(e, codes, B, T)) return rc;
  if (!out) return e.fail(Q3TTS_BAD_SHAPE, "rvq_decode: null output");
  DeviceGuard g(e.device);
  const int Q = e.cfg.num_quantizers;
  cudaError_t err = rvq_decode((const long long*)codes, (long long)Q * T, T, 1, e.rvq_tables, nullptr, out, B, Q, T, e.cfg.codebook_size,
                               e.cfg.codebook_dim, e.bf16, (cudaStream_t)stream, 0, nullptr, e.d_bad_code);
  ++e.launches;
  if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("rvq_decode: ") + cudaGetErrorString(err));
  return e.check_codes ? engine_check_codes(&e, (cudaStream_t)stream) : Q3TTS_OK;
}

int q3tts_forward(q3tts_handle* h, const int64_t* codes, int B, int T, float* wav, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (int rc = ready(e, codes, B, T)) return rc;
  if (!wav) return e.fail(Q3TTS_BAD_SHAPE, "forward: null output");
  DeviceGuard g(e.device);
  const int Q = e.cfg.num_quantizers;
  const long long n = (long long)T * e.total_up - e.lookahead;
  if (n <= 0) return e.fail(Q3TTS_BAD_SHAPE, "forward: sequence too short");
  if (int rc = engine_forward_batched(&e, (const long long*)codes, (long long)Q * T, T, 1, B, T, wav, n, 0, (cudaStream_t)stream)) return rc;
  return e.check_codes ? engine_check_codes(&e, (cudaStream_t)stream) : Q3TTS_OK;
}

int q3tts_chunked_decode(q3tts_handle* h, const int64_t* codes, int B, int T, int chunk, int left_ctx, float* wav, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (int rc = ready(e, codes, B, T)) return rc;
  if (!wav || chunk < 1 || left_ctx < 0) return e.fail(Q3TTS_BAD_SHAPE, "chunked_decode: bad arguments");
  DeviceGuard g(e.device);
  const int Q = e.cfg.num_quantizers;
  const long long total = q3tts_chunked_num_samples(h, T, chunk, left_ctx);
  if (int rc = chunked(e, (const long long*)codes, (long long)Q * T, T, 1, B, T, chunk, left_ctx, wav, total, (cudaStream_t)stream)) return rc;
  return e.check_codes ? engine_check_codes(&e, (cudaStream_t)stream) : Q3TTS_OK;
}

// codes16 != 0: codes_host is int16 [B,T,Q]; pcm16 != 0: wav_host is int16 PCM (strides in elements of the output type)
// skip_samples (host, optional): leading samples of each item that are dropped (never stored); lengths[b] is what remains.
static int decode_host_impl(q3tts_handle* h, const void* codes_host, int codes16, int B, int T, void* wav_host, int pcm16, int64_t wav_stride,
                            int64_t* lengths, const int64_t* skip_samples, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (int rc = ready(e, codes_host, B, T)) return rc;
  if (!wav_host || !lengths) return e.fail(Q3TTS_BAD_SHAPE, "decode_host: null output");
  DeviceGuard g(e.device);
  cudaStream_t s = (cudaStream_t)stream;
  const int Q = e.cfg.num_quantizers;
  const int chunk = 300, left_ctx = 25;  // Qwen3TTSTokenizerV2Model.decode uses chunked_decode's defaults (modeling :1012)
  const long long total = q3tts_chunked_num_samples(h, T, chunk, left_ctx);
  const size_t ncodes = (size_t)B * T * Q;
  // One host pass over the codes before anything is launched: the range check the reference gets from F.embedding (an IndexError on
  // CPU, a device assert on CUDA; modeling :675-678) and audio_lengths = (audio_codes[..., 0] > 0).sum(1) * decode_upsample_rate,
  // then a[:l] (modeling :1014-1015).
  const long long ncb = e.cfg.codebook_size;
  bool any_skip = false;
  long long min_skip = -1;
  for (int b = 0; b < B; ++b) {
    long long cnt = 0;
    for (int t = 0; t < T; ++t) {
      const size_t i = ((size_t)b * T + t) * Q;
      for (int k = 0; k < Q; ++k) {
        const long long c = codes16 ? (long long)static_cast<const int16_t*>(codes_host)[i + k] : (long long)static_cast<const int64_t*>(codes_host)[i + k];
        if (c < 0 || c >= ncb)
          return e.fail(Q3TTS_CODE_OUT_OF_RANGE, "decode_host: code " + std::to_string(c) + " at [" + std::to_string(b) + "," + std::to_string(t) + "," +
                                                      std::to_string(k) + "] is outside [0, " + std::to_string(ncb) + ")");
        if (k == 0) cnt += c > 0 ? 1 : 0;
      }
    }
    long long l = cnt * e.total_up;
    if (l > total) l = total;
    if (l < 0) l = 0;
    long long cut = skip_samples ? skip_samples[b] : 0;
    if (cut < 0) return e.fail(Q3TTS_BAD_SHAPE, "decode_host: negative skip_samples");
    if (cut > l) cut = l;
    any_skip = any_skip || cut > 0;
    if (min_skip < 0 || cut < min_skip) min_skip = cut;
    lengths[b] = l - cut;
  }
  if (min_skip < 0) min_skip = 0;
  if (any_skip) {
    if ((size_t)B > e.item_skip_elems) {
      cudaStreamSynchronize(s);
      if (e.item_skip_stage) cudaFree(e.item_skip_stage);
      e.item_skip_stage = nullptr; e.item_skip_elems = 0;
      if (cudaMalloc((void**)&e.item_skip_stage, (size_t)B * sizeof(long long)) != cudaSuccess) { cudaGetLastError(); return e.fail(Q3TTS_OOM, "decode_host: skip staging"); }
      e.item_skip_elems = (size_t)B;
    }
    // the caller's array may be pageable and is only borrowed for this call: clamp into a small vector and copy synchronously
    std::vector<long long> cuts((size_t)B);
    for (int b = 0; b < B; ++b) { const long long c = skip_samples[b]; cuts[b] = c < 0 ? 0 : c; }
    if (cudaMemcpyAsync(e.item_skip_stage, cuts.data(), (size_t)B * sizeof(long long), cudaMemcpyHostToDevice, s) != cudaSuccess ||
        cudaStreamSynchronize(s) != cudaSuccess)
      return e.fail(Q3TTS_CUDA_ERROR, "decode_host: skip upload failed");
  }
  if (ncodes > e.codes_stage_elems) {
    cudaStreamSynchronize(s);
    if (e.codes_stage) cudaFree(e.codes_stage);
    e.codes_stage = nullptr; e.codes_stage_elems = 0;
    if (cudaMalloc((void**)&e.codes_stage, ncodes * sizeof(long long)) != cudaSuccess) { cudaGetLastError(); return e.fail(Q3TTS_OOM, "decode_host: codes staging"); }
    e.codes_stage_elems = ncodes;
  }
  if (codes16 && ncodes > e.codes16_stage_elems) {
    cudaStreamSynchronize(s);
    if (e.codes16_stage) cudaFree(e.codes16_stage);
    e.codes16_stage = nullptr; e.codes16_stage_elems = 0;
    if (cudaMalloc((void**)&e.codes16_stage, ncodes * sizeof(short)) != cudaSuccess) { cudaGetLastError(); return e.fail(Q3TTS_OOM, "decode_host: int16 codes staging"); }
    e.codes16_stage_elems = ncodes;
  }
  const size_t nwav = (size_t)B * (size_t)(total > 0 ? total : 0);
  // Pinned (page-locked) result buffer with room for whole rows: the output conv stores straight into it over PCIe (unified addressing
  // gives the device an alias of the host pointer), so the waveform never exists in HBM and there is no separate device-to-host copy.
  void* direct = nullptr;
  {
    static const bool off = Q3_ENV("Q3TTS_NO_DIRECT_D2H") != nullptr;
    cudaPointerAttributes at;
    if (!off && wav_stride >= total && cudaPointerGetAttributes(&at, wav_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer)
      direct = at.devicePointer;
    else
      cudaGetLastError();
  }
  if (!direct && nwav > e.wav_stage_elems) {
    cudaStreamSynchronize(s);
    if (e.wav_stage) cudaFree(e.wav_stage);
    e.wav_stage = nullptr; e.wav_stage_elems = 0;
    if (cudaMalloc((void**)&e.wav_stage, nwav * sizeof(float)) != cudaSuccess) { cudaGetLastError(); return e.fail(Q3TTS_OOM, "decode_host: wav staging"); }
    e.wav_stage_elems = nwav;
  }
  cudaError_t err;
  if (codes16) {
    err = cudaMemcpyAsync(e.codes16_stage, codes_host, ncodes * sizeof(short), cudaMemcpyHostToDevice, s);
    if (err == cudaSuccess) err = widen_codes_i16(e.codes16_stage, e.codes_stage, (long long)ncodes, s);
  } else {
    err = cudaMemcpyAsync(e.codes_stage, codes_host, ncodes * sizeof(long long), cudaMemcpyHostToDevice, s);
  }
  if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("decode_host: H2D: ") + cudaGetErrorString(err));
  if (total > 0) {
    // codes are [B,T,Q] here (wrapper layout): strides b = T*Q, k = 1, t = Q
    e.out_pcm16 = pcm16 != 0;  // the staging buffer is then int16 [B][total]
    float* dst = direct ? (float*)direct : e.wav_stage;
    const long long dst_stride = direct ? wav_stride : total;
    int rc = chunked(e, e.codes_stage, (long long)T * Q, 1, Q, B, T, chunk, left_ctx, dst, dst_stride, s, any_skip ? e.item_skip_stage : nullptr, min_skip);
    e.out_pcm16 = false;
    if (rc) return rc;
    // one strided copy for the whole batch: rows are cut to their length by the caller's views, so the columns between a
    // row's length and the longest row may be written too (the host buffer has wav_stride >= that many columns per row)
    long long maxlen = 0;
    for (int b = 0; b < B; ++b) maxlen = lengths[b] > maxlen ? lengths[b] : maxlen;
    if (maxlen > wav_stride) return e.fail(Q3TTS_BAD_SHAPE, "decode_host: wav_stride smaller than an output row");
    if (maxlen > 0 && !direct) {
      const size_t es = pcm16 ? sizeof(short) : sizeof(float);
      err = cudaMemcpy2DAsync(wav_host, (size_t)wav_stride * es, e.wav_stage, (size_t)total * es, (size_t)maxlen * es, (size_t)B, cudaMemcpyDeviceToHost, s);
      if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("decode_host: D2H: ") + cudaGetErrorString(err));
    }
  }
  err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("decode_host: ") + cudaGetErrorString(err));
  return Q3TTS_OK;
}

int q3tts_decode_host(q3tts_handle* h, const int64_t* codes_host, int B, int T, float* wav_host, int64_t wav_stride, int64_t* lengths,
                      const int64_t* skip_samples, void* stream) {
  return decode_host_impl(h, codes_host, 0, B, T, wav_host, 0, wav_stride, lengths, skip_samples, stream);
}

int q3tts_decode_host_pcm16(q3tts_handle* h, const int16_t* codes_host, int B, int T, int16_t* pcm_host, int64_t pcm_stride, int64_t* lengths,
                            const int64_t* skip_samples, void* stream) {
  return decode_host_impl(h, codes_host, 1, B, T, pcm_host, 1, pcm_stride, lengths, skip_samples, stream);
}

int q3tts_set_check_codes(q3tts_handle* h, int on) {
  if (!h) return Q3TTS_BAD_CONFIG;
  h->e.check_codes = on != 0;
  return Q3TTS_OK;
}

int q3tts_forward_probe(q3tts_handle* h, const int64_t* codes, int B, int T, const char* stage, float* out, int64_t capacity, int* C, int* L,
                        void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (int rc = ready(e, codes, B, T)) return rc;
  if (!stage || !out) return e.fail(Q3TTS_BAD_SHAPE, "forward_probe: null argument");
  DeviceGuard g(e.device);
  Plan* plan = nullptr;
  if (int rc = engine_get_plan(&e, B, T, &plan)) return rc;
  const int Q = e.cfg.num_quantizers;
  return engine_run(&e, plan, (const long long*)codes, (long long)Q * T, T, 1, nullptr, 0, 0, stage, out, capacity, C, L, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------ streaming
int q3tts_stream_reserve(q3tts_handle* h, int max_streams, int max_frames_per_step) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (!e.finalized) return e.fail(Q3TTS_NOT_READY, "stream_reserve: weights not finalized");
  if (max_streams < 1 || max_frames_per_step < 1) return e.fail(Q3TTS_BAD_SHAPE, "stream_reserve: sizes must be positive");
  DeviceGuard g(e.device);
  return engine_stream_reserve(&e, max_streams, max_frames_per_step);
}

static int stream_clear_slot(Engine& e, int id, void* stream) {
  cudaStream_t s = (cudaStream_t)stream;
  if (cudaMemsetAsync(e.stream_state + (size_t)id * e.state_bytes, 0, (size_t)e.state_bytes, s) != cudaSuccess ||
      cudaMemsetAsync(e.d_slot_pos + id, 0, sizeof(int), s) != cudaSuccess)
    return e.fail(Q3TTS_CUDA_ERROR, "stream: state reset failed");
  e.slot_frames[id] = 0;
  return Q3TTS_OK;
}

int q3tts_stream_open(q3tts_handle* h, int* stream_id, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (!stream_id) return e.fail(Q3TTS_BAD_SHAPE, "stream_open: null output");
  if (!e.max_slots) return e.fail(Q3TTS_BAD_STREAM, "stream_open: call q3tts_stream_reserve first");
  DeviceGuard g(e.device);
  for (int i = 0; i < e.max_slots; ++i)
    if (e.slot_frames[i] < 0) {
      *stream_id = i;
      return stream_clear_slot(e, i, stream);
    }
  return e.fail(Q3TTS_BAD_STREAM, "stream_open: all reserved stream slots are in use");
}

int q3tts_stream_reset(q3tts_handle* h, int stream_id, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (stream_id < 0 || stream_id >= e.max_slots || e.slot_frames[stream_id] < 0) return e.fail(Q3TTS_BAD_STREAM, "stream_reset: unknown stream id");
  DeviceGuard g(e.device);
  return stream_clear_slot(e, stream_id, stream);
}

int q3tts_stream_close(q3tts_handle* h, int stream_id) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (stream_id < 0 || stream_id >= e.max_slots || e.slot_frames[stream_id] < 0) return e.fail(Q3TTS_BAD_STREAM, "stream_close: unknown stream id");
  e.slot_frames[stream_id] = -1;
  return Q3TTS_OK;
}

int64_t q3tts_stream_frames(q3tts_handle* h, int stream_id) {
  if (!h || stream_id < 0 || stream_id >= h->e.max_slots) return -1;
  return h->e.slot_frames[stream_id];
}

int64_t q3tts_stream_state_bytes(q3tts_handle* h) {
  if (!h) return 0;
  return h->e.state_bytes + (int64_t)2 * h->e.cfg.num_hidden_layers * h->e.kv_ring * h->e.cfg.hidden_size * 2;
}

int q3tts_stream_step(q3tts_handle* h, const int32_t* stream_ids, int S, const int64_t* codes, int P, float* wav, int64_t wav_stride, int32_t* n_valid,
                      void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (!e.finalized) return e.fail(Q3TTS_NOT_READY, "weights not finalized");
  if (!stream_ids || !codes || !wav) return e.fail(Q3TTS_BAD_SHAPE, "stream_step: null argument");
  if (wav_stride < (int64_t)P * e.total_up) return e.fail(Q3TTS_BAD_SHAPE, "stream_step: wav_stride smaller than P * upsample");
  DeviceGuard g(e.device);
  const int Q = e.cfg.num_quantizers;
  if (int rc = engine_stream_step(&e, stream_ids, S, (const long long*)codes, (long long)Q * P, P, 1, P, wav, wav_stride, n_valid, (cudaStream_t)stream))
    return rc;
  return e.check_codes ? engine_check_codes(&e, (cudaStream_t)stream) : Q3TTS_OK;  // opt-in: the check synchronises the stream
}

}  // extern "C"
