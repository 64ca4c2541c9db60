/* libq3tts_decode.so — C-ABI of the B200-native Qwen3-TTS-Tokenizer-12Hz decode path.
 *
 * The reference (kritsanan1/Qwen3-TTS) has no FFI for this path: its boundary is a set of Python
 * classes.  Each entry point below replaces the body of one reference method; the Python shim in
 * qwen3_tts_b200/ keeps the reference's class names and signatures and calls these through ctypes
 * (INTEGRATION.md shows the binding).  "modeling" =
 * qwen_tts/core/tokenizer_12hz/modeling_qwen3_tts_tokenizer_v2.py, "wrapper" =
 * qwen_tts/inference/qwen3_tts_tokenizer.py.
 *
 * Conventions: every function returns 0 on success or a negative q3tts_status; the message of the
 * last failure on a handle is q3tts_last_error(h).  Plain pointers and sizes only.  A handle is
 * bound to one CUDA device and is NOT thread-safe (one handle per device per worker).  Device
 * pointers are borrowed for the duration of the call; every launch goes to the caller's stream
 * (cudaStream_t passed as void*; NULL = default stream).  No allocation happens after
 * q3tts_reserve() unless a call needs more frames than were reserved.
 */
#ifndef Q3TTS_DECODE_H_
#define Q3TTS_DECODE_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct q3tts_handle q3tts_handle;

typedef enum {
  Q3TTS_OK = 0,
  Q3TTS_BAD_SHAPE = -1,         /* Python: ValueError (modeling :869-870) */
  Q3TTS_BAD_CONFIG = -2,        /* Python: ValueError */
  Q3TTS_CODE_OUT_OF_RANGE = -3, /* Python: IndexError (F.embedding would assert) */
  Q3TTS_CUDA_ERROR = -4,        /* Python: RuntimeError */
  Q3TTS_OOM = -5,               /* Python: MemoryError */
  Q3TTS_BAD_STREAM = -6,        /* Python: KeyError */
  Q3TTS_BAD_KEY = -7,           /* Python: KeyError (unknown / missing state_dict key) */
  Q3TTS_NOT_READY = -8          /* Python: RuntimeError (weights not finalized) */
} q3tts_status;

/* Fields the reference decoder reads from Qwen3TTSTokenizerV2DecoderConfig
 * (configuration_qwen3_tts_tokenizer_v2.py:72-114) plus codebook_dim / head_dim, which reach the
 * reference through **kwargs (modeling :288, :830). */
typedef struct {
  int32_t codebook_size;      /* 2048 */
  int32_t codebook_dim;       /* 512 (RVQ rows are codebook_dim/2 wide) */
  int32_t hidden_size;        /* 1024 */
  int32_t latent_dim;         /* 1024 */
  int32_t num_attention_heads; /* 16 */
  int32_t num_key_value_heads; /* 16 */
  int32_t head_dim;           /* 64 */
  int32_t sliding_window;     /* 72 */
  int32_t intermediate_size;  /* 3072 */
  int32_t num_hidden_layers;  /* 8 */
  int32_t num_quantizers;     /* 16 */
  int32_t decoder_dim;        /* 1536 */
  int32_t num_upsample_rates;
  int32_t upsample_rates[8];  /* 8,5,4,3 */
  int32_t num_upsampling_ratios;
  int32_t upsampling_ratios[4]; /* 2,2 */
  int32_t max_position_embeddings; /* 8000 */
  float rms_norm_eps;         /* 1e-5 */
  float rope_theta;           /* 10000 */
} q3tts_config;

/* compute_mode of q3tts_finalize_weights.  Bit 0 = tensor-core operand type (0 fp16, 1 bf16; accumulation is fp32 in both).  The
 * upper bits pick how many operands are carried as hi + lo pairs of 16-bit values (fp32-class products from 2-3 MMAs; DESIGN.md
 * "Precision"): default = the main signal path of the early stages (RVQ -> pre_conv -> input_proj, final norm -> output_proj ->
 * upsample convs -> decoder[0]; ~3 % of the FLOPs, ~40 % of the rounding noise), _PLAIN = none (every operand one 16-bit value),
 * _SPLIT_ALL = every GEMM in front of the vocoder blocks. */
enum {
  Q3TTS_MODE_FP16 = 0, Q3TTS_MODE_BF16 = 1,
  Q3TTS_MODE_FP16_PLAIN = 2, Q3TTS_MODE_BF16_PLAIN = 3,
  Q3TTS_MODE_FP16_SPLIT_ALL = 4, Q3TTS_MODE_BF16_SPLIT_ALL = 5
};
enum { Q3TTS_ENGINE_TC = 0, Q3TTS_ENGINE_SIMT = 1 }; /* SIMT = debug cross-check kernels, not a product path */

/* Lifetime.  Replaces Qwen3TTSTokenizerV2Decoder.__init__ (modeling :824-866). */
int q3tts_create(const q3tts_config* cfg, int device, q3tts_handle** out);
void q3tts_destroy(q3tts_handle* h);
const char* q3tts_last_error(q3tts_handle* h);

/* Weights.  Replaces nn.Module.load_state_dict for the decoder: `key` is a reference
 * Qwen3TTSTokenizerV2Decoder.state_dict() key (SURVEY §8(b) lists them), `data` is fp32, host
 * (on_device = 0) or device (on_device = 1) memory, copied before the call returns. */
int q3tts_load_tensor(q3tts_handle* h, const char* key, const float* data, int ndim, const int64_t* shape, int on_device);
/* Repack everything into tensor-core operand layouts (16-bit K-major, folded RVQ tables, exp'd
 * Snake parameters).  Fails with Q3TTS_BAD_KEY if a tensor is missing or mis-shaped. */
int q3tts_finalize_weights(q3tts_handle* h, int compute_mode);
int q3tts_reserve(q3tts_handle* h, int max_frames_per_launch);
int q3tts_set_engine(q3tts_handle* h, int engine);

/* Sizes (pure host arithmetic). */
int64_t q3tts_forward_num_samples(q3tts_handle* h, int T);                          /* T*1920 - 555 */
int64_t q3tts_chunked_num_samples(q3tts_handle* h, int T, int chunk, int left_ctx); /* T*1920 - 555*nchunks */

/* SplitResidualVectorQuantizer.decode (modeling :814-820): codes [B,Q,T] int64 (device) ->
 * out [B,codebook_dim,T] fp32 (device). */
int q3tts_rvq_decode(q3tts_handle* h, const int64_t* codes, int B, int T, float* out, void* stream);

/* Qwen3TTSTokenizerV2Decoder.forward (modeling :868-883): codes [B,Q,T] int64 (device) ->
 * wav [B, T*1920-555] fp32 (device), clamped to [-1,1]. */
int q3tts_forward(q3tts_handle* h, const int64_t* codes, int B, int T, float* wav, void* stream);

/* Qwen3TTSTokenizerV2Decoder.chunked_decode (modeling :885-895): codes [B,Q,T] (device) ->
 * wav [B, q3tts_chunked_num_samples] fp32 (device). */
int q3tts_chunked_decode(q3tts_handle* h, const int64_t* codes, int B, int T, int chunk, int left_ctx, float* wav, void* stream);

/* Qwen3TTSTokenizerV2Model.decode + the wrapper's H2D/D2H (modeling :1012-1015, wrapper :329, :364):
 * codes [B,T,Q] int64 in HOST memory (padded with 0 like pad_sequence) -> wav_host [B, wav_stride]
 * fp32 HOST memory, lengths[b] = min((codes[b,:,0] > 0).sum()*upsample, chunked length).  The first
 * lengths[b] samples of row b are its waveform; columns up to the longest row's length may be written
 * too (one strided copy for the batch).  Synchronises the stream before returning.
 * Every code is range-checked on the host first: Q3TTS_CODE_OUT_OF_RANGE (Python IndexError) where the reference's F.embedding
 * raises (modeling :675-678), before anything is launched.
 * skip_samples (host, [B], may be NULL): the voice-clone prefix cut of Qwen3TTSModel.generate_voice_clone (model wrapper :612-631)
 * folded into the decode.  The first skip_samples[b] samples of item b are never stored (the output conv skips them, chunks that
 * lie wholly inside every item's cut are not computed); row b starts at its sample skip_samples[b] and lengths[b] is what is left. */
int q3tts_decode_host(q3tts_handle* h, const int64_t* codes_host, int B, int T, float* wav_host, int64_t wav_stride, int64_t* lengths,
                      const int64_t* skip_samples, void* stream);

/* Same call with the wire formats of SURVEY 8f-3: int16 code frames in (a 2048-entry codebook index fits; the reference
 * keeps codes as JSON lists / numpy payloads, finetuning/prepare_data.py:49, examples/test_tokenizer_12hz.py:54-58) and PCM16
 * out, pcm = lrintf(clamp(wav, -1, 1) * 32767): what `sf.write(path, wav, sr)` stores for a float waveform (examples :33).
 * The conversion is fused into the output conv, so host traffic is 4x / 2x smaller.  Strides in int16 elements. */
int q3tts_decode_host_pcm16(q3tts_handle* h, const int16_t* codes_host, int B, int T, int16_t* pcm_host, int64_t pcm_stride, int64_t* lengths,
                            const int64_t* skip_samples, void* stream);

/* Opt-in range check for the entry points that take DEVICE codes (rvq_decode, forward, chunked_decode, stream_step).  The gather
 * kernel never faults on a bad index (it decodes a clamped one and raises a device flag); with the check on, these calls read the
 * flag after their launches - which synchronises the stream - and return Q3TTS_CODE_OUT_OF_RANGE, as the reference's F.embedding
 * asserts (modeling :675-678).  Off by default: the hot path stays asynchronous ("check on request only"). */
int q3tts_set_check_codes(q3tts_handle* h, int on);

/* Encode side, first stage built: the residual-vector-quantiser search at the end of
 * Qwen3TTSTokenizerV2Model.encode (modeling :960-990 -> transformers MimiSplitResidualVectorQuantizer.encode): latents
 * [B,codebook_dim,T] fp32 (device; the Mimi encoder's output) -> codes [B,Q,T] int64 (device), Q = num_quantizers (the reference
 * keeps the first encoder_valid_num_quantizers = 16, modeling :982).  Needs the reference model's `encoder.quantizer.*` tensors
 * loaded before q3tts_finalize_weights (semantic / acoustic input_proj.weight, layers.N.codebook.{embed_sum, cluster_usage});
 * Q3TTS_NOT_READY otherwise. */
int q3tts_rvq_encode(q3tts_handle* h, const float* latents, int B, int T, int64_t* codes, void* stream);

/* The whole encode path.  Replaces Qwen3TTSTokenizerV2Model.encode (modeling :960-990), i.e. MimiModel.encode of the
 * `transformers` package as the reference instantiates it (Qwen3TTSTokenizerV2Encoder, modeling :898-907): SEANet conv stack
 * (k7 conv 1 -> 64; four [ResnetBlock, ELU, stride-r conv], r = 4, 5, 6, 8; ELU, k3 conv -> 512), 8-layer causal sliding-window
 * transformer (LayerNorm, RoPE, GELU MLP, layer scale), replicate-padded stride-2 conv to 12.5 Hz, RVQ search.  Needs the reference
 * model's `encoder.*` tensors (encoder.encoder.layers.*, encoder.encoder_transformer.layers.*, encoder.downsample.conv.weight,
 * encoder.quantizer.*) loaded before q3tts_finalize_weights; Q3TTS_NOT_READY otherwise.  Channel counts, strides and the layer count
 * are read from the tensor shapes; what shapes cannot tell comes from the q3tts_encoder_config struct, whose defaults are MimiConfig()'s.
 * Arithmetic: every GEMM runs on the tensor cores with BOTH operands carried as fp16 hi + lo pairs (3 MMAs per product, fp32
 * accumulate: ~22 bits per operand) whatever the handle's compute mode; activations stay fp32 in HBM.  The output is integer: the
 * bar is identical indices wherever the float64 oracle's two best codes are not tied to fp32 rounding (tests/test_gpu_encode.py).
 * Memory: the encode workspace (~1 KB per sample of a launch; batches above 16 M samples run as several launches) is allocated by the
 * first call that needs it and kept; q3tts_reserve() does not cover it.  Range: operands are fp16 pairs, so |activation| must stay
 * below 65504 inside the encoder (true by a wide margin for audio in [-1, 1] through LayerNorm'd / ELU'd stages; not checked). */
typedef struct {
  int32_t num_attention_heads;     /* 8 (head_dim must be 64) */
  int32_t sliding_window;          /* 250 */
  int32_t max_position_embeddings; /* 8000 frames at 25 Hz per item */
  float rope_theta;                /* 10000 */
  float norm_eps;                  /* 1e-5 */
} q3tts_encoder_config;
int q3tts_set_encoder_config(q3tts_handle* h, const q3tts_encoder_config* cfg);   /* before q3tts_finalize_weights */
/* frames MimiModel.encode emits for `samples` input samples (every strided conv rounds up): ceil-chain over 4, 5, 6, 8, 2 */
int64_t q3tts_encode_num_frames(q3tts_handle* h, int64_t samples);
/* wav [B, wav_stride] fp32 (device; row b = item b's T samples, zero-padded like the reference's feature extractor pads) ->
 * codes [B, Q, F] int64 (device), F = q3tts_encode_num_frames(T).  latents (optional, device, [B, codebook_dim, F] fp32): the
 * quantiser's input, for parity checks.  The caller cuts item b to ceil(valid_samples[b] / 1920) frames (modeling :983). */
int q3tts_encode(q3tts_handle* h, const float* wav, int B, int64_t T, int64_t wav_stride, int64_t* codes, float* latents, void* stream);
/* Host form: wav_host [B, wav_stride] fp32 -> codes_host [B, F, Q] int64 (the layout q3tts_decode_host reads); valid_samples (host,
 * [B], may be NULL = T) -> lengths[b] = ceil(valid_samples[b] / 1920) frames of row b are item b's codes.  Synchronises.
 * A page-locked wav_host is read in place by the first conv over PCIe (no separate host-to-device copy); a pageable one is staged. */
int q3tts_encode_host(q3tts_handle* h, const float* wav_host, int B, int64_t T, int64_t wav_stride, const int64_t* valid_samples,
                      int64_t* codes_host, int64_t* lengths, void* stream);

/* The same gather-sum on int16 code frames (the wire format of q3tts_decode_host_pcm16) with NO allocation anywhere in the call.
 * No counterpart in the reference; added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid it).  codes [B,Q,T] int16 (device) -> out [B,codebook_dim,T]
 * fp32 (device), equal to q3tts_rvq_decode on the widened codes.  The int64 staging the gather kernel reads lives in a file-scope
 * __device__ array of Q3TTS_STATIC_MAX_CODES entries (4 MiB, held by every GPU that loads the library until the process exits):
 * B*Q*T beyond that returns Q3TTS_BAD_SHAPE.  All calls on a device share the array and it keeps the last call's contents (each
 * call overwrites exactly what it reads); calls that overlap on the GPU - two streams, two handles on one device - race on it. */
#define Q3TTS_STATIC_MAX_CODES (512 * 1024)
int64_t q3tts_static_max_codes(void);
int q3tts_rvq_decode_i16_static(q3tts_handle* h, const int16_t* codes, int B, int T, float* out, void* stream);

/* ---- Streaming: stateful per-packet decode.  The reference has NO counterpart: its only streaming-shaped call is
 * chunked_decode's `forward(codes[..., s-25:e])[..., 25*1920:]` (modeling :889-893), which re-decodes 25 frames of left
 * context per chunk and is only approximately equal to the whole-sequence decode.  These entry points keep, per stream, the
 * history rows of every causal conv, the last input of every transposed conv and a 72-window K/V cache per transformer
 * layer, so that the concatenation of the step outputs EQUALS Decoder.forward(whole sequence) (same arithmetic, same
 * rounding; oracle = modeling :868-883).  A stream's first step emits P*1920-555 samples (the look-ahead of the four
 * k = 2s transposed convs), every later step P*1920; the last 555 samples of the sequence are never emitted, as in the
 * reference.  Used by the talker's Dual-Track packet generation (model wrapper :838 hands codes to decode()).
 *
 * q3tts_stream_reserve sizes everything (state, K/V rings, workspace for max_streams x max_frames_per_step <= 57); no
 * allocation happens in q3tts_stream_step for shapes within the reservation. */
int q3tts_stream_reserve(q3tts_handle* h, int max_streams, int max_frames_per_step);
int q3tts_stream_open(q3tts_handle* h, int* stream_id, void* stream);
int q3tts_stream_reset(q3tts_handle* h, int stream_id, void* stream);
int q3tts_stream_close(q3tts_handle* h, int stream_id);
int64_t q3tts_stream_frames(q3tts_handle* h, int stream_id);   /* frames decoded so far, -1 = not open */
int64_t q3tts_stream_state_bytes(q3tts_handle* h);              /* device bytes of state per stream */
/* One step: stream_ids[S] (host), codes [S,Q,P] int64 (device), wav [S, wav_stride] fp32 (device, wav_stride >= P*1920),
 * n_valid[S] (host, optional) = samples written per row.  All launches go to `stream`; returns without synchronising (the stream
 * ids travel through a small ring of pinned staging blocks guarded by events, never through a stream synchronise). */
int q3tts_stream_step(q3tts_handle* h, const int32_t* stream_ids, int S, const int64_t* codes, int P, float* wav, int64_t wav_stride,
                      int32_t* n_valid, void* stream);

/* Parity probe: run forward() up to the stage named `stage` (an oracle tap name such as
 * "pre_conv", "pre_transformer", "upsample.0.1", "decoder.2.block.3", "decoder.0.snake") and write
 * that activation as fp32 [B,C,L] to `out` (device, `capacity` floats).  *C and *L receive the shape. */
int q3tts_forward_probe(q3tts_handle* h, const int64_t* codes, int B, int T, const char* stage, float* out, int64_t capacity, int* C, int* L,
                        void* stream);

/* Bench support: number of kernel launches issued by this handle since creation. */
int64_t q3tts_launch_count(q3tts_handle* h);

/* Per-op profile.  While profiling is on, a CUDA-event pair is recorded around every op on the launch stream and the profile
 * ACCUMULATES over every engine launch (all chunks / sub-batches / stream plans) until q3tts_get_profile reads it:
 * q3tts_get_profile synchronises the device, returns the number of ops recorded since the last read (fill at most max_ops
 * entries) and clears the profile.  types[i] = kernel: 0 RVQ gather-sum, 1 conv_gemm_tc_kernel (1-tap GEMMs), 2 RMSNorm,
 * 3 attention, 4 dwconv + LayerNorm, 5 output conv, 6 unit_kernel (fused residual unit), 7 ws_conv_kernel, 8 slab_conv_kernel
 * (multi-tap convs), 9 streaming attention, 10 stream state copies, 11 fp32 -> [hi | lo] split; encode path: 12 first conv, 13 activation +
 * split pass, 14 LayerNorm, 15 fp32 attention, 16 latents tap, 17 column split, 18 argmin + residual update; flops[i] = ALGORITHMIC 2 * MACs
 * (the extra MMAs of split-precision operands are not counted), bytes[i] = minimum HBM bytes (operands and outputs once);
 * dims (optional, 4 per op) = rows over all batch items, N, input channels, taps for GEMM ops. */
int q3tts_set_profile(q3tts_handle* h, int on);
int q3tts_get_profile(q3tts_handle* h, int max_ops, int32_t* types, float* ms, double* flops, double* bytes, int32_t* dims);

#ifdef __cplusplus
}
#endif
#endif /* Q3TTS_DECODE_H_ */
