// q3tts_rvq_decode_i16_static: the RVQ gather-sum on int16 code frames with NO allocation anywhere in the call.
//
// The int16 wire format needs scratch: the gather kernel reads int64 codes, so the int16 frames are widened into a
// staging buffer first (q3tts_decode_host_pcm16 keeps that buffer in cudaMalloc'ed memory that grows on demand).  Callers that
// cannot call cudaMalloc - code under CUDA graph capture, tight serving loops, harnesses that forbid it - get this variant, whose
// staging lives in a file-scope __device__ array of a fixed, modest size.  Every GPU that loads the library holds the array until
// the process exits.
//
// Rules that follow from that:
//  * In HOST code the array's name is a host-side stand-in, not a device address: the address a kernel can use comes from
//    cudaGetSymbolAddress, queried once per device with that device current (static_scratch_setup below, run by
//    q3tts_finalize_weights), and is passed to the kernels.
//  * Every call on a device shares the one array, and it keeps the last call's contents.  Nothing here needs to start at zero: the
//    widening kernel overwrites exactly the B*Q*T entries the gather kernel then reads.  Calls that overlap on the GPU (two streams,
//    or two handles on the same device) RACE on it: serialise them.
//  * Sizes beyond the maximum are rejected (Q3TTS_BAD_SHAPE), never truncated.
#include <string>

#include "engine.h"
#include "kernels.h"

namespace {

constexpr long long kMaxCodes = Q3TTS_STATIC_MAX_CODES;  // 512 Ki codes = 32768 frames of 16 codebooks (43.7 min of audio) = 4 MiB as int64
static __device__ __align__(256) long long g_codes_scratch[kMaxCodes];

constexpr int kMaxDevices = 64;
long long* g_scratch_addr[kMaxDevices] = {nullptr};  // per device ordinal: device address of g_codes_scratch on THAT device

}  // namespace

namespace q3 {

// Set-up, once per device, with that device current (engine_finalize calls it beside the other per-device set-ups): resolves the
// device address of the scratch array on THIS device.  The entry point below only reads the cached address, so it makes no
// runtime query of its own either.
int static_scratch_setup(int device) {
  if (device < 0 || device >= kMaxDevices) return -1;
  if (g_scratch_addr[device]) return 0;
  void* p = nullptr;
  const cudaError_t err = cudaGetSymbolAddress(&p, g_codes_scratch);
  if (err != cudaSuccess || !p) { cudaGetLastError(); return -2; }
  g_scratch_addr[device] = static_cast<long long*>(p);
  return 0;
}

}  // namespace q3

extern "C" {

int64_t q3tts_static_max_codes(void) { return kMaxCodes; }

int q3tts_rvq_decode_i16_static(q3tts_handle* h, const int16_t* codes, int B, int T, float* out, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  q3::Engine& e = h->e;
  if (!e.finalized) return e.fail(Q3TTS_NOT_READY, "weights not finalized");
  if (!codes || !out || B < 1 || T < 1) return e.fail(Q3TTS_BAD_SHAPE, "rvq_decode_i16_static: codes must be [B>=1, Q, T>=1] and out non-null");
  const int Q = e.cfg.num_quantizers;
  const long long n = (long long)B * Q * T;
  if (n > kMaxCodes)
    return e.fail(Q3TTS_BAD_SHAPE, "rvq_decode_i16_static: B*Q*T = " + std::to_string(n) + " codes exceed the static scratch of " + std::to_string(kMaxCodes) +
                                       " codes (q3tts_static_max_codes); split the call or use q3tts_rvq_decode");
  q3::DeviceGuard g(e.device);
  long long* scratch = (e.device >= 0 && e.device < kMaxDevices) ? g_scratch_addr[e.device] : nullptr;
  if (!scratch) return e.fail(Q3TTS_NOT_READY, "rvq_decode_i16_static: the static scratch was not set up on this device (q3tts_finalize_weights does it)");
  cudaStream_t s = (cudaStream_t)stream;
  cudaError_t err = q3::widen_codes_i16(reinterpret_cast<const short*>(codes), scratch, n, s);
  if (err == cudaSuccess)
    err = q3::rvq_decode(scratch, (long long)Q * T, T, 1, e.rvq_tables, nullptr, out, B, Q, T, e.cfg.codebook_size, e.cfg.codebook_dim, e.bf16, s, 0, nullptr,
                         e.d_bad_code);
  e.launches += 2;
  if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("rvq_decode_i16_static: ") + cudaGetErrorString(err));
  return Q3TTS_OK;
}

}  // extern "C"
