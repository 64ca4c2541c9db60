// extern "C" surface of libq3tts_decode.so (declarations + reference citations: include/q3tts_decode.h).
#include <cstring>
#include <new>
#include <vector>

#include "engine.h"
#include "kernels.h"

using namespace q3;

namespace {

int chunk_count(int T, int chunk) { return T <= 0 ? 0 : (T + chunk - 1) / chunk; }

}  // namespace

extern "C" {

int q3tts_create(const q3tts_config* cfg, int device, q3tts_handle** out) {
  if (!cfg || !out) return Q3TTS_BAD_CONFIG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    cudaGetLastError();
    return Q3TTS_CUDA_ERROR;  // no CPU fallback: this library needs a CUDA device
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return Q3TTS_CUDA_ERROR;
  if (prop.major != 10) return Q3TTS_CUDA_ERROR;  // sm_100a only: tcgen05 / TMEM kernels
  if (cfg->num_quantizers < 2 || cfg->num_quantizers > 64 || cfg->num_upsample_rates < 1 || cfg->num_upsample_rates > 8 ||
      cfg->num_upsampling_ratios < 0 || cfg->num_upsampling_ratios > 4 || cfg->codebook_size < 1 || cfg->num_hidden_layers < 0)
    return Q3TTS_BAD_CONFIG;
  q3tts_handle* h = new (std::nothrow) q3tts_handle();
  if (!h) return Q3TTS_OOM;
  h->e.cfg = *cfg;
  h->e.device = device;
  *out = h;
  return Q3TTS_OK;
}

void q3tts_destroy(q3tts_handle* h) {
  if (!h) return;
  {
    DeviceGuard g(h->e.device);
    cudaDeviceSynchronize();
    engine_free(&h->e);
  }
  delete h;
}

const char* q3tts_last_error(q3tts_handle* h) { return h ? h->e.err.c_str() : "null handle"; }

int q3tts_load_tensor(q3tts_handle* h, const char* key, const float* data, int ndim, const int64_t* shape, int on_device) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (!key || !data || ndim < 1 || ndim > 4 || !shape) return e.fail(Q3TTS_BAD_SHAPE, "load_tensor: bad arguments");
  DeviceGuard g(e.device);
  RawTensor t;
  t.numel = 1;
  for (int i = 0; i < ndim; ++i) {
    if (shape[i] <= 0) return e.fail(Q3TTS_BAD_SHAPE, std::string("load_tensor: non-positive dim for ") + key);
    t.shape.push_back(shape[i]);
    t.numel *= shape[i];
  }
  if (cudaMalloc((void**)&t.d, (size_t)t.numel * sizeof(float)) != cudaSuccess) {
    cudaGetLastError();
    return e.fail(Q3TTS_OOM, std::string("load_tensor: cudaMalloc failed for ") + key);
  }
  cudaError_t err = cudaMemcpy(t.d, data, (size_t)t.numel * sizeof(float), on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice);
  if (err != cudaSuccess) {
    cudaFree(t.d);
    return e.fail(Q3TTS_CUDA_ERROR, std::string("load_tensor: copy failed: ") + cudaGetErrorString(err));
  }
  auto it = e.raw.find(key);
  if (it != e.raw.end()) {
    cudaFree(it->second.d);
    e.raw.erase(it);
  }
  e.raw[key] = t;
  e.finalized = false;
  return Q3TTS_OK;
}

int q3tts_finalize_weights(q3tts_handle* h, int compute_mode) {
  if (!h) return Q3TTS_BAD_CONFIG;
  DeviceGuard g(h->e.device);
  return engine_finalize(&h->e, compute_mode);
}

int q3tts_reserve(q3tts_handle* h, int max_frames_per_launch) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  if (max_frames_per_launch < 1) return e.fail(Q3TTS_BAD_SHAPE, "reserve: max_frames_per_launch must be positive");
  if (!e.finalized) return e.fail(Q3TTS_NOT_READY, "reserve: weights not finalized");
  DeviceGuard g(e.device);
  e.max_frames = max_frames_per_launch;
  // worst case is one long item: size for (1, max_frames)
  const size_t need = engine_plan_bytes(&e, 1, max_frames_per_launch);
  if (need > e.ws_bytes) {
    cudaDeviceSynchronize();
    if (e.ws) cudaFree(e.ws);
    e.ws = nullptr; e.ws_bytes = 0;
    e.plans.clear();
    if (cudaMalloc((void**)&e.ws, need) != cudaSuccess) {
      cudaGetLastError();
      return e.fail(Q3TTS_OOM, "reserve: workspace cudaMalloc failed");
    }
    e.ws_bytes = need;
  }
  return Q3TTS_OK;
}

int q3tts_set_engine(q3tts_handle* h, int engine) {
  if (!h) return Q3TTS_BAD_CONFIG;
  if (engine != Q3TTS_ENGINE_TC && engine != Q3TTS_ENGINE_SIMT) return h->e.fail(Q3TTS_BAD_CONFIG, "unknown engine");
  if (h->e.engine != engine) h->e.plans.clear();  // fused ops are TC-only: plans depend on the engine
  h->e.engine = engine;
  return Q3TTS_OK;
}

int64_t q3tts_forward_num_samples(q3tts_handle* h, int T) {
  if (!h || T <= 0) return 0;
  const int64_t n = (int64_t)T * h->e.total_up - h->e.lookahead;
  return n > 0 ? n : 0;
}

int64_t q3tts_chunked_num_samples(q3tts_handle* h, int T, int chunk, int left_ctx) {
  (void)left_ctx;
  if (!h || T <= 0 || chunk <= 0) return 0;
  return (int64_t)T * h->e.total_up - (int64_t)h->e.lookahead * chunk_count(T, chunk);
}

int64_t q3tts_launch_count(q3tts_handle* h) { return h ? h->e.launches : 0; }

int q3tts_set_profile(q3tts_handle* h, int on) {
  if (!h) return Q3TTS_BAD_CONFIG;
  h->e.profile = on != 0;
  prof_clear(&h->e);
  return Q3TTS_OK;
}

int q3tts_get_profile(q3tts_handle* h, int max_ops, int32_t* types, float* ms, double* flops, double* bytes, int32_t* dims) {
  if (!h) return Q3TTS_BAD_CONFIG;
  Engine& e = h->e;
  DeviceGuard g(e.device);
  if (cudaDeviceSynchronize() != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, "get_profile: sync failed");
  const int n = (int)e.prof_type.size();
  for (int i = 0; i < n && i < max_ops; ++i) {
    float t = 0.f;
    cudaEventElapsedTime(&t, e.prof_events[2 * i], e.prof_events[2 * i + 1]);
    types[i] = e.prof_type[i]; ms[i] = t; flops[i] = e.prof_flops[i]; bytes[i] = e.prof_bytes[i];
    if (dims) for (int d = 0; d < 4; ++d) dims[4 * i + d] = e.prof_dims[4 * i + d];
  }
  prof_clear(&e);  // read-and-clear: the next launches start a new profile
  return n;
}

}  // extern "C"
