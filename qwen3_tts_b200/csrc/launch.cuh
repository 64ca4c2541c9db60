// Kernel launches with programmatic dependent launch (PDL, sm_90+): a kernel launched through launch_k may START while the previous
// kernel of the stream is still draining - its blocks are scheduled once every block of the previous kernel has executed
// griddepcontrol.launch_dependents (pdl_trigger(), the first statement of every kernel here) or exited - and runs its prologue
// (barrier init, TMEM allocation, descriptor prefetch, parameter staging) in that shadow.  It must not touch memory the previous
// kernels produce, or memory they still read, before griddepcontrol.wait (pdl_wait()), which returns when all of them have
// completed and their writes are visible.  Rule kept in this library: ONLY kernels that execute pdl_wait() are launched through
// launch_k; every other kernel uses a plain launch and therefore serialises as usual.
#pragma once
#include <cuda_runtime.h>

#include <utility>

namespace q3 {

// PDL is used where it was measured to help: the streaming step (100 short launches: 2.89 -> 2.73 ms at 256 streams, 1.43 -> 1.26 ms
// for one stream, same-box A/B tools/ab_pdl.py).  It is OFF for bulk forwards: their kernels are persistent, one CTA per SM for
// milliseconds, and early scheduling of the next kernel's CTAs made the 64 x 10 s step 4 % SLOWER (45.3 -> 47.2 ms).  The engine opens a
// PdlScope around the launches of a streaming plan; q3tts_debug_set_pdl(0) disables it everywhere.
bool pdl_enabled();        // scope open on this thread AND the process-wide switch (default on) set
void pdl_set(bool on);
struct PdlScope {
  bool prev;
  explicit PdlScope(bool on);
  ~PdlScope();
};

template <typename... P, typename... A>
inline cudaError_t launch_k(void (*kernel)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, int cluster, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[2];
  unsigned n = 0;
  if (cluster > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = (unsigned)cluster; attr[n].val.clusterDim.y = 1; attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr; cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<A>(args)...);
}

}  // namespace q3
