#include "kernels.h"
#include "ptx.cuh"

namespace q3 {

static inline int nblocks(long long n, int threads, int cap = 148 * 32) {
  long long b = (n + threads - 1) / threads;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return (int)b;
}

#define Q3_DISPATCH(bf16, ...)                                   \
  do {                                                            \
    if (bf16) { using T = __nv_bfloat16; __VA_ARGS__; }           \
    else { using T = __half; __VA_ARGS__; }                       \
  } while (0)

// ============================================================================================ packers
// One packed weight element of row `row`, input channel ci (of cin): plain (m = 1), or carried as hi + lo for the split-precision
// GEMMs (gemm_params.h): m = 2 -> [hi | lo], m = 3 -> [hi | hi | lo] (the K' order that meets A = [a_hi | a_lo | a_hi]).
// cg (K group, 0 = cin): the row is cin / cg groups, each packed on its own as above - [hi(g) | hi(g) | lo(g)] per group - so that a
// launch can walk ONE group (a shorter accumulation chain, encode_plan.cu) through a tensor map with the full row pitch.
template <typename T>
__device__ __forceinline__ void put_w(T* __restrict__ dst, long long row, int ci, int cin, int m, float v, int cg = 0) {
  T* r = dst + row * (long long)(m * cin);
  if (cg > 0 && cg < cin) { r += (long long)(ci / cg) * (m * cg); ci %= cg; cin = cg; }
  const T hi = Elem<T>::from_f(v);
  r[ci] = hi;
  if (m > 1) {
    const T lo = Elem<T>::from_f(v - Elem<T>::to_f(hi));
    if (m == 2) r[cin + ci] = lo;
    else { r[cin + ci] = hi; r[2 * cin + ci] = lo; }
  }
}
template <typename T>
__global__ void k_pack_conv(const float* __restrict__ src, T* __restrict__ dst, int cout, int cin, int k, int m, int cg) {
  const long long n = (long long)cout * cin * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % cin);
    const int co = (int)((i / cin) % cout);
    const int j = (int)(i / ((long long)cin * cout));
    put_w<T>(dst, (long long)j * cout + co, ci, cin, m, src[((long long)co * cin + ci) * k + j], cg);
  }
}
cudaError_t pack_conv_weight(const float* src, void* dst, int cout, int cin, int k, bool bf16, cudaStream_t s, int split, int group) {
  const long long n = (long long)cout * cin * k;
  Q3_DISPATCH(bf16, (k_pack_conv<T><<<nblocks(n, 256), 256, 0, s>>>(src, (T*)dst, cout, cin, k, split, group)));
  return cudaGetLastError();
}

template <typename T>
__global__ void k_pack_transconv(const float* __restrict__ src, T* __restrict__ dst, int cin, int cout, int k, int st, int m) {
  const int taps = k / st;  // 2 (k = 2*st) or 1 (k = st)
  const long long n = (long long)taps * st * cout * cin;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % cin);
    const long long r = i / cin;
    const int co = (int)(r % cout);
    const int p = (int)((r / cout) % st);
    const int tap = (int)(r / ((long long)cout * st));
    // tap 0 multiplies x[q]; for k = 2*st that is kernel index p + st, tap 1 (x[q+1]) is index p
    const int kk = (taps == 2) ? (tap == 0 ? p + st : p) : p;
    put_w<T>(dst, r, ci, cin, m, src[((long long)ci * cout + co) * k + kk]);
  }
}
cudaError_t pack_transconv_weight(const float* src, void* dst, int cin, int cout, int k, int st, bool bf16, cudaStream_t s, int split) {
  const long long n = (long long)cin * cout * k;
  Q3_DISPATCH(bf16, (k_pack_transconv<T><<<nblocks(n, 256), 256, 0, s>>>(src, (T*)dst, cin, cout, k, st, split)));
  return cudaGetLastError();
}

template <typename T>
__global__ void k_pack_interleave2(const float* __restrict__ a, const float* __restrict__ b, T* __restrict__ dst, int n, int k, int m) {
  const long long tot = 2ll * n * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < tot; i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % k);
    const long long r = i / k;
    const float* src = (r & 1) ? b : a;
    put_w<T>(dst, r, c, k, m, src[(r >> 1) * k + c]);
  }
}
cudaError_t pack_interleave2(const float* a, const float* b, void* dst, int n, int k, bool bf16, cudaStream_t s, int split) {
  Q3_DISPATCH(bf16, (k_pack_interleave2<T><<<nblocks(2ll * n * k, 256), 256, 0, s>>>(a, b, (T*)dst, n, k, split)));
  return cudaGetLastError();
}

template <typename T>
__global__ void k_pack_cast(const float* __restrict__ src, T* __restrict__ dst, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    dst[i] = Elem<T>::from_f(src[i]);
}
cudaError_t pack_cast(const float* src, void* dst, long long n, bool bf16, cudaStream_t s) {
  Q3_DISPATCH(bf16, (k_pack_cast<T><<<nblocks(n, 256), 256, 0, s>>>(src, (T*)dst, n)));
  return cudaGetLastError();
}

// [rows][k] fp32 -> [rows][split * k] 16-bit (linear weights; split as in put_w)
template <typename T>
__global__ void k_pack_rows(const float* __restrict__ src, T* __restrict__ dst, long long rows, int k, int m, int cg) {
  const long long n = rows * k;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    put_w<T>(dst, i / k, (int)(i % k), k, m, src[i], cg);
}
cudaError_t pack_rows(const float* src, void* dst, long long rows, int k, bool bf16, cudaStream_t s, int split, int group) {
  Q3_DISPATCH(bf16, (k_pack_rows<T><<<nblocks(rows * k, 256), 256, 0, s>>>(src, (T*)dst, rows, k, split, group)));
  return cudaGetLastError();
}

// Strided Conv1d [cout][c][2r] (stride r, causal) as a two-tap GEMM over super-rows of r input rows (encode path): dst [2][cout][m * r * c],
// tap = j / r, K index = (j % r) * c + ci.
template <typename T>
__global__ void k_pack_strided_conv(const float* __restrict__ src, T* __restrict__ dst, int cout, int c, int r, int m, int cg) {
  const long long n = 2ll * r * cout * c;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int ci = (int)(i % c);
    long long q = i / c;
    const int p = (int)(q % r); q /= r;
    const int co = (int)(q % cout);
    const int tap = (int)(q / cout);
    put_w<T>(dst, (long long)tap * cout + co, p * c + ci, r * c, m, src[((long long)co * c + ci) * (2 * r) + tap * r + p], cg);
  }
}
cudaError_t pack_strided_conv_weight(const float* src, void* dst, int cout, int c, int r, bool bf16, cudaStream_t s, int split, int group) {
  Q3_DISPATCH(bf16, (k_pack_strided_conv<T><<<nblocks(2ll * r * cout * c, 256), 256, 0, s>>>(src, (T*)dst, cout, c, r, split, group)));
  return cudaGetLastError();
}

__global__ void k_pack_snake(const float* alpha, const float* beta, float* a, float* ib, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    // tables of the cosine form of SnakeBeta (epilogue.cuh snake()): 2*exp(alpha) and (1/2) / (exp(beta) + 1e-9)
    a[i] = 2.0f * expf(alpha[i]);
    ib[i] = 0.5f / (expf(beta[i]) + 1e-9f);
  }
}
cudaError_t pack_snake(const float* alpha, const float* beta, float* a, float* ib, int n, cudaStream_t s) {
  k_pack_snake<<<(n + 127) / 128, 128, 0, s>>>(alpha, beta, a, ib, n);
  return cudaGetLastError();
}

__global__ void k_pack_dw(const float* __restrict__ src, float* __restrict__ dst, int C, int k) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < C * k) {
    const int c = i % C, j = i / C;
    dst[i] = src[c * k + j];
  }
}
cudaError_t pack_dw_weight(const float* src, float* dst, int C, int k, cudaStream_t s) {
  k_pack_dw<<<(C * k + 127) / 128, 128, 0, s>>>(src, dst, C, k);
  return cudaGetLastError();
}

// folded[code][d] = sum_j (es[code][j] / max(usage[code], 1e-5)) * wout[d][j]; one block per code.
__global__ void k_fold_codebook(const float* __restrict__ es, const float* __restrict__ usage, const float* __restrict__ wout,
                                float* __restrict__ folded, int dq, int d) {
  extern __shared__ float row[];
  const int code = blockIdx.x;
  const float u = fmaxf(usage[code], 1e-5f);
  for (int j = threadIdx.x; j < dq; j += blockDim.x) row[j] = es[(long long)code * dq + j] / u;
  __syncthreads();
  for (int o = threadIdx.x; o < d; o += blockDim.x) {
    const float* w = wout + (long long)o * dq;
    float acc = 0.f;
    for (int j = 0; j < dq; ++j) acc = fmaf(row[j], w[j], acc);
    folded[(long long)code * d + o] = acc;
  }
}
cudaError_t fold_codebook(const float* es, const float* usage, const float* wout, float* folded, int ncodes, int dq, int d, cudaStream_t s) {
  k_fold_codebook<<<ncodes, 256, dq * sizeof(float), s>>>(es, usage, wout, folded, dq, d);
  return cudaGetLastError();
}

}  // namespace q3
