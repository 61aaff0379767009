// Transformer-side kernels for the GPT-2 / MLP configs of BASELINE.json (absent from the reference):
// LayerNorm fwd/bwd, GELU(tanh) fwd/bwd, residual add, token+position embedding fwd/bwd, causal softmax fwd/bwd for the
// attention score matrices produced by the batched tcgen05 GEMM, plus small bf16 helpers (ReLU backward, column sums,
// casts, scaled copies).  All templated on the storage type (fp32 or bf16), math in fp32.  sm_100a.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/ddpx_kernels.h"

namespace {

typedef __nv_bfloat16 bf16;

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<bf16>(const bf16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<bf16>(bf16* p, float v) { *p = __float2bfloat16(v); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
// block-wide sum of two values at once (blockDim.x <= 1024)
__device__ __forceinline__ void block_sum2(float& a, float& b, float* sh) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  a = warp_sum(a); b = warp_sum(b);
  __syncthreads();
  if (lane == 0) { sh[w] = a; sh[32 + w] = b; }
  __syncthreads();
  float x = (threadIdx.x < nw) ? sh[threadIdx.x] : 0.f;
  float y = (threadIdx.x < nw) ? sh[32 + threadIdx.x] : 0.f;
  if (w == 0) {
    x = warp_sum(x); y = warp_sum(y);
    if (lane == 0) { sh[0] = x; sh[32] = y; }
  }
  __syncthreads();
  a = sh[0]; b = sh[32];
}

inline int cap_blocks(long long n, int cap = 148 * 16) { return static_cast<int>(n < 1 ? 1 : (n > cap ? cap : n)); }

// ------------------------------------------------------------------------------------------------ LayerNorm
// one block per row (grid-stride); saves mean / rstd (fp32) for the backward
template <typename T>
__global__ void layernorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ gamma, const T* __restrict__ beta,
                                     T* __restrict__ y, float* __restrict__ mean, float* __restrict__ rstd, long long rows,
                                     int C, float eps) {
  __shared__ float sh[64];
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const T* xr = x + r * C;
    float s = 0.f, q = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) { const float v = ldf(xr + c); s += v; q += v * v; }
    block_sum2(s, q, sh);
    const float m = s / C;
    const float rs = rsqrtf(fmaxf(q / C - m * m, 0.f) + eps);
    if (threadIdx.x == 0) { mean[r] = m; rstd[r] = rs; }
    T* yr = y + r * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x)
      stf(yr + c, (ldf(xr + c) - m) * rs * ldf(gamma + c) + ldf(beta + c));
  }
}

// dx per row + per-block partial dgamma/dbeta accumulated into fp32 [C] buffers with atomics (pre-zeroed)
template <typename T>
__global__ void layernorm_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, const T* __restrict__ gamma,
                                     const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                                     float* __restrict__ dgamma, float* __restrict__ dbeta, long long rows, int C) {
  extern __shared__ float acc[];   // [2][C] per-block partial dgamma / dbeta
  __shared__ float sh[64];
  for (int c = threadIdx.x; c < 2 * C; c += blockDim.x) acc[c] = 0.f;
  __syncthreads();
  for (long long r = blockIdx.x; r < rows; r += gridDim.x) {
    const T* xr = x + r * C;
    const T* dr = dy + r * C;
    const float m = mean[r], rs = rstd[r];
    float s1 = 0.f, s2 = 0.f;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float g = ldf(dr + c) * ldf(gamma + c);
      const float xh = (ldf(xr + c) - m) * rs;
      s1 += g; s2 += g * xh;
    }
    block_sum2(s1, s2, sh);
    const float a = s1 / C, b = s2 / C;
    T* dxr = dx + r * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float d = ldf(dr + c);
      const float xh = (ldf(xr + c) - m) * rs;
      stf(dxr + c, rs * (d * ldf(gamma + c) - a - xh * b));
      acc[c] += d * xh;        // each thread owns its columns: no race
      acc[C + c] += d;
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += blockDim.x) {
    atomicAdd(dgamma + c, acc[c]);
    atomicAdd(dbeta + c, acc[C + c]);
  }
}

// ------------------------------------------------------------------------------------------------ GELU / add / casts
__device__ __forceinline__ float gelu_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  return 0.5f * x * (1.f + tanhf(k0 * (x + k1 * x * x * x)));
}
__device__ __forceinline__ float gelu_grad_f(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  const float u = k0 * (x + k1 * x * x * x);
  const float t = tanhf(u);
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k0 * (1.f + 3.f * k1 * x * x);
}
template <typename T>
__global__ void gelu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long long n) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    stf(y + i, gelu_f(ldf(x + i)));
}
template <typename T>
__global__ void gelu_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, T* __restrict__ dx, long long n) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    stf(dx + i, ldf(dy + i) * gelu_grad_f(ldf(x + i)));
}
template <typename T>
__global__ void relu_bwd_t_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, long long n) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    stf(dx + i, ldf(y + i) > 0.f ? ldf(dy + i) : 0.f);
}
template <typename T>
__global__ void add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, long long n) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    stf(out + i, ldf(a + i) + ldf(b + i));
}
template <typename TS, typename TD>
__global__ void cast_scale_kernel(const TS* __restrict__ src, TD* __restrict__ dst, long long n, float scale) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    stf(dst + i, ldf(src + i) * scale);
}
template <typename T>
__global__ void colsum_t_kernel(const T* __restrict__ x, float* __restrict__ out, long long rows, int cols) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (long long r = blockIdx.y; r < rows; r += gridDim.y) s += ldf(x + r * cols + c);
  atomicAdd(out + c, s);
}
// inverted dropout with a counter-based hash RNG: keep iff hash(seed, i) >= p * 2^32
__device__ __forceinline__ uint32_t hash32(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return static_cast<uint32_t>(x);
}
template <typename T>
__global__ void dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, float p, float scale, uint64_t seed) {
  const uint32_t thr = static_cast<uint32_t>(fminf(p, 0.999999f) * 4294967296.0f);
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    stf(y + i, hash32(seed * 0x9E3779B97F4A7C15ULL + static_cast<uint64_t>(i)) >= thr ? ldf(x + i) * scale : 0.f);
}

}  // namespace

#define LAUNCH_OK() (cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH)
#define ST(s) static_cast<cudaStream_t>(s)
#define EW_GRID(n) cap_blocks(((n) + 255) / 256)

extern "C" int ddpx_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                                  long long rows, int C, float eps, int is_bf16, void* stream) {
  const int block = C >= 1024 ? 256 : 128;
  const int grid = cap_blocks(rows, 148 * 8);
  if (is_bf16)
    layernorm_fwd_kernel<bf16><<<grid, block, 0, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(gamma),
                                                            static_cast<const bf16*>(beta), static_cast<bf16*>(y), mean, rstd,
                                                            rows, C, eps);
  else
    layernorm_fwd_kernel<float><<<grid, block, 0, ST(stream)>>>(static_cast<const float*>(x), static_cast<const float*>(gamma),
                                                             static_cast<const float*>(beta), static_cast<float*>(y), mean,
                                                             rstd, rows, C, eps);
  return LAUNCH_OK();
}

// dgamma / dbeta are fp32 [C] accumulators that must be zero on entry
extern "C" int ddpx_layernorm_bwd(const void* x, const void* dy, const void* gamma, const float* mean, const float* rstd,
                                  void* dx, float* dgamma, float* dbeta, long long rows, int C, int is_bf16, void* stream) {
  const int block = C >= 1024 ? 256 : 128;
  const int grid = cap_blocks(rows, 148 * 2);
  const size_t smem = static_cast<size_t>(2) * C * sizeof(float);
  if (smem > 48 * 1024) return DDPX_ERR_SHAPE;
  if (is_bf16)
    layernorm_bwd_kernel<bf16><<<grid, block, smem, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy),
                                                               static_cast<const bf16*>(gamma), mean, rstd,
                                                               static_cast<bf16*>(dx), dgamma, dbeta, rows, C);
  else
    layernorm_bwd_kernel<float><<<grid, block, smem, ST(stream)>>>(static_cast<const float*>(x), static_cast<const float*>(dy),
                                                                static_cast<const float*>(gamma), mean, rstd,
                                                                static_cast<float*>(dx), dgamma, dbeta, rows, C);
  return LAUNCH_OK();
}

extern "C" int ddpx_gelu_fwd(const void* x, void* y, long long n, int is_bf16, void* stream) {
  if (is_bf16) gelu_fwd_kernel<bf16><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<bf16*>(y), n);
  else gelu_fwd_kernel<float><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const float*>(x), static_cast<float*>(y), n);
  return LAUNCH_OK();
}
extern "C" int ddpx_gelu_bwd(const void* x, const void* dy, void* dx, long long n, int is_bf16, void* stream) {
  if (is_bf16)
    gelu_bwd_kernel<bf16><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy),
                                                           static_cast<bf16*>(dx), n);
  else
    gelu_bwd_kernel<float><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const float*>(x), static_cast<const float*>(dy),
                                                            static_cast<float*>(dx), n);
  return LAUNCH_OK();
}
extern "C" int ddpx_relu_bwd_bf16(const void* y, const void* dy, void* dx, long long n, void* stream) {
  relu_bwd_t_kernel<bf16><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const bf16*>(y), static_cast<const bf16*>(dy),
                                                           static_cast<bf16*>(dx), n);
  return LAUNCH_OK();
}
extern "C" int ddpx_add(const void* a, const void* b, void* out, long long n, int is_bf16, void* stream) {
  if (is_bf16)
    add_kernel<bf16><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const bf16*>(a), static_cast<const bf16*>(b),
                                                      static_cast<bf16*>(out), n);
  else
    add_kernel<float><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const float*>(a), static_cast<const float*>(b),
                                                       static_cast<float*>(out), n);
  return LAUNCH_OK();
}
// dst = src * scale with a dtype change: kind 0 = f32->bf16, 1 = bf16->f32, 2 = bf16->bf16
extern "C" int ddpx_cast_scale(const void* src, void* dst, long long n, float scale, int kind, void* stream) {
  if (kind == 0)
    cast_scale_kernel<float, bf16><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const float*>(src), static_cast<bf16*>(dst), n, scale);
  else if (kind == 1)
    cast_scale_kernel<bf16, float><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const bf16*>(src), static_cast<float*>(dst), n, scale);
  else
    cast_scale_kernel<bf16, bf16><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const bf16*>(src), static_cast<bf16*>(dst), n, scale);
  return LAUNCH_OK();
}
// out[c] = sum_r x[r][c] for bf16 x (fp32 out; zeroed here)
extern "C" int ddpx_colsum_bf16(const void* x, float* out, long long rows, int cols, void* stream) {
  int gy = static_cast<int>(rows / 64);
  if (gy < 1) gy = 1;
  if (gy > 256) gy = 256;
  cudaMemsetAsync(out, 0, sizeof(float) * cols, ST(stream));
  dim3 grid((cols + 127) / 128, gy);
  colsum_t_kernel<bf16><<<grid, 128, 0, ST(stream)>>>(static_cast<const bf16*>(x), out, rows, cols);
  return LAUNCH_OK();
}
extern "C" int ddpx_dropout(const void* x, void* y, long long n, float p, unsigned long long seed, int is_bf16, void* stream) {
  const float scale = 1.f / (1.f - p);
  if (is_bf16)
    dropout_kernel<bf16><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<bf16*>(y), n, p, scale, seed);
  else
    dropout_kernel<float><<<EW_GRID(n), 256, 0, ST(stream)>>>(static_cast<const float*>(x), static_cast<float*>(y), n, p, scale, seed);
  return LAUNCH_OK();
}
