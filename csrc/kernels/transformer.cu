// Transformer-side kernels for the GPT-2 / MLP configs of BASELINE.json (absent from the reference):
// LayerNorm fwd/bwd, GELU(tanh) fwd/bwd, residual add, bf16 helpers (ReLU backward, column sums, casts, scaled copies),
// dropout.  Templated on the storage type (fp32 or bf16), math in fp32, 16-byte vector accesses, sm_100a.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>
#include "vec.cuh"
#include "../../include/ddpx_kernels.h"

namespace {

using namespace ddpx;
typedef __nv_bfloat16 bf16;

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<bf16>(const bf16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stf(T* p, float v);
template <> __device__ __forceinline__ void stf<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stf<bf16>(bf16* p, float v) { *p = __float2bfloat16(v); }

inline int cap_blocks(long long n, int cap = 148 * 16) { return static_cast<int>(n < 1 ? 1 : (n > cap ? cap : n)); }

// ------------------------------------------------------------------------------------------------ LayerNorm
// One warp per row, 16-byte loads; C must be a multiple of the vector width.  mean / rstd (fp32) are saved.
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_fwd_kernel(const T* __restrict__ x, const T* __restrict__ gamma, const T* __restrict__ beta, T* __restrict__ y,
                     float* __restrict__ mean, float* __restrict__ rstd, long long rows, int C, float eps) {
  constexpr int V = VecOf<T>::N;
  const int lane = threadIdx.x & 31;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const int nvec = C / V;
  for (long long r = warp0; r < rows; r += nwarps) {
    const T* xr = x + r * C;
    float s = 0.f, q = 0.f;
    for (int i = lane; i < nvec; i += 32) {
      float f[V];
      vload(xr + i * V, f);
#pragma unroll
      for (int k = 0; k < V; ++k) { s += f[k]; q += f[k] * f[k]; }
    }
    s = warp_sum_f(s); q = warp_sum_f(q);
    const float m = s / C;
    const float rs = rsqrtf(fmaxf(q / C - m * m, 0.f) + eps);
    if (lane == 0) { mean[r] = m; rstd[r] = rs; }
    T* yr = y + r * C;
    for (int i = lane; i < nvec; i += 32) {
      float f[V], g[V], b[V];
      vload(xr + i * V, f); vload(gamma + i * V, g); vload(beta + i * V, b);
#pragma unroll
      for (int k = 0; k < V; ++k) f[k] = (f[k] - m) * rs * g[k] + b[k];
      vstore(yr + i * V, f);
    }
  }
}

// Backward: one warp per row; each lane keeps the dgamma / dbeta partial sums of its own columns in registers over all
// the rows its warp visits (NV vectors per lane), then the block folds them through shared memory and issues one
// atomic per column into the fp32 accumulators (pre-zeroed).
template <typename T, int NV>
__global__ void __launch_bounds__(256, (NV <= 2 ? 2 : 1))
layernorm_bwd_kernel(const T* __restrict__ x, const T* __restrict__ dy, const T* __restrict__ gamma,
                     const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                     float* __restrict__ dgamma, float* __restrict__ dbeta, long long rows, int C,
                     float* __restrict__ partial, const T* __restrict__ dres) {
  constexpr int V = VecOf<T>::N;
  // A row is loaded ONCE, as raw 16-byte vectors issued back to back (predicated, no branches in between), and kept packed
  // in registers for the second pass; the residual gradient's vectors are in flight during the first pass.  With a
  // branch and a load-use pair per vector the NV loads of a pass were serialised - NV full memory round trips per pass,
  // 27 us for a 2048 x 1600 call whose traffic is worth 4 us.  Rows wider than 8 vectors per lane (C > 2048 in bf16) would
  // not fit the register file this way and take the vector-at-a-time path.
  constexpr bool kBatch = NV <= 8;
  extern __shared__ float acc[];   // [2][C]
  const int lane = threadIdx.x & 31;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const int nvec = C / V;
  float ag[NV][V], ab[NV][V];
#pragma unroll
  for (int j = 0; j < NV; ++j)
#pragma unroll
    for (int k = 0; k < V; ++k) { ag[j][k] = 0.f; ab[j][k] = 0.f; }
  const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
  for (long long r = warp0; r < rows; r += nwarps) {
    const T* xr = x + r * C;
    const T* dr = dy + r * C;
    const float m = mean[r], rs = rstd[r];
    float s1 = 0.f, s2 = 0.f;
    T* dxr = dx + r * C;
    if constexpr (kBatch) {
      uint4 rx[NV], rd[NV], rr[NV];
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int i = lane + j * 32;
        const bool in = i < nvec;
        rx[j] = in ? __ldg(reinterpret_cast<const uint4*>(xr + i * V)) : zero4;
        rd[j] = in ? __ldg(reinterpret_cast<const uint4*>(dr + i * V)) : zero4;
      }
      if (dres != nullptr) {
#pragma unroll
        for (int j = 0; j < NV; ++j) {
          const int i = lane + j * 32;
          rr[j] = (i < nvec) ? __ldg(reinterpret_cast<const uint4*>(dres + r * C + i * V)) : zero4;
        }
      }
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int i = lane + j * 32;
        float f[V], d[V], g[V];
        vunpack<T>(rx[j], f); vunpack<T>(rd[j], d);
        vunpack<T>(i < nvec ? __ldg(reinterpret_cast<const uint4*>(gamma + i * V)) : zero4, g);
#pragma unroll
        for (int k = 0; k < V; ++k) {            // lanes past the row hold zeros: d = 0 keeps every sum unchanged
          const float xh = (f[k] - m) * rs;
          const float gd = d[k] * g[k];
          s1 += gd; s2 += gd * xh;
          ag[j][k] += d[k] * xh;
          ab[j][k] += d[k];
        }
      }
      s1 = warp_sum_f(s1); s2 = warp_sum_f(s2);
      const float a = s1 / C, b = s2 / C;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int i = lane + j * 32;
        if (i < nvec) {
          float f[V], d[V], g[V];
          vunpack<T>(rx[j], f); vunpack<T>(rd[j], d);
          vload(gamma + i * V, g);               // L1 hit
#pragma unroll
          for (int k = 0; k < V; ++k) {
            const float xh = (f[k] - m) * rs;
            f[k] = rs * (d[k] * g[k] - a - xh * b);
          }
          if (dres != nullptr) {                  // the gradient arriving over the residual connection around this LayerNorm
            float e[V];
            vunpack<T>(rr[j], e);
#pragma unroll
            for (int k = 0; k < V; ++k) f[k] += e[k];
          }
          vstore(dxr + i * V, f);
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int i = lane + j * 32;
        if (i < nvec) {
          float f[V], d[V], g[V];
          vload(xr + i * V, f); vload(dr + i * V, d); vload(gamma + i * V, g);
#pragma unroll
          for (int k = 0; k < V; ++k) {
            const float xh = (f[k] - m) * rs;
            const float gd = d[k] * g[k];
            s1 += gd; s2 += gd * xh;
            ag[j][k] += d[k] * xh;
            ab[j][k] += d[k];
          }
        }
      }
      s1 = warp_sum_f(s1); s2 = warp_sum_f(s2);
      const float a = s1 / C, b = s2 / C;
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int i = lane + j * 32;
        if (i < nvec) {
          float f[V], d[V], g[V];
          vload(xr + i * V, f); vload(dr + i * V, d); vload(gamma + i * V, g);
#pragma unroll
          for (int k = 0; k < V; ++k) {
            const float xh = (f[k] - m) * rs;
            f[k] = rs * (d[k] * g[k] - a - xh * b);
          }
          if (dres != nullptr) {
            float e[V];
            vload(dres + r * C + i * V, e);
#pragma unroll
            for (int k = 0; k < V; ++k) f[k] += e[k];
          }
          vstore(dxr + i * V, f);
        }
      }
    }
  }
  // fold the warps' register sums into acc one warp at a time (plain vector read-modify-write: shared-memory float
  // atomics compile to compare-and-swap spin loops, and 2 * NV * V of them per lane with eight warps on the same
  // addresses cost more than the rest of the kernel)
  const int nw = blockDim.x >> 5, wid = threadIdx.x >> 5;
  for (int w = 0; w < nw; ++w) {
    if (wid == w) {
#pragma unroll
      for (int j = 0; j < NV; ++j) {
        const int i = lane + j * 32;
        if (i < nvec) {
#pragma unroll
          for (int k = 0; k < V; k += 4) {
            float4* pg = reinterpret_cast<float4*>(&acc[i * V + k]);
            float4* pb = reinterpret_cast<float4*>(&acc[C + i * V + k]);
            float4 g4 = make_float4(ag[j][k], ag[j][k + 1], ag[j][k + 2], ag[j][k + 3]);
            float4 b4 = make_float4(ab[j][k], ab[j][k + 1], ab[j][k + 2], ab[j][k + 3]);
            if (w > 0) {
              const float4 og = *pg, ob = *pb;
              g4.x += og.x; g4.y += og.y; g4.z += og.z; g4.w += og.w;
              b4.x += ob.x; b4.y += ob.y; b4.z += ob.z; b4.w += ob.w;
            }
            *pg = g4; *pb = b4;
          }
        }
      }
    }
    __syncthreads();
  }
  if (partial != nullptr) {
    // the block's sums go to its own row of the scratch matrix [grid][2C] with plain stores; ln_partial_reduce_kernel adds
    // the rows up.  (Reductions from every block onto the same C addresses cost ~30 us per call, whatever the tensor size.)
    float* row = partial + static_cast<size_t>(blockIdx.x) * 2 * C;
    for (int c = threadIdx.x * 4; c < 2 * C; c += blockDim.x * 4)
      *reinterpret_cast<float4*>(row + c) = make_float4(acc[c], acc[c + 1], acc[c + 2], acc[c + 3]);
    return;
  }
  // one 16-byte vector reduction per four columns (C % 4 == 0; dgamma / dbeta are 16-byte aligned: checked on the host)
  for (int c = threadIdx.x * 4; c < C; c += blockDim.x * 4) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dgamma + c), "f"(acc[c]), "f"(acc[c + 1]), "f"(acc[c + 2]),
                 "f"(acc[c + 3]) : "memory");
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dbeta + c), "f"(acc[C + c]), "f"(acc[C + c + 1]),
                 "f"(acc[C + c + 2]), "f"(acc[C + c + 3]) : "memory");
  }
}

// ---- the same backward as two kernels with more parallelism (rows up to 8 vectors per lane): the column sums are what
// force the fused kernel to keep 2 * NV * V accumulators per lane alive across rows and to run at eight warps per SM.
// (1) dx alone: one row per warp at a time, nothing carried between rows, 16+ warps per SM.
template <typename T, int NV>
__global__ void __launch_bounds__(256, 2)
layernorm_bwd_dx_kernel(const T* __restrict__ x, const T* __restrict__ dy, const T* __restrict__ gamma,
                        const float* __restrict__ mean, const float* __restrict__ rstd, T* __restrict__ dx,
                        long long rows, int C, const T* __restrict__ dres) {
  constexpr int V = VecOf<T>::N;
  const int lane = threadIdx.x & 31;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const int nvec = C / V;
  const uint4 zero4 = make_uint4(0u, 0u, 0u, 0u);
  for (long long r = warp0; r < rows; r += nwarps) {
    const T* xr = x + r * C;
    const T* dr = dy + r * C;
    const float m = mean[r], rs = rstd[r];
    uint4 rx[NV], rd[NV], rg[NV], rr[NV];
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = lane + j * 32;
      const bool in = i < nvec;
      rx[j] = in ? __ldg(reinterpret_cast<const uint4*>(xr + i * V)) : zero4;
      rd[j] = in ? __ldg(reinterpret_cast<const uint4*>(dr + i * V)) : zero4;
      rg[j] = in ? __ldg(reinterpret_cast<const uint4*>(gamma + i * V)) : zero4;
      rr[j] = (in && dres != nullptr) ? __ldg(reinterpret_cast<const uint4*>(dres + r * C + i * V)) : zero4;
    }
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      float f[V], d[V], g[V];
      vunpack<T>(rx[j], f); vunpack<T>(rd[j], d); vunpack<T>(rg[j], g);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const float gd = d[k] * g[k];
        s1 += gd; s2 += gd * ((f[k] - m) * rs);
      }
    }
    s1 = warp_sum_f(s1); s2 = warp_sum_f(s2);
    const float a = s1 / C, b = s2 / C;
    T* dxr = dx + r * C;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = lane + j * 32;
      if (i < nvec) {
        float f[V], d[V], g[V], e[V];
        vunpack<T>(rx[j], f); vunpack<T>(rd[j], d); vunpack<T>(rg[j], g); vunpack<T>(rr[j], e);
#pragma unroll
        for (int k = 0; k < V; ++k) f[k] = rs * (d[k] * g[k] - a - ((f[k] - m) * rs) * b) + e[k];
        vstore(dxr + i * V, f);
      }
    }
  }
}
// (2) partial[by][c] = sum over the block's rows of dy * x-hat, partial[by][C + c] = sum of dy: block = 32 column vectors x
// 8 row lanes, rows split over gridDim.y; ln_partial_reduce_kernel adds the gridDim.y rows up
template <typename T>
__global__ void __launch_bounds__(256)
layernorm_bwd_colsum_kernel(const T* __restrict__ x, const T* __restrict__ dy, const float* __restrict__ mean,
                            const float* __restrict__ rstd, float* __restrict__ partial, long long rows, int C) {
  constexpr int V = VecOf<T>::N;
  __shared__ float shg[8][32][V + 1], shb[8][32][V + 1];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int cv = blockIdx.x * 32 + tx;               // column vector
  float sg[V], sb[V];
#pragma unroll
  for (int k = 0; k < V; ++k) { sg[k] = 0.f; sb[k] = 0.f; }
  if (cv * V < C) {
#pragma unroll 4
    for (long long r = static_cast<long long>(blockIdx.y) * 8 + ty; r < rows; r += static_cast<long long>(gridDim.y) * 8) {
      float f[V], d[V];
      vload(x + r * C + cv * V, f); vload(dy + r * C + cv * V, d);
      const float m = mean[r], rs = rstd[r];
#pragma unroll
      for (int k = 0; k < V; ++k) { sg[k] += d[k] * ((f[k] - m) * rs); sb[k] += d[k]; }
    }
  }
#pragma unroll
  for (int k = 0; k < V; ++k) { shg[ty][tx][k] = sg[k]; shb[ty][tx][k] = sb[k]; }
  __syncthreads();
  // 256 threads fold the 8 row lanes of the block's 32 * V columns
  for (int t = threadIdx.x; t < 32 * V; t += 256) {
    const int g = t / V, k = t - g * V;
    const int c = (blockIdx.x * 32 + g) * V + k;
    if (c < C) {
      float a = 0.f, b = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { a += shg[j][g][k]; b += shb[j][g][k]; }
      float* row = partial + static_cast<size_t>(blockIdx.y) * 2 * C;
      row[c] = a; row[C + c] = b;
    }
  }
}

// out[c] = sum over the nrows rows of partial[nrows][n2] (n2 = 2C: dgamma | dbeta).  Block = 32 columns x 8 row groups:
// a thread's loads are independent (a single thread walking all rows pays an L2 round trip per few loads).
__global__ void __launch_bounds__(256) ln_partial_reduce_kernel(const float* __restrict__ partial, int nrows, int n2, int C,
                                                               float* __restrict__ dgamma, float* __restrict__ dbeta) {
  __shared__ float red[8][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + tx;
  float s = 0.f;
  if (c < n2) {
#pragma unroll 8
    for (int r = ty; r < nrows; r += 8) s += __ldcg(partial + static_cast<size_t>(r) * n2 + c);
  }
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && c < n2) {
#pragma unroll
    for (int j = 1; j < 8; ++j) s += red[j][tx];
    if (c < C) dgamma[c] = s; else dbeta[c - C] = s;
  }
}

// ------------------------------------------------------------------------------------------------ elementwise
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

enum { OP_GELU_FWD = 0, OP_GELU_BWD = 1, OP_RELU_BWD = 2, OP_ADD = 3 };

// out = op(a[, b]) over n elements; vector body + scalar tail
template <typename T, int OP>
__global__ void __launch_bounds__(256)
ew_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ out, long long n) {
  constexpr int V = VecOf<T>::N;
  const long long nvec = n / V;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i = tid; i < nvec; i += nth) {
    float fa[V], fb[V];
    vload(a + i * V, fa);
    if (OP != OP_GELU_FWD) vload(b + i * V, fb);
#pragma unroll
    for (int k = 0; k < V; ++k) {
      if (OP == OP_GELU_FWD) fa[k] = gelu_f(fa[k]);
      else if (OP == OP_GELU_BWD) fa[k] = fb[k] * gelu_grad_f(fa[k]);   // a = pre-activation, b = dy
      else if (OP == OP_RELU_BWD) fa[k] = fa[k] > 0.f ? fb[k] : 0.f;    // a = y, b = dy
      else fa[k] = fa[k] + fb[k];
    }
    vstore(out + i * V, fa);
  }
  for (long long i = nvec * V + tid; i < n; i += nth) {
    const float x = ldf(a + i);
    const float y = (OP != OP_GELU_FWD) ? ldf(b + i) : 0.f;
    float r;
    if (OP == OP_GELU_FWD) r = gelu_f(x);
    else if (OP == OP_GELU_BWD) r = y * gelu_grad_f(x);
    else if (OP == OP_RELU_BWD) r = x > 0.f ? y : 0.f;
    else r = x + y;
    stf(out + i, r);
  }
}

template <typename TS, typename TD>
__global__ void cast_scale_kernel(const TS* __restrict__ src, TD* __restrict__ dst, long long n, float scale) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    stf(dst + i, ldf(src + i) * scale);
}
// column sums of a bf16 matrix: block = 32 columns x 8 row lanes, rows split over gridDim.y, fp32 atomics
__global__ void colsum_bf16_kernel(const bf16* __restrict__ x, float* __restrict__ out, long long rows, int cols) {
  __shared__ float sh[8][33];
  const int c = blockIdx.x * 32 + threadIdx.x;
  float s = 0.f;
  if (c < cols)
    for (long long r = static_cast<long long>(blockIdx.y) * 8 + threadIdx.y; r < rows; r += static_cast<long long>(gridDim.y) * 8)
      s += __bfloat162float(x[r * cols + c]);
  sh[threadIdx.y][threadIdx.x] = s;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) t += sh[j][threadIdx.x];
    atomicAdd(out + c, t);
  }
}
// vector flavour (cols % 8 == 0, 16-byte aligned rows): a thread owns 8 adjacent columns (one 16-byte load per row),
// block = 32 column-groups x 8 row lanes = 256 columns wide
__global__ void colsum_bf16_vec_kernel(const bf16* __restrict__ x, float* __restrict__ out, long long rows, int cols) {
  __shared__ float sh[8][32][9];
  const int cg = blockIdx.x * 32 + threadIdx.x;      // column group (8 columns)
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = 0.f;
  if (cg * 8 < cols)
    for (long long r = static_cast<long long>(blockIdx.y) * 8 + threadIdx.y; r < rows; r += static_cast<long long>(gridDim.y) * 8) {
      float f[8];
      vload(x + r * cols + cg * 8, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) s[k] += f[k];
    }
#pragma unroll
  for (int k = 0; k < 8; ++k) sh[threadIdx.y][threadIdx.x][k] = s[k];
  __syncthreads();
  // 256 threads fold the 8 row lanes of the block's 256 columns: one atomic per column
  const int t = threadIdx.y * 32 + threadIdx.x;
  const int g = t >> 3, k = t & 7;
  const int c = (blockIdx.x * 32 + g) * 8 + k;
  if (c < cols) {
    float a = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) a += sh[j][g][k];
    atomicAdd(out + c, a);
  }
}
// the same for up to four matrices with the same number of rows in one launch (the four bias gradients of a transformer
// block): blockIdx.x walks the concatenated column groups, the table says which matrix a block belongs to
struct ColsumTable { const bf16* x[4]; float* out[4]; int cols[4]; int first_block[5]; int count; };
__global__ void colsum_bf16_multi_kernel(const ColsumTable t, long long rows) {
  __shared__ float sh[8][32][9];
  // (compile-time indices only: a runtime index into the by-value table would make the compiler copy it to local memory)
  const bf16* __restrict__ x = t.x[0];
  float* out = t.out[0];
  int cols = t.cols[0], first = 0;
#pragma unroll
  for (int i = 1; i < 4; ++i)
    if (i < t.count && static_cast<int>(blockIdx.x) >= t.first_block[i]) { x = t.x[i]; out = t.out[i]; cols = t.cols[i]; first = t.first_block[i]; }
  const int bx = static_cast<int>(blockIdx.x) - first;
  const int cg = bx * 32 + threadIdx.x;              // column group (8 columns)
  float s[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) s[k] = 0.f;
  if (cg * 8 < cols) {
#pragma unroll 8
    for (long long r = static_cast<long long>(blockIdx.y) * 8 + threadIdx.y; r < rows; r += static_cast<long long>(gridDim.y) * 8) {
      float f[8];
      vload(x + r * cols + cg * 8, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) s[k] += f[k];
    }
  }
#pragma unroll
  for (int k = 0; k < 8; ++k) sh[threadIdx.y][threadIdx.x][k] = s[k];
  __syncthreads();
  const int tt = threadIdx.y * 32 + threadIdx.x;
  const int g = tt >> 3, k = tt & 7;
  const int c = (bx * 32 + g) * 8 + k;
  if (c < cols) {
    float a = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) a += sh[j][g][k];
    // explicit global reduction: through the table's runtime-indexed pointer a plain atomicAdd compiles to the generic form
    asm volatile("red.global.add.f32 [%0], %1;" ::"l"(out + c), "f"(a) : "memory");
  }
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

template <typename T>
int launch_ln_bwd_dx(int nv, int grid, cudaStream_t s, const T* x, const T* dy, const T* g, const float* mean, const float* rstd,
                     T* dx, long long rows, int C, const T* dres) {
#define LN_CASE(NVV) case NVV: layernorm_bwd_dx_kernel<T, NVV><<<grid, 256, 0, s>>>(x, dy, g, mean, rstd, dx, rows, C, dres); break;
  switch (nv) {
    LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(5) LN_CASE(6) LN_CASE(7) LN_CASE(8)
    default: return DDPX_ERR_SHAPE;
  }
#undef LN_CASE
  return DDPX_OK;
}

template <typename T>
int launch_ln_bwd(int nv, int grid, size_t smem, cudaStream_t s, const T* x, const T* dy, const T* g, const float* mean,
                  const float* rstd, T* dx, float* dgamma, float* dbeta, long long rows, int C, float* partial, const T* dres) {
#define LN_CASE(NVV) case NVV: layernorm_bwd_kernel<T, NVV><<<grid, 256, smem, s>>>(x, dy, g, mean, rstd, dx, dgamma, dbeta, rows, C, partial, dres); break;
  switch (nv) {
    LN_CASE(1) LN_CASE(2) LN_CASE(3) LN_CASE(4) LN_CASE(5) LN_CASE(6) LN_CASE(7) LN_CASE(8)
    LN_CASE(9) LN_CASE(10) LN_CASE(11) LN_CASE(12) LN_CASE(13) LN_CASE(14) LN_CASE(15) LN_CASE(16)
    default: return DDPX_ERR_SHAPE;
  }
#undef LN_CASE
  return DDPX_OK;
}

}  // namespace

#define LAUNCH_OK() (cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH)
#define ST(s) static_cast<cudaStream_t>(s)
#define EW_GRID(n, V) cap_blocks(((n) / (V) + 255) / 256)

extern "C" int ddpx_layernorm_fwd(const void* x, const void* gamma, const void* beta, void* y, float* mean, float* rstd,
                                  long long rows, int C, float eps, int is_bf16, void* stream) {
  const int V = is_bf16 ? 8 : 4;
  if (C % V) return DDPX_ERR_SHAPE;
  const int grid = cap_blocks((rows + 7) / 8, 148 * 8);
  if (is_bf16)
    layernorm_fwd_kernel<bf16><<<grid, 256, 0, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(gamma),
                                                          static_cast<const bf16*>(beta), static_cast<bf16*>(y), mean, rstd,
                                                          rows, C, eps);
  else
    layernorm_fwd_kernel<float><<<grid, 256, 0, ST(stream)>>>(static_cast<const float*>(x), static_cast<const float*>(gamma),
                                                           static_cast<const float*>(beta), static_cast<float*>(y), mean,
                                                           rstd, rows, C, eps);
  return LAUNCH_OK();
}

// dgamma / dbeta: fp32 [C].  Without scratch they are accumulators that must be zero on entry (every block reduces into
// them); with `scratch` (fp32, at least ddpx_layernorm_bwd_scratch_floats(rows, C) elements) the blocks store partial sums
// there, a second small kernel adds them up and OVERWRITES dgamma / dbeta - no zeroing needed and ~25 us less per call.
static int ln_bwd_grid(long long rows, int nv) { return cap_blocks((rows + 7) / 8, nv <= 2 ? 296 : 148); }

extern "C" long long ddpx_layernorm_bwd_scratch_floats(long long rows, int C) {
  return static_cast<long long>(ln_bwd_grid(rows, 1)) * 2 * C;          // the larger of the two grids
}

// `dres` (optional, same shape and dtype as dx, not aliasing it) is added to dx: the gradient of the residual branch around
// the LayerNorm, so that no separate add pass runs.
extern "C" int ddpx_layernorm_bwd_res(const void* x, const void* dy, const void* gamma, const float* mean, const float* rstd,
                                      void* dx, float* dgamma, float* dbeta, long long rows, int C, int is_bf16, float* scratch,
                                      const void* dres, void* stream) {
  const int V = is_bf16 ? 8 : 4;
  if (C % V) return DDPX_ERR_SHAPE;
  const int nv = (C / V + 31) / 32;
  const size_t smem = static_cast<size_t>(2) * C * sizeof(float);
  if (nv > 16 || smem > 48 * 1024) return DDPX_ERR_SHAPE;
  if ((reinterpret_cast<uintptr_t>(dgamma) | reinterpret_cast<uintptr_t>(dbeta) | reinterpret_cast<uintptr_t>(scratch)) & 15)
    return DDPX_ERR_ALIGN;
  static int split = -1;            // DDPX_LN_SPLIT=0: the single fused kernel also when scratch is given (A/B timing)
  if (split < 0) { const char* e = getenv("DDPX_LN_SPLIT"); split = (e && e[0] == '0') ? 0 : 1; }
  // worth it when the fused kernel's warps would see only a few rows each (their 2 * NV * V accumulators are set up and
  // folded per warp whatever the row count): GPT-2 XL's 2048-row calls 18.9 -> ~12 us, 144.4 instead of 146.6 ms per step;
  // at 8192 rows (seven rows per warp) the fused kernel is as fast
  if (scratch != nullptr && nv <= 8 && split && rows < 148 * 8 * 4) {
    // two kernels: dx with one row per warp and nothing carried between rows (high occupancy), then the column sums as a
    // tiled reduction over rows; the scratch rows are those of the second kernel's gridDim.y (<= 64 <= the documented size)
    const int grid_dx = cap_blocks((rows + 7) / 8, 148 * 8);
    int rc2;
    if (is_bf16)
      rc2 = launch_ln_bwd_dx<bf16>(nv, grid_dx, ST(stream), static_cast<const bf16*>(x), static_cast<const bf16*>(dy),
                                   static_cast<const bf16*>(gamma), mean, rstd, static_cast<bf16*>(dx), rows, C,
                                   static_cast<const bf16*>(dres));
    else
      rc2 = launch_ln_bwd_dx<float>(nv, grid_dx, ST(stream), static_cast<const float*>(x), static_cast<const float*>(dy),
                                    static_cast<const float*>(gamma), mean, rstd, static_cast<float*>(dx), rows, C,
                                    static_cast<const float*>(dres));
    if (rc2 != DDPX_OK) return rc2;
    if (cudaGetLastError() != cudaSuccess) return DDPX_ERR_LAUNCH;
    int gy = static_cast<int>((rows + 63) / 64);       // >= 8 rows per thread
    if (gy > 64) gy = 64;
    if (gy > ln_bwd_grid(rows, 1)) gy = ln_bwd_grid(rows, 1);      // never more rows than the scratch is sized for
    if (gy < 1) gy = 1;
    const dim3 grid_cs((C / V + 31) / 32, gy);
    if (is_bf16)
      layernorm_bwd_colsum_kernel<bf16><<<grid_cs, 256, 0, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<const bf16*>(dy),
                                                                         mean, rstd, scratch, rows, C);
    else
      layernorm_bwd_colsum_kernel<float><<<grid_cs, 256, 0, ST(stream)>>>(static_cast<const float*>(x), static_cast<const float*>(dy),
                                                                          mean, rstd, scratch, rows, C);
    if (cudaGetLastError() != cudaSuccess) return DDPX_ERR_LAUNCH;
    ln_partial_reduce_kernel<<<(2 * C + 31) / 32, 256, 0, ST(stream)>>>(scratch, gy, 2 * C, C, dgamma, dbeta);
    return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
  }
  // every warp carries register accumulators across its rows; one or two blocks per SM
  const int grid = ln_bwd_grid(rows, nv);
  int rc;
  if (is_bf16)
    rc = launch_ln_bwd<bf16>(nv, grid, smem, ST(stream), static_cast<const bf16*>(x), static_cast<const bf16*>(dy),
                             static_cast<const bf16*>(gamma), mean, rstd, static_cast<bf16*>(dx), dgamma, dbeta, rows, C, scratch,
                             static_cast<const bf16*>(dres));
  else
    rc = launch_ln_bwd<float>(nv, grid, smem, ST(stream), static_cast<const float*>(x), static_cast<const float*>(dy),
                              static_cast<const float*>(gamma), mean, rstd, static_cast<float*>(dx), dgamma, dbeta, rows, C, scratch,
                              static_cast<const float*>(dres));
  if (rc != DDPX_OK) return rc;
  if (cudaGetLastError() != cudaSuccess) return DDPX_ERR_LAUNCH;
  if (scratch != nullptr) {
    ln_partial_reduce_kernel<<<(2 * C + 31) / 32, 256, 0, ST(stream)>>>(scratch, grid, 2 * C, C, dgamma, dbeta);
    if (cudaGetLastError() != cudaSuccess) return DDPX_ERR_LAUNCH;
  }
  return DDPX_OK;
}

extern "C" int ddpx_layernorm_bwd_ws(const void* x, const void* dy, const void* gamma, const float* mean, const float* rstd,
                                     void* dx, float* dgamma, float* dbeta, long long rows, int C, int is_bf16, float* scratch,
                                     void* stream) {
  return ddpx_layernorm_bwd_res(x, dy, gamma, mean, rstd, dx, dgamma, dbeta, rows, C, is_bf16, scratch, nullptr, stream);
}

extern "C" int ddpx_layernorm_bwd(const void* x, const void* dy, const void* gamma, const float* mean, const float* rstd,
                                  void* dx, float* dgamma, float* dbeta, long long rows, int C, int is_bf16, void* stream) {
  return ddpx_layernorm_bwd_ws(x, dy, gamma, mean, rstd, dx, dgamma, dbeta, rows, C, is_bf16, nullptr, stream);
}

template <int OP>
static int launch_ew(const void* a, const void* b, void* out, long long n, int is_bf16, void* stream) {
  if (is_bf16)
    ew_kernel<bf16, OP><<<EW_GRID(n, 8), 256, 0, ST(stream)>>>(static_cast<const bf16*>(a), static_cast<const bf16*>(b),
                                                            static_cast<bf16*>(out), n);
  else
    ew_kernel<float, OP><<<EW_GRID(n, 4), 256, 0, ST(stream)>>>(static_cast<const float*>(a), static_cast<const float*>(b),
                                                             static_cast<float*>(out), n);
  return LAUNCH_OK();
}
static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int ddpx_gelu_fwd(const void* x, void* y, long long n, int is_bf16, void* stream) {
  if (!aligned16(x) || !aligned16(y)) return DDPX_ERR_ALIGN;
  return launch_ew<OP_GELU_FWD>(x, nullptr, y, n, is_bf16, stream);
}
extern "C" int ddpx_gelu_bwd(const void* x, const void* dy, void* dx, long long n, int is_bf16, void* stream) {
  if (!aligned16(x) || !aligned16(dy) || !aligned16(dx)) return DDPX_ERR_ALIGN;
  return launch_ew<OP_GELU_BWD>(x, dy, dx, n, is_bf16, stream);
}
extern "C" int ddpx_relu_bwd_bf16(const void* y, const void* dy, void* dx, long long n, void* stream) {
  if (!aligned16(y) || !aligned16(dy) || !aligned16(dx)) return DDPX_ERR_ALIGN;
  return launch_ew<OP_RELU_BWD>(y, dy, dx, n, 1, stream);
}
extern "C" int ddpx_add(const void* a, const void* b, void* out, long long n, int is_bf16, void* stream) {
  if (!aligned16(a) || !aligned16(b) || !aligned16(out)) return DDPX_ERR_ALIGN;
  return launch_ew<OP_ADD>(a, b, out, n, is_bf16, stream);
}
// dst = src * scale with a dtype change: kind 0 = f32->bf16, 1 = bf16->f32, 2 = bf16->bf16
extern "C" int ddpx_cast_scale(const void* src, void* dst, long long n, float scale, int kind, void* stream) {
  const int grid = cap_blocks((n + 255) / 256);
  if (kind == 0)
    cast_scale_kernel<float, bf16><<<grid, 256, 0, ST(stream)>>>(static_cast<const float*>(src), static_cast<bf16*>(dst), n, scale);
  else if (kind == 1)
    cast_scale_kernel<bf16, float><<<grid, 256, 0, ST(stream)>>>(static_cast<const bf16*>(src), static_cast<float*>(dst), n, scale);
  else
    cast_scale_kernel<bf16, bf16><<<grid, 256, 0, ST(stream)>>>(static_cast<const bf16*>(src), static_cast<bf16*>(dst), n, scale);
  return LAUNCH_OK();
}
// out[c] = sum_r x[r][c] for bf16 x (fp32 out; zeroed here)
extern "C" int ddpx_colsum_bf16(const void* x, float* out, long long rows, int cols, void* stream) {
  int gy = static_cast<int>(rows / 128);
  if (gy < 1) gy = 1;
  if (gy > 64) gy = 64;
  cudaMemsetAsync(out, 0, sizeof(float) * cols, ST(stream));
  dim3 block(32, 8);
  if (cols % 8 == 0 && (reinterpret_cast<uintptr_t>(x) & 15) == 0) {
    dim3 grid((cols / 8 + 31) / 32, gy);
    colsum_bf16_vec_kernel<<<grid, block, 0, ST(stream)>>>(static_cast<const bf16*>(x), out, rows, cols);
    return LAUNCH_OK();
  }
  dim3 grid((cols + 31) / 32, gy);
  colsum_bf16_kernel<<<grid, block, 0, ST(stream)>>>(static_cast<const bf16*>(x), out, rows, cols);
  return LAUNCH_OK();
}
// out[i][c] += sum_r x[i][r][c] for n <= 4 bf16 matrices [rows][cols[i]] (cols[i] % 8 == 0, 16-byte aligned) in ONE launch;
// the outputs are fp32 accumulators the caller has zeroed (they may be slices of one buffer cleared with one memset)
extern "C" int ddpx_colsum_bf16_multi(int n, const void* const* x, float* const* out, long long rows, const int* cols,
                                      void* stream) {
  if (n < 1 || n > 4) return DDPX_ERR_ARG;
  ColsumTable t{};
  t.count = n;
  int blocks = 0;
  for (int i = 0; i < n; ++i) {
    if (cols[i] % 8 || (reinterpret_cast<uintptr_t>(x[i]) & 15)) return DDPX_ERR_ALIGN;
    t.x[i] = static_cast<const bf16*>(x[i]); t.out[i] = out[i]; t.cols[i] = cols[i];
    t.first_block[i] = blocks;
    blocks += (cols[i] / 8 + 31) / 32;
  }
  t.first_block[n] = blocks;
  int gy = static_cast<int>(rows / 128);
  if (gy < 1) gy = 1;
  if (gy > 64) gy = 64;
  colsum_bf16_multi_kernel<<<dim3(blocks, gy), dim3(32, 8), 0, ST(stream)>>>(t, rows);
  return LAUNCH_OK();
}
extern "C" int ddpx_dropout(const void* x, void* y, long long n, float p, unsigned long long seed, int is_bf16, void* stream) {
  const float scale = 1.f / (1.f - p);
  const int grid = cap_blocks((n + 255) / 256);
  if (is_bf16)
    dropout_kernel<bf16><<<grid, 256, 0, ST(stream)>>>(static_cast<const bf16*>(x), static_cast<bf16*>(y), n, p, scale, seed);
  else
    dropout_kernel<float><<<grid, 256, 0, ST(stream)>>>(static_cast<const float*>(x), static_cast<float*>(y), n, p, scale, seed);
  return LAUNCH_OK();
}
