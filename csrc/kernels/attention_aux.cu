// Attention / embedding auxiliaries for GPT-2 (BASELINE.json configs 3-4), sm_100a.
//   * token + position embedding gather (forward) and its scatter-add / batch-sum backward
//   * causal softmax forward (in place on the score matrix written by the batched tcgen05 GEMM) and backward
// Attention itself is  S = scale * Q K^T (tcgen05 batched GEMM, tiles above the diagonal skipped)  ->  causal softmax
// (this file)  ->  O = P V (tcgen05 batched GEMM, K loop cut at the diagonal); the backward is four more batched GEMMs
// around softmax_bwd.  No counterpart in the reference (a CNN); bar = torch SDPA.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "vec.cuh"
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

// out[b,t,:] = wte[idx[b,t],:] + wpe[t,:]
template <typename T>
__global__ void embedding_fwd_kernel(const long long* __restrict__ idx, const T* __restrict__ wte, const T* __restrict__ wpe,
                                     T* __restrict__ out, long long BT, int T_, int C) {
  for (long long r = blockIdx.x; r < BT; r += gridDim.x) {
    const long long tok = idx[r];
    const int t = static_cast<int>(r % T_);
    const T* a = wte + tok * C;
    const T* b = wpe + static_cast<long long>(t) * C;
    T* o = out + r * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) stf(o + c, ldf(a + c) + ldf(b + c));
  }
}
// dwte[idx[b,t],:] += dy[b,t,:] (fp32 accumulators, atomics);  dwpe[t,:] += sum_b dy[b,t,:]
template <typename T>
__global__ void embedding_bwd_kernel(const long long* __restrict__ idx, const T* __restrict__ dy, float* __restrict__ dwte,
                                     float* __restrict__ dwpe, long long BT, int T_, int C) {
  for (long long r = blockIdx.x; r < BT; r += gridDim.x) {
    const long long tok = idx[r];
    const int t = static_cast<int>(r % T_);
    const T* d = dy + r * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      const float v = ldf(d + c);
      atomicAdd(dwte + tok * C + c, v);
      atomicAdd(dwpe + static_cast<long long>(t) * C + c, v);
    }
  }
}

// Causal softmax in place: row (bh, i) of S[BH][T][T] keeps columns j <= i, P = softmax over them, zeros elsewhere.
// One warp per row, 16-byte vector accesses (T must be a multiple of the vector width); the row is read once, kept in
// registers (NV vectors per lane, T <= 32*NV*V), and written once.
template <typename T, int NV>
__global__ void __launch_bounds__(256) causal_softmax_fwd_kernel(T* __restrict__ s, long long rows, int T_) {
  constexpr int V = ddpx::VecOf<T>::N;
  const int lane = threadIdx.x & 31;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const int nvec = T_ / V;
  for (long long r = warp0; r < rows; r += nwarps) {
    const int i = static_cast<int>(r % T_);
    T* row = s + r * T_;
    float f[NV][V];
    float mx = -INFINITY;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vi = lane + j * 32;
      if (vi < nvec && vi * V <= i) {
        ddpx::vload(row + vi * V, f[j]);
#pragma unroll
        for (int k = 0; k < V; ++k) {
          if (vi * V + k > i) f[j][k] = -INFINITY;
          mx = fmaxf(mx, f[j][k]);
        }
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) f[j][k] = -INFINITY;
      }
    }
    mx = ddpx::warp_max_f(mx);
    float se = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j)
#pragma unroll
      for (int k = 0; k < V; ++k) { f[j][k] = __expf(f[j][k] - mx); se += f[j][k]; }
    se = ddpx::warp_sum_f(se);
    const float inv = 1.f / se;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vi = lane + j * 32;
      if (vi < nvec) {
#pragma unroll
        for (int k = 0; k < V; ++k) f[j][k] *= inv;
        ddpx::vstore(row + vi * V, f[j]);
      }
    }
  }
}
// dS = P * (dP - sum_j dP_j P_j) * scale, in place on dP; zeros above the diagonal
template <typename T, int NV>
__global__ void __launch_bounds__(256)
causal_softmax_bwd_kernel(const T* __restrict__ p, T* __restrict__ dp, long long rows, int T_, float scale) {
  constexpr int V = ddpx::VecOf<T>::N;
  const int lane = threadIdx.x & 31;
  const long long warp0 = (static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = (static_cast<long long>(gridDim.x) * blockDim.x) >> 5;
  const int nvec = T_ / V;
  for (long long r = warp0; r < rows; r += nwarps) {
    const int i = static_cast<int>(r % T_);
    const T* pr = p + r * T_;
    T* dr = dp + r * T_;
    float fp[NV][V], fd[NV][V];
    float dot = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vi = lane + j * 32;
      if (vi < nvec && vi * V <= i) {
        ddpx::vload(pr + vi * V, fp[j]);
        ddpx::vload(dr + vi * V, fd[j]);
#pragma unroll
        for (int k = 0; k < V; ++k) {
          if (vi * V + k > i) { fp[j][k] = 0.f; fd[j][k] = 0.f; }
          dot += fp[j][k] * fd[j][k];
        }
      } else {
#pragma unroll
        for (int k = 0; k < V; ++k) { fp[j][k] = 0.f; fd[j][k] = 0.f; }
      }
    }
    dot = ddpx::warp_sum_f(dot);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vi = lane + j * 32;
      if (vi < nvec) {
#pragma unroll
        for (int k = 0; k < V; ++k) fd[j][k] = fp[j][k] * (fd[j][k] - dot) * scale;
        ddpx::vstore(dr + vi * V, fd[j]);
      }
    }
  }
}

inline int cap_blocks(long long n, int cap = 148 * 16) { return static_cast<int>(n < 1 ? 1 : (n > cap ? cap : n)); }

}  // namespace

#define LAUNCH_OK() (cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH)
#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int ddpx_embedding_fwd(const long long* idx, const void* wte, const void* wpe, void* out, long long BT, int T_,
                                  int C, int is_bf16, void* stream) {
  const int grid = cap_blocks(BT, 148 * 16);
  if (is_bf16)
    embedding_fwd_kernel<bf16><<<grid, 128, 0, ST(stream)>>>(idx, static_cast<const bf16*>(wte), static_cast<const bf16*>(wpe),
                                                          static_cast<bf16*>(out), BT, T_, C);
  else
    embedding_fwd_kernel<float><<<grid, 128, 0, ST(stream)>>>(idx, static_cast<const float*>(wte), static_cast<const float*>(wpe),
                                                           static_cast<float*>(out), BT, T_, C);
  return LAUNCH_OK();
}
// dwte [V][C] and dwpe [n_pos][C] are fp32 accumulators (zeroed by the caller or holding a gradient to add to)
extern "C" int ddpx_embedding_bwd(const long long* idx, const void* dy, float* dwte, float* dwpe, long long BT, int T_, int C,
                                  int is_bf16, void* stream) {
  const int grid = cap_blocks(BT, 148 * 16);
  if (is_bf16)
    embedding_bwd_kernel<bf16><<<grid, 128, 0, ST(stream)>>>(idx, static_cast<const bf16*>(dy), dwte, dwpe, BT, T_, C);
  else
    embedding_bwd_kernel<float><<<grid, 128, 0, ST(stream)>>>(idx, static_cast<const float*>(dy), dwte, dwpe, BT, T_, C);
  return LAUNCH_OK();
}
template <typename T>
static int launch_softmax(bool bwd, const T* p, T* s, long long rows, int T_, float scale, cudaStream_t st) {
  constexpr int V = ddpx::VecOf<T>::N;
  if (T_ % V) return DDPX_ERR_SHAPE;
  const int nv = (T_ / V + 31) / 32;
  const int grid = cap_blocks((rows + 7) / 8, 148 * 16);
#define SM_CASE(NVV)                                                                                   \
  case NVV:                                                                                            \
    if (bwd) causal_softmax_bwd_kernel<T, NVV><<<grid, 256, 0, st>>>(p, s, rows, T_, scale);           \
    else causal_softmax_fwd_kernel<T, NVV><<<grid, 256, 0, st>>>(s, rows, T_);                         \
    break;
  switch (nv) {
    SM_CASE(1) SM_CASE(2) SM_CASE(3) SM_CASE(4) SM_CASE(5) SM_CASE(6) SM_CASE(7) SM_CASE(8)
    default: return DDPX_ERR_SHAPE;     // T <= 2048 (bf16) / 1024 (fp32)
  }
#undef SM_CASE
  return LAUNCH_OK();
}
extern "C" int ddpx_causal_softmax_fwd(void* s, long long rows, int T_, int is_bf16, void* stream) {
  if (is_bf16) return launch_softmax<bf16>(false, nullptr, static_cast<bf16*>(s), rows, T_, 1.f, ST(stream));
  return launch_softmax<float>(false, nullptr, static_cast<float*>(s), rows, T_, 1.f, ST(stream));
}
extern "C" int ddpx_causal_softmax_bwd(const void* p, void* dp, long long rows, int T_, float scale, int is_bf16, void* stream) {
  if (is_bf16) return launch_softmax<bf16>(true, static_cast<const bf16*>(p), static_cast<bf16*>(dp), rows, T_, scale, ST(stream));
  return launch_softmax<float>(true, static_cast<const float*>(p), static_cast<float*>(dp), rows, T_, scale, ST(stream));
}
