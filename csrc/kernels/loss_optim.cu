// Loss and optimizer kernels, sm_100a.
//   * softmax cross-entropy, forward + backward in one kernel (loss mean over rows and dlogits = (softmax-onehot)/rows)
//     - replaces F.cross_entropy at /root/reference/multigpu.py:95; fp32 or bf16 logits
//   * MSE loss forward + backward (upstream-tutorial loss, BASELINE.json config #1; W18)
//   * multi-tensor SGD(momentum, weight decay, nesterov) and AdamW in ONE launch for any number of tensors
//     - replaces torch's _foreach SGD (torch/optim/sgd.py:346-380) behind optimizer.step() at multigpu.py:97 (K14, W23)
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string.h>
#include "vec.cuh"
#include "../../include/ddpx_kernels.h"

namespace {

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float block_reduce(float v, float* sh, bool is_max) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = is_max ? warp_max(v) : warp_sum(v);
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  float r = (threadIdx.x < nw) ? sh[threadIdx.x] : (is_max ? -INFINITY : 0.f);
  if (w == 0) {
    r = is_max ? warp_max(r) : warp_sum(r);
    if (lane == 0) sh[0] = r;
  }
  __syncthreads();
  return sh[0];
}

template <typename T> __device__ __forceinline__ float to_f(T v);
template <> __device__ __forceinline__ float to_f<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

// one block per row; targets int64; ignore_index rows contribute nothing.  loss_sum accumulates sum(-log p)/rows_valid.
template <typename T>
__global__ void softmax_ce_kernel(const T* __restrict__ logits, const long long* __restrict__ targets, T* __restrict__ dlogits,
                                  float* __restrict__ loss_sum, long long rows, int cols, long long ld, float inv_rows,
                                  long long ignore_index, float grad_scale) {
  __shared__ float sh[33];
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* x = logits + row * ld;
    const long long tgt = targets[row];
    float mx = -INFINITY;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) mx = fmaxf(mx, to_f(x[c]));
    mx = block_reduce(mx, sh, true);
    float se = 0.f;
    for (int c = threadIdx.x; c < cols; c += blockDim.x) se += __expf(to_f(x[c]) - mx);
    se = block_reduce(se, sh, false);
    const bool valid = (tgt != ignore_index) && tgt >= 0 && tgt < cols;
    if (threadIdx.x == 0 && valid && loss_sum != nullptr) {
      const float lp = to_f(x[tgt]) - mx - __logf(se);
      atomicAdd(loss_sum, -lp * inv_rows);
    }
    if (dlogits != nullptr) {
      T* d = dlogits + row * ld;
      const float inv_se = 1.f / se;
      const float sc = valid ? inv_rows * grad_scale : 0.f;
      for (int c = threadIdx.x; c < cols; c += blockDim.x) {
        float pr = __expf(to_f(x[c]) - mx) * inv_se;
        if (c == tgt) pr -= 1.f;
        d[c] = from_f<T>(pr * sc);
      }
    }
    __syncthreads();
  }
}

// ---- two-kernel form used by the autograd ops: the forward reads the logits ONCE (online max / sum-of-exp, 16-byte
// loads) and keeps one log-sum-exp per row; the backward reads them once more and writes
// dlogits = (softmax - onehot) * g / rows with the upstream scalar gradient g taken from device memory (no host sync,
// no separate scaling pass).  Columns [cols, ld) (vocabulary padding) get a zero gradient.
template <typename T>
__global__ void __launch_bounds__(512)
ce_fwd_kernel(const T* __restrict__ logits, const long long* __restrict__ targets, float* __restrict__ lse,
              float* __restrict__ loss_sum, long long rows, int cols, long long ld, float inv_rows, long long ignore_index,
              int vec_ok) {
  constexpr int V = sizeof(T) == 2 ? 8 : 4;
  __shared__ float sh[66];
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* x = logits + row * ld;
    float mx = -INFINITY, se = 0.f;
    const int nvec = vec_ok ? cols / V : 0;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      const uint4 u = *reinterpret_cast<const uint4*>(x + i * V);
      float f[V];
      if (sizeof(T) == 2) {
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) { f[2 * k] = __uint_as_float(w[k] << 16); f[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u); }
      } else {
        f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
      }
      float m2 = mx;
#pragma unroll
      for (int k = 0; k < V; ++k) m2 = fmaxf(m2, f[k]);
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < V; ++k) acc += __expf(f[k] - m2);
      se = se * __expf(mx - m2) + acc;
      mx = m2;
    }
    for (int c = nvec * V + threadIdx.x; c < cols; c += blockDim.x) {
      const float v = to_f(x[c]);
      const float m2 = fmaxf(mx, v);
      se = se * __expf(mx - m2) + __expf(v - m2);
      mx = m2;
    }
    // combine (max, sum) pairs across the block
    const float bm = block_reduce(mx, sh, true);
    const float contrib = (mx == -INFINITY) ? 0.f : se * __expf(mx - bm);
    const float bs = block_reduce(contrib, sh + 33, false);
    if (threadIdx.x == 0) {
      const float l = bm + __logf(bs);
      lse[row] = l;
      const long long tgt = targets[row];
      if (tgt != ignore_index && tgt >= 0 && tgt < cols && loss_sum != nullptr) atomicAdd(loss_sum, (l - to_f(x[tgt])) * inv_rows);
    }
    __syncthreads();
  }
}

template <typename T>
__global__ void __launch_bounds__(512)
ce_bwd_kernel(const T* __restrict__ logits, const long long* __restrict__ targets, const float* __restrict__ lse,
              const float* __restrict__ g, T* __restrict__ dlogits, long long rows, int cols, long long ld, float inv_rows,
              long long ignore_index, int vec_ok) {
  constexpr int V = sizeof(T) == 2 ? 8 : 4;
  const float gs = (g != nullptr ? *g : 1.f) * inv_rows;
  for (long long row = blockIdx.x; row < rows; row += gridDim.x) {
    const T* x = logits + row * ld;
    T* d = dlogits + row * ld;
    const long long tgt = targets[row];
    const bool valid = (tgt != ignore_index) && tgt >= 0 && tgt < cols;
    const float sc = valid ? gs : 0.f;
    const float l = lse[row];
    const int nvec = vec_ok ? cols / V : 0;
    for (int i = threadIdx.x; i < nvec; i += blockDim.x) {
      const uint4 u = *reinterpret_cast<const uint4*>(x + i * V);
      float f[V];
      if (sizeof(T) == 2) {
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int k = 0; k < 4; ++k) { f[2 * k] = __uint_as_float(w[k] << 16); f[2 * k + 1] = __uint_as_float(w[k] & 0xffff0000u); }
      } else {
        f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
      }
#pragma unroll
      for (int k = 0; k < V; ++k) {
        float pr = __expf(f[k] - l);
        if (i * V + k == tgt) pr -= 1.f;
        f[k] = pr * sc;
      }
      uint4 o;
      if (sizeof(T) == 2) {
        uint32_t w[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) { __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * k], f[2 * k + 1]); w[k] = *reinterpret_cast<uint32_t*>(&h); }
        o = make_uint4(w[0], w[1], w[2], w[3]);
      } else {
        o = make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
      }
      *reinterpret_cast<uint4*>(d + i * V) = o;
    }
    for (int c = nvec * V + threadIdx.x; c < ld; c += blockDim.x) {
      float pr = 0.f;
      if (c < cols) {
        pr = __expf(to_f(x[c]) - l);
        if (c == tgt) pr -= 1.f;
      }
      d[c] = from_f<T>(pr * sc);
    }
  }
}

// dlogits *= g (chain rule when the upstream gradient of the scalar loss is not 1)
template <typename T>
__global__ void scale_inplace_kernel(T* __restrict__ p, long long n, const float* __restrict__ g) {
  const float s = *g;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    p[i] = from_f<T>(to_f(p[i]) * s);
}

__global__ void mse_kernel(const float* __restrict__ pred, const float* __restrict__ target, float* __restrict__ dpred,
                           float* __restrict__ loss_sum, long long n, float inv_n) {
  __shared__ float sh[33];
  float acc = 0.f;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const float d = pred[i] - target[i];
    acc += d * d;
    if (dpred != nullptr) dpred[i] = 2.f * d * inv_n;
  }
  acc = block_reduce(acc, sh, false);
  if (threadIdx.x == 0 && loss_sum != nullptr) atomicAdd(loss_sum, acc * inv_n);
}

// ------------------------------------------------------------------------------------------------
// multi-tensor optimizers: tensor table passed by value in the kernel parameters (no device-side table, no sync)
// ------------------------------------------------------------------------------------------------
constexpr int kMaxTensors = 96;
constexpr int kChunk = 4096;  // elements per block-iteration

struct TensorTable {
  void* p[kMaxTensors];
  const void* g[kMaxTensors];
  void* s0[kMaxTensors];       // SGD: momentum buffer; AdamW: exp_avg
  void* s1[kMaxTensors];       // AdamW: exp_avg_sq
  int chunk_start[kMaxTensors + 1];   // prefix sum of chunk counts
  long long numel[kMaxTensors];
  int n;
};

__device__ __forceinline__ int find_tensor(const TensorTable& t, int chunk) {
  int lo = 0, hi = t.n - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (t.chunk_start[mid] <= chunk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// torch SGD semantics (sgd.py:346-380): g += wd*p; first step buf = g else buf = mu*buf + (1-dampening)*g;
// nesterov: g = g + mu*buf else g = buf; p -= lr*g.   State is updated even when lr == 0 (LambdaLR's first step).
template <typename T>
__global__ void sgd_multi_kernel(const __grid_constant__ TensorTable t, float lr, float momentum, float dampening, float wd,
                                 int nesterov, int first_step, float grad_scale) {
  const int total_chunks = t.chunk_start[t.n];
  for (int chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const int ti = find_tensor(t, chunk);
    const long long base = static_cast<long long>(chunk - t.chunk_start[ti]) * kChunk;
    const long long end = min(base + kChunk, t.numel[ti]);
    T* __restrict__ p = static_cast<T*>(t.p[ti]);
    const T* __restrict__ g = static_cast<const T*>(t.g[ti]);
    T* __restrict__ buf = static_cast<T*>(t.s0[ti]);
    constexpr int V = ddpx::VecOf<T>::N;
    const bool al = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(buf)) & 15) == 0;
    const long long nvec = al ? (end - base) / V : 0;
    for (long long vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
      const long long i = base + vi * V;
      float pv[V], gv[V], bv[V];
      ddpx::vload(p + i, pv); ddpx::vload(g + i, gv);
      if (momentum != 0.f && !first_step) ddpx::vload(buf + i, bv);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        float gg = gv[k] * grad_scale + wd * pv[k];
        if (momentum != 0.f) {
          const float b = first_step ? gg : momentum * bv[k] + (1.f - dampening) * gg;
          bv[k] = b;
          gg = nesterov ? gg + momentum * b : b;
        }
        pv[k] -= lr * gg;
      }
      if (momentum != 0.f) ddpx::vstore(buf + i, bv);
      ddpx::vstore(p + i, pv);
    }
    for (long long i = base + nvec * V + threadIdx.x; i < end; i += blockDim.x) {
      float pv = to_f(p[i]);
      float gv = to_f(g[i]) * grad_scale + wd * pv;
      if (momentum != 0.f) {
        float b = first_step ? gv : momentum * to_f(buf[i]) + (1.f - dampening) * gv;
        buf[i] = from_f<T>(b);
        gv = nesterov ? gv + momentum * b : b;
      }
      p[i] = from_f<T>(pv - lr * gv);
    }
  }
}

// dst[i] = src[i] * scale for a table of tensors in ONE launch (gradient reducer: every gradient of a bucket goes into
// its slot of the flat communication buffer, pre-divided by the world size).  t.p = destinations, t.g = sources.
template <typename T>
__global__ void scale_copy_multi_kernel(const __grid_constant__ TensorTable t, float scale) {
  const int total_chunks = t.chunk_start[t.n];
  for (int chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const int ti = find_tensor(t, chunk);
    const long long base = static_cast<long long>(chunk - t.chunk_start[ti]) * kChunk;
    const long long end = min(base + kChunk, t.numel[ti]);
    T* __restrict__ d = static_cast<T*>(t.p[ti]);
    const T* __restrict__ g = static_cast<const T*>(t.g[ti]);
    constexpr int V = ddpx::VecOf<T>::N;
    const bool al = ((reinterpret_cast<uintptr_t>(d) | reinterpret_cast<uintptr_t>(g)) & 15) == 0;
    const long long nvec = al ? (end - base) / V : 0;
    for (long long vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
      const long long i = base + vi * V;
      float v[V];
      ddpx::vload(g + i, v);
#pragma unroll
      for (int k = 0; k < V; ++k) v[k] *= scale;
      ddpx::vstore(d + i, v);
    }
    for (long long i = base + nvec * V + threadIdx.x; i < end; i += blockDim.x) d[i] = from_f<T>(to_f(g[i]) * scale);
  }
}

// torch AdamW semantics: p *= 1 - lr*wd; m = b1*m + (1-b1)*g; v = b2*v + (1-b2)*g^2;
// p -= lr/bc1 * m / (sqrt(v)/sqrt(bc2) + eps)
template <typename T>
__global__ void adamw_multi_kernel(const __grid_constant__ TensorTable t, float lr, float beta1, float beta2, float eps, float wd,
                                   float bc1, float bc2_sqrt, float grad_scale) {
  const int total_chunks = t.chunk_start[t.n];
  for (int chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const int ti = find_tensor(t, chunk);
    const long long base = static_cast<long long>(chunk - t.chunk_start[ti]) * kChunk;
    const long long end = min(base + kChunk, t.numel[ti]);
    T* __restrict__ p = static_cast<T*>(t.p[ti]);
    const T* __restrict__ g = static_cast<const T*>(t.g[ti]);
    T* __restrict__ m = static_cast<T*>(t.s0[ti]);
    T* __restrict__ v = static_cast<T*>(t.s1[ti]);
    constexpr int V = ddpx::VecOf<T>::N;
    const bool al = ((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
                      reinterpret_cast<uintptr_t>(v)) & 15) == 0;
    const long long nvec = al ? (end - base) / V : 0;
    for (long long vi = threadIdx.x; vi < nvec; vi += blockDim.x) {
      const long long i = base + vi * V;
      float pv[V], gv[V], mv[V], vv[V];
      ddpx::vload(p + i, pv); ddpx::vload(g + i, gv); ddpx::vload(m + i, mv); ddpx::vload(v + i, vv);
#pragma unroll
      for (int k = 0; k < V; ++k) {
        const float gg = gv[k] * grad_scale;
        pv[k] *= (1.f - lr * wd);
        mv[k] = beta1 * mv[k] + (1.f - beta1) * gg;
        vv[k] = beta2 * vv[k] + (1.f - beta2) * gg * gg;
        pv[k] -= (lr / bc1) * mv[k] / (sqrtf(vv[k]) / bc2_sqrt + eps);
      }
      ddpx::vstore(m + i, mv); ddpx::vstore(v + i, vv); ddpx::vstore(p + i, pv);
    }
    for (long long i = base + nvec * V + threadIdx.x; i < end; i += blockDim.x) {
      const float gv = to_f(g[i]) * grad_scale;
      float pv = to_f(p[i]) * (1.f - lr * wd);
      const float mv = beta1 * to_f(m[i]) + (1.f - beta1) * gv;
      const float vv = beta2 * to_f(v[i]) + (1.f - beta2) * gv * gv;
      m[i] = from_f<T>(mv);
      v[i] = from_f<T>(vv);
      pv -= (lr / bc1) * mv / (sqrtf(vv) / bc2_sqrt + eps);
      p[i] = from_f<T>(pv);
    }
  }
}

int fill_table(TensorTable& t, int n, void* const* params, void* const* grads, void* const* s0, void* const* s1,
               const long long* numel) {
  if (n > kMaxTensors) return DDPX_ERR_LIMIT;
  memset(&t, 0, sizeof(t));
  t.n = n;
  int chunks = 0;
  for (int i = 0; i < n; ++i) {
    t.p[i] = params[i];
    t.g[i] = grads[i];
    t.s0[i] = s0 ? s0[i] : nullptr;
    t.s1[i] = s1 ? s1[i] : nullptr;
    t.numel[i] = numel[i];
    t.chunk_start[i] = chunks;
    chunks += static_cast<int>((numel[i] + kChunk - 1) / kChunk);
  }
  t.chunk_start[n] = chunks;
  return DDPX_OK;
}

}  // namespace

#define LAUNCH_OK() (cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH)

extern "C" int ddpx_softmax_ce(const void* logits, const long long* targets, void* dlogits, float* loss_sum, long long rows,
                               int cols, long long ld, int bf16, long long ignore_index, float grad_scale, void* stream) {
  if (rows <= 0 || cols <= 0) return DDPX_ERR_ARG;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (loss_sum != nullptr) cudaMemsetAsync(loss_sum, 0, sizeof(float), s);
  const int block = cols >= 4096 ? 512 : (cols >= 256 ? 256 : 64);
  const int grid = rows < 148 * 16 ? static_cast<int>(rows) : 148 * 16;
  const float inv_rows = 1.f / static_cast<float>(rows);
  if (bf16)
    softmax_ce_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(static_cast<const __nv_bfloat16*>(logits), targets,
                                                           static_cast<__nv_bfloat16*>(dlogits), loss_sum, rows, cols, ld,
                                                           inv_rows, ignore_index, grad_scale);
  else
    softmax_ce_kernel<float><<<grid, block, 0, s>>>(static_cast<const float*>(logits), targets, static_cast<float*>(dlogits),
                                                   loss_sum, rows, cols, ld, inv_rows, ignore_index, grad_scale);
  return LAUNCH_OK();
}

extern "C" int ddpx_ce_fwd(const void* logits, const long long* targets, float* lse, float* loss_sum, long long rows,
                           int cols, long long ld, int bf16, long long ignore_index, void* stream) {
  if (rows <= 0 || cols <= 0) return DDPX_ERR_ARG;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (loss_sum != nullptr) cudaMemsetAsync(loss_sum, 0, sizeof(float), s);
  const int es = bf16 ? 2 : 4;
  const int vec_ok = ((ld * es) % 16 == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0) ? 1 : 0;
  const int block = cols >= 8192 ? 512 : (cols >= 512 ? 256 : 64);
  const int grid = rows < 148 * 16 ? static_cast<int>(rows) : 148 * 16;
  const float inv_rows = 1.f / static_cast<float>(rows);
  if (bf16)
    ce_fwd_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(static_cast<const __nv_bfloat16*>(logits), targets, lse, loss_sum, rows,
                                                       cols, ld, inv_rows, ignore_index, vec_ok);
  else
    ce_fwd_kernel<float><<<grid, block, 0, s>>>(static_cast<const float*>(logits), targets, lse, loss_sum, rows, cols, ld,
                                               inv_rows, ignore_index, vec_ok);
  return LAUNCH_OK();
}

extern "C" int ddpx_ce_bwd(const void* logits, const long long* targets, const float* lse, const float* g, void* dlogits,
                           long long rows, int cols, long long ld, int bf16, long long ignore_index, void* stream) {
  if (rows <= 0 || cols <= 0) return DDPX_ERR_ARG;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int es = bf16 ? 2 : 4;
  const int vec_ok = ((ld * es) % 16 == 0 && (reinterpret_cast<uintptr_t>(logits) & 15) == 0 &&
                      (reinterpret_cast<uintptr_t>(dlogits) & 15) == 0) ? 1 : 0;
  const int block = cols >= 8192 ? 512 : (cols >= 512 ? 256 : 64);
  const int grid = rows < 148 * 16 ? static_cast<int>(rows) : 148 * 16;
  const float inv_rows = 1.f / static_cast<float>(rows);
  if (bf16)
    ce_bwd_kernel<__nv_bfloat16><<<grid, block, 0, s>>>(static_cast<const __nv_bfloat16*>(logits), targets, lse, g,
                                                       static_cast<__nv_bfloat16*>(dlogits), rows, cols, ld, inv_rows,
                                                       ignore_index, vec_ok);
  else
    ce_bwd_kernel<float><<<grid, block, 0, s>>>(static_cast<const float*>(logits), targets, lse, g, static_cast<float*>(dlogits),
                                               rows, cols, ld, inv_rows, ignore_index, vec_ok);
  return LAUNCH_OK();
}

extern "C" int ddpx_scale_by_device_scalar(void* p, long long n, const float* g, int bf16, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  long long b = (n + 255) / 256;
  if (b > 148 * 16) b = 148 * 16;
  if (bf16) scale_inplace_kernel<__nv_bfloat16><<<static_cast<int>(b), 256, 0, s>>>(static_cast<__nv_bfloat16*>(p), n, g);
  else scale_inplace_kernel<float><<<static_cast<int>(b), 256, 0, s>>>(static_cast<float*>(p), n, g);
  return LAUNCH_OK();
}

extern "C" int ddpx_mse(const float* pred, const float* target, float* dpred, float* loss_sum, long long n, void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (loss_sum != nullptr) cudaMemsetAsync(loss_sum, 0, sizeof(float), s);
  long long b = (n + 255) / 256;
  if (b > 148 * 4) b = 148 * 4;
  mse_kernel<<<static_cast<int>(b), 256, 0, s>>>(pred, target, dpred, loss_sum, n, 1.f / static_cast<float>(n));
  return LAUNCH_OK();
}

extern "C" int ddpx_sgd_multi(int n, void* const* params, void* const* grads, void* const* momentum_bufs,
                              const long long* numel, float lr, float momentum, float dampening, float weight_decay,
                              int nesterov, int first_step, float grad_scale, int is_bf16, void* stream) {
  for (int off = 0; off < n; off += kMaxTensors) {
    const int m = (n - off) < kMaxTensors ? (n - off) : kMaxTensors;
    TensorTable t;
    int rc = fill_table(t, m, params + off, grads + off, momentum_bufs ? momentum_bufs + off : nullptr, nullptr, numel + off);
    if (rc != DDPX_OK) return rc;
    int grid = t.chunk_start[m] < 148 * 8 ? t.chunk_start[m] : 148 * 8;
    if (grid < 1) continue;
    if (is_bf16)
      sgd_multi_kernel<__nv_bfloat16><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(t, lr, momentum, dampening, weight_decay,
                                                                                         nesterov, first_step, grad_scale);
    else
      sgd_multi_kernel<float><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(t, lr, momentum, dampening, weight_decay,
                                                                                 nesterov, first_step, grad_scale);
  }
  return LAUNCH_OK();
}

extern "C" int ddpx_adamw_multi(int n, void* const* params, void* const* grads, void* const* exp_avg, void* const* exp_avg_sq,
                                const long long* numel, float lr, float beta1, float beta2, float eps, float weight_decay,
                                long long step, float grad_scale, int is_bf16, void* stream) {
  const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
  const float bc2_sqrt = sqrtf(1.f - powf(beta2, static_cast<float>(step)));
  for (int off = 0; off < n; off += kMaxTensors) {
    const int m = (n - off) < kMaxTensors ? (n - off) : kMaxTensors;
    TensorTable t;
    int rc = fill_table(t, m, params + off, grads + off, exp_avg + off, exp_avg_sq + off, numel + off);
    if (rc != DDPX_OK) return rc;
    int grid = t.chunk_start[m] < 148 * 8 ? t.chunk_start[m] : 148 * 8;
    if (grid < 1) continue;
    if (is_bf16)
      adamw_multi_kernel<__nv_bfloat16><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(t, lr, beta1, beta2, eps, weight_decay,
                                                                                           bc1, bc2_sqrt, grad_scale);
    else
      adamw_multi_kernel<float><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(t, lr, beta1, beta2, eps, weight_decay, bc1,
                                                                                   bc2_sqrt, grad_scale);
  }
  return LAUNCH_OK();
}

extern "C" int ddpx_scale_copy_multi(int n, void* const* dst, void* const* src, const long long* numel, float scale, int is_bf16,
                                     void* stream) {
  for (int off = 0; off < n; off += kMaxTensors) {
    const int m = (n - off) < kMaxTensors ? (n - off) : kMaxTensors;
    TensorTable t;
    int rc = fill_table(t, m, dst + off, src + off, nullptr, nullptr, numel + off);
    if (rc != DDPX_OK) return rc;
    int grid = t.chunk_start[m] < 148 * 8 ? t.chunk_start[m] : 148 * 8;
    if (grid < 1) continue;
    if (is_bf16) scale_copy_multi_kernel<__nv_bfloat16><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(t, scale);
    else scale_copy_multi_kernel<float><<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(t, scale);
  }
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}
