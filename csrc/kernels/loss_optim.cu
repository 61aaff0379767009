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
    for (long long i = base + threadIdx.x; i < end; i += blockDim.x) {
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
    for (long long i = base + threadIdx.x; i < end; i += blockDim.x) {
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
