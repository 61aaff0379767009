// BatchNorm2d (+ReLU, +2x2 max-pool) forward / backward on NHWC fp32 activations, sm_100a.
//
// Reference parity: nn.BatchNorm2d + nn.ReLU(True) + nn.MaxPool2d(2) at /root/reference/multigpu.py:54-59
// (SURVEY.md K8-K10, W21-W22).  Design: the batch statistics arrive as per-channel sum / sum-of-squares that the
// convolution epilogue already accumulated (ddpx_conv3x3_fprop), so the forward is ONE pass (normalise + affine +
// ReLU + optional pool) and the backward is two passes that recompute the ReLU mask and the pool arg-max from the
// saved pre-BN tensor instead of storing them.
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/ddpx_kernels.h"

namespace {

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }

// ------------------------------------------------------------------------------------------------
// finalize: sum / sumsq -> mean, rstd; running-stat update (momentum, unbiased var); num_batches_tracked += 1
// ------------------------------------------------------------------------------------------------
__global__ void bn_finalize_kernel(const float* __restrict__ sum, const float* __restrict__ sq, float count,
                                   float* __restrict__ mean, float* __restrict__ rstd, float* running_mean,
                                   float* running_var, long long* num_batches_tracked, float momentum, float eps,
                                   int C) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
    float m = sum[c] / count;
    float var = fmaxf(sq[c] / count - m * m, 0.f);
    mean[c] = m;
    rstd[c] = rsqrtf(var + eps);
    if (running_mean != nullptr) {
      float unbiased = count > 1.f ? var * (count / (count - 1.f)) : var;
      running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
      running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
    }
  }
  if (c == 0 && num_batches_tracked != nullptr) *num_batches_tracked += 1;
}

// eval mode: mean = running_mean, rstd = 1/sqrt(running_var + eps)
__global__ void bn_eval_stats_kernel(const float* __restrict__ rm, const float* __restrict__ rv, float eps,
                                     float* __restrict__ mean, float* __restrict__ rstd, int C) {
  int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c < C) {
    mean[c] = rm[c];
    rstd[c] = rsqrtf(rv[c] + eps);
  }
}

// plain per-channel statistics (for inputs that did not come out of the fused conv epilogue)
__global__ void bn_stats_kernel(const float* __restrict__ x, long long P, int C, float* __restrict__ sum,
                                float* __restrict__ sq) {
  const int groups = C >> 2;
  const int lanes = blockDim.x / groups;  // pixel lanes per block
  const int g = threadIdx.x % groups;
  const int pl = threadIdx.x / groups;
  if (pl >= lanes) return;
  float4 s = make_float4(0, 0, 0, 0), q = make_float4(0, 0, 0, 0);
  for (long long p = static_cast<long long>(blockIdx.x) * lanes + pl; p < P; p += static_cast<long long>(gridDim.x) * lanes) {
    float4 v = ld4(x + p * C + g * 4);
    s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    q.x += v.x * v.x; q.y += v.y * v.y; q.z += v.z * v.z; q.w += v.w * v.w;
  }
  atomicAdd(sum + g * 4 + 0, s.x); atomicAdd(sum + g * 4 + 1, s.y);
  atomicAdd(sum + g * 4 + 2, s.z); atomicAdd(sum + g * 4 + 3, s.w);
  atomicAdd(sq + g * 4 + 0, q.x); atomicAdd(sq + g * 4 + 1, q.y);
  atomicAdd(sq + g * 4 + 2, q.z); atomicAdd(sq + g * 4 + 3, q.w);
}

__device__ __forceinline__ float4 bn_relu4(float4 v, float4 sc, float4 sh, bool relu) {
  float4 o;
  o.x = fmaf(v.x, sc.x, sh.x); o.y = fmaf(v.y, sc.y, sh.y);
  o.z = fmaf(v.z, sc.z, sh.z); o.w = fmaf(v.w, sc.w, sh.w);
  if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
  return o;
}
__device__ __forceinline__ float4 max4(float4 a, float4 b) {
  return make_float4(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), fmaxf(a.w, b.w));
}

// forward apply: y = [pool2x2](relu(x * scale + shift)),  scale = gamma*rstd, shift = beta - mean*scale
template <bool kPool>
__global__ void bn_apply_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ mean,
                                const float* __restrict__ rstd, const float* __restrict__ gamma,
                                const float* __restrict__ beta, int Nimg, int H, int W, int C, int relu) {
  const int groups = C >> 2;
  const int Ho = kPool ? H / 2 : H, Wo = kPool ? W / 2 : W;
  const long long total = static_cast<long long>(Nimg) * Ho * Wo * groups;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % groups);
    const long long po = i / groups;
    const int c = g * 4;
    const float4 m = ld4(mean + c), r = ld4(rstd + c), ga = ld4(gamma + c), be = ld4(beta + c);
    float4 sc = make_float4(ga.x * r.x, ga.y * r.y, ga.z * r.z, ga.w * r.w);
    float4 sh = make_float4(be.x - m.x * sc.x, be.y - m.y * sc.y, be.z - m.z * sc.z, be.w - m.w * sc.w);
    if (!kPool) {
      st4(y + po * C + c, bn_relu4(ld4(x + po * C + c), sc, sh, relu));
    } else {
      const int wo = static_cast<int>(po % Wo);
      const long long t = po / Wo;
      const int ho = static_cast<int>(t % Ho);
      const long long n = t / Ho;
      const float* base = x + ((n * H + 2 * ho) * W + 2 * wo) * C + c;
      float4 a = bn_relu4(ld4(base), sc, sh, relu);
      float4 b = bn_relu4(ld4(base + C), sc, sh, relu);
      float4 d = bn_relu4(ld4(base + static_cast<long long>(W) * C), sc, sh, relu);
      float4 e = bn_relu4(ld4(base + static_cast<long long>(W) * C + C), sc, sh, relu);
      st4(y + po * C + c, max4(max4(a, b), max4(d, e)));
    }
  }
}

// forward apply straight from the convolution epilogue's (sum, sum of squares): every block derives scale / shift for its
// channels itself, so no separate finalize launch sits between the convolution and this pass (8 launches and ~64 us per
// VGG step, most of it launch gaps); block 0 also writes mean / rstd for the backward and updates the running
// statistics.  training = 0: statistics are the running ones (eval mode), nothing is updated.
template <bool kPool>
__global__ void bn_apply_stats_kernel(const float* __restrict__ x, float* __restrict__ y, const float* __restrict__ sum,
                                      const float* __restrict__ sq, float count, float eps, float momentum,
                                      const float* __restrict__ gamma, const float* __restrict__ beta,
                                      float* running_mean, float* running_var, long long* num_batches_tracked,
                                      float* __restrict__ mean_out, float* __restrict__ rstd_out, int training, int Nimg,
                                      int H, int W, int C, int relu) {
  const int groups = C >> 2;
  const int Ho = kPool ? H / 2 : H, Wo = kPool ? W / 2 : W;
  const long long total = static_cast<long long>(Nimg) * Ho * Wo * groups;
  const float inv = 1.f / count;
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      float m, var;
      if (training) {
        m = sum[c] * inv;
        var = fmaxf(sq[c] * inv - m * m, 0.f);
        if (running_mean != nullptr) {
          const float unbiased = count > 1.f ? var * (count / (count - 1.f)) : var;
          running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * m;
          running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
        }
      } else {
        m = running_mean[c];
        var = running_var[c];
      }
      mean_out[c] = m;
      rstd_out[c] = rsqrtf(var + eps);
    }
    if (threadIdx.x == 0 && training && num_batches_tracked != nullptr) *num_batches_tracked += 1;
  }
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % groups);
    const long long po = i / groups;
    const int c = g * 4;
    float4 m, v;
    if (training) {
      const float4 s4 = ld4(sum + c), q4 = ld4(sq + c);
      m = make_float4(s4.x * inv, s4.y * inv, s4.z * inv, s4.w * inv);
      v = make_float4(fmaxf(q4.x * inv - m.x * m.x, 0.f), fmaxf(q4.y * inv - m.y * m.y, 0.f),
                      fmaxf(q4.z * inv - m.z * m.z, 0.f), fmaxf(q4.w * inv - m.w * m.w, 0.f));
    } else {
      // block 0 only READS the running statistics in eval mode, so every block may read them here
      m = ld4(running_mean + c); v = ld4(running_var + c);
    }
    const float4 ga = ld4(gamma + c), be = ld4(beta + c);
    const float4 sc = make_float4(ga.x * rsqrtf(v.x + eps), ga.y * rsqrtf(v.y + eps), ga.z * rsqrtf(v.z + eps),
                                  ga.w * rsqrtf(v.w + eps));
    const float4 sh = make_float4(be.x - m.x * sc.x, be.y - m.y * sc.y, be.z - m.z * sc.z, be.w - m.w * sc.w);
    if (!kPool) {
      st4(y + po * C + c, bn_relu4(ld4(x + po * C + c), sc, sh, relu));
    } else {
      const int wo = static_cast<int>(po % Wo);
      const long long t = po / Wo;
      const int ho = static_cast<int>(t % Ho);
      const long long n = t / Ho;
      const float* base = x + ((n * H + 2 * ho) * W + 2 * wo) * C + c;
      float4 a = bn_relu4(ld4(base), sc, sh, relu);
      float4 b = bn_relu4(ld4(base + C), sc, sh, relu);
      float4 d = bn_relu4(ld4(base + static_cast<long long>(W) * C), sc, sh, relu);
      float4 e = bn_relu4(ld4(base + static_cast<long long>(W) * C + C), sc, sh, relu);
      st4(y + po * C + c, max4(max4(a, b), max4(d, e)));
    }
  }
}

// ------------------------------------------------------------------------------------------------
// backward pass 1: per-channel sum(dz) and sum(dz * xhat), dz = gradient at the BN output after routing the
// incoming gradient back through (pool arg-max,) ReLU mask.  Accumulated with atomics into pre-zeroed buffers.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void route_pool(float za, float zb, float zd, float ze, int& arg, float& zmax) {
  // first maximum in window scan order (torch MaxPool2d tie-break)
  arg = 0; zmax = za;
  if (zb > zmax) { zmax = zb; arg = 1; }
  if (zd > zmax) { zmax = zd; arg = 2; }
  if (ze > zmax) { zmax = ze; arg = 3; }
}

// kU output pixels per thread per trip, every load issued before the first use (the kernel is pure bandwidth), few large
// blocks, and ONE 16-byte vector reduction per channel group per block: the first version (one pixel per trip, 1184
// blocks x 1024 scalar atomics onto the same 1024 addresses) spent 33 us on the 17 MB layers and ran the large ones at
// 4.1 TB/s.
template <bool kPool, int kU>
__global__ void __launch_bounds__(256, 3)
bn_bwd_reduce_kernel(const float* __restrict__ x, const float* __restrict__ dout,
                     const float* __restrict__ mean, const float* __restrict__ rstd,
                     const float* __restrict__ gamma, const float* __restrict__ beta, int Nimg, int H,
                     int W, int C, int relu, float* __restrict__ sum_dz, float* __restrict__ sum_dzx) {
  const int groups = C >> 2;
  const int lanes = blockDim.x / groups;
  const int g = threadIdx.x % groups;
  const int pl = threadIdx.x / groups;
  const bool active = pl < lanes;          // every thread reaches the barrier below
  const int c = g * 4;
  const float4 m4 = ld4(mean + c), r4 = ld4(rstd + c), g4 = ld4(gamma + c), b4 = ld4(beta + c);
  const float m[4] = {m4.x, m4.y, m4.z, m4.w}, r[4] = {r4.x, r4.y, r4.z, r4.w};
  const float ga[4] = {g4.x, g4.y, g4.z, g4.w}, be[4] = {b4.x, b4.y, b4.z, b4.w};
  float s1[4] = {0, 0, 0, 0}, s2[4] = {0, 0, 0, 0};
  const int Ho = kPool ? H / 2 : H, Wo = kPool ? W / 2 : W;
  const long long Pout = active ? static_cast<long long>(Nimg) * Ho * Wo : 0;
  const long long stride = static_cast<long long>(gridDim.x) * lanes;
  for (long long p0 = static_cast<long long>(blockIdx.x) * lanes + pl; p0 < Pout; p0 += stride * kU) {
    float4 dv[kU], xv[kU][kPool ? 4 : 1];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long po = p0 + u * stride;
      if (po < Pout) {
        dv[u] = ld4(dout + po * C + c);
        if (!kPool) {
          xv[u][0] = ld4(x + po * C + c);
        } else {
          const int wo = static_cast<int>(po % Wo);
          const long long t = po / Wo;
          const int ho = static_cast<int>(t % Ho);
          const long long n = t / Ho;
          const float* base = x + ((n * H + 2 * ho) * W + 2 * wo) * C + c;
          xv[u][0] = ld4(base); xv[u][1] = ld4(base + C);
          xv[u][2] = ld4(base + static_cast<long long>(W) * C); xv[u][3] = ld4(base + static_cast<long long>(W) * C + C);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      if (p0 + u * stride >= Pout) break;
      const float d[4] = {dv[u].x, dv[u].y, dv[u].z, dv[u].w};
      if (!kPool) {
        const float xs[4] = {xv[u][0].x, xv[u][0].y, xv[u][0].z, xv[u][0].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float xh = (xs[k] - m[k]) * r[k];
          const float z = fmaf(ga[k], xh, be[k]);
          const float dz = (!relu || z > 0.f) ? d[k] : 0.f;
          s1[k] += dz;
          s2[k] += dz * xh;
        }
      } else {
        const float xa[4] = {xv[u][0].x, xv[u][0].y, xv[u][0].z, xv[u][0].w}, xb[4] = {xv[u][1].x, xv[u][1].y, xv[u][1].z, xv[u][1].w};
        const float xd[4] = {xv[u][2].x, xv[u][2].y, xv[u][2].z, xv[u][2].w}, xe[4] = {xv[u][3].x, xv[u][3].y, xv[u][3].z, xv[u][3].w};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float ha = (xa[k] - m[k]) * r[k], hb = (xb[k] - m[k]) * r[k];
          const float hd = (xd[k] - m[k]) * r[k], he = (xe[k] - m[k]) * r[k];
          int arg; float zmax;
          route_pool(fmaf(ga[k], ha, be[k]), fmaf(ga[k], hb, be[k]), fmaf(ga[k], hd, be[k]), fmaf(ga[k], he, be[k]), arg, zmax);
          const float dz = (!relu || zmax > 0.f) ? d[k] : 0.f;
          const float xh = arg == 0 ? ha : (arg == 1 ? hb : (arg == 2 ? hd : he));
          s1[k] += dz;
          s2[k] += dz * xh;
        }
      }
    }
  }
  // block-level reduce across pixel lanes through shared memory, then one vector reduction per channel group per block
  extern __shared__ float red[];
  float4* r1 = reinterpret_cast<float4*>(red);
  float4* r2 = r1 + blockDim.x;
  r1[threadIdx.x] = make_float4(s1[0], s1[1], s1[2], s1[3]);
  r2[threadIdx.x] = make_float4(s2[0], s2[1], s2[2], s2[3]);
  __syncthreads();
  if (pl == 0 && active) {
    float4 a = make_float4(0, 0, 0, 0), b = make_float4(0, 0, 0, 0);
    for (int l = 0; l < lanes; ++l) {
      const float4 u = r1[l * groups + g], v = r2[l * groups + g];
      a.x += u.x; a.y += u.y; a.z += u.z; a.w += u.w;
      b.x += v.x; b.y += v.y; b.z += v.z; b.w += v.w;
    }
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(sum_dz + c), "f"(a.x), "f"(a.y), "f"(a.z), "f"(a.w) : "memory");
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(sum_dzx + c), "f"(b.x), "f"(b.y), "f"(b.z), "f"(b.w) : "memory");
  }
}

// backward pass 2: dx = gamma*rstd*(dz - mean(dz) - xhat*mean(dz*xhat)); training mode.
// eval mode (use_batch_stats = 0): dx = gamma*rstd*dz.
template <bool kPool>
__global__ void bn_bwd_apply_kernel(const float* __restrict__ x, const float* __restrict__ dout,
                                    const float* __restrict__ mean, const float* __restrict__ rstd,
                                    const float* __restrict__ gamma, const float* __restrict__ beta,
                                    const float* __restrict__ sum_dz, const float* __restrict__ sum_dzx, float inv_count,
                                    int Nimg, int H, int W, int C, int relu, int use_batch_stats, float* __restrict__ dx) {
  const int groups = C >> 2;
  const int Ho = kPool ? H / 2 : H, Wo = kPool ? W / 2 : W;
  const long long total = static_cast<long long>(Nimg) * Ho * Wo * groups;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int g = static_cast<int>(i % groups);
    const long long po = i / groups;
    const int c = g * 4;
    const float4 m4 = ld4(mean + c), r4 = ld4(rstd + c), g4 = ld4(gamma + c), b4 = ld4(beta + c);
    const float4 sd4 = ld4(sum_dz + c), sx4 = ld4(sum_dzx + c);
    const float m[4] = {m4.x, m4.y, m4.z, m4.w}, r[4] = {r4.x, r4.y, r4.z, r4.w};
    const float ga[4] = {g4.x, g4.y, g4.z, g4.w}, be[4] = {b4.x, b4.y, b4.z, b4.w};
    float md[4] = {sd4.x * inv_count, sd4.y * inv_count, sd4.z * inv_count, sd4.w * inv_count};
    float mx[4] = {sx4.x * inv_count, sx4.y * inv_count, sx4.z * inv_count, sx4.w * inv_count};
    if (!use_batch_stats) {
#pragma unroll
      for (int k = 0; k < 4; ++k) { md[k] = 0.f; mx[k] = 0.f; }
    }
    const float4 d4 = ld4(dout + po * C + c);
    const float d[4] = {d4.x, d4.y, d4.z, d4.w};
    if (!kPool) {
      const float4 x4 = ld4(x + po * C + c);
      const float xv[4] = {x4.x, x4.y, x4.z, x4.w};
      float o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float xh = (xv[k] - m[k]) * r[k];
        const float z = fmaf(ga[k], xh, be[k]);
        const float dz = (!relu || z > 0.f) ? d[k] : 0.f;
        o[k] = ga[k] * r[k] * (dz - md[k] - xh * mx[k]);
      }
      st4(dx + po * C + c, make_float4(o[0], o[1], o[2], o[3]));
    } else {
      const int wo = static_cast<int>(po % Wo);
      const long long t = po / Wo;
      const int ho = static_cast<int>(t % Ho);
      const long long n = t / Ho;
      const long long off = ((n * H + 2 * ho) * W + 2 * wo) * C + c;
      const long long offs[4] = {off, off + C, off + static_cast<long long>(W) * C, off + static_cast<long long>(W) * C + C};
      float xv[4][4];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const float4 v = ld4(x + offs[q]);
        xv[q][0] = v.x; xv[q][1] = v.y; xv[q][2] = v.z; xv[q][3] = v.w;
      }
      float o[4][4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        float h[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) h[q] = (xv[q][k] - m[k]) * r[k];
        int arg; float zmax;
        route_pool(fmaf(ga[k], h[0], be[k]), fmaf(ga[k], h[1], be[k]), fmaf(ga[k], h[2], be[k]), fmaf(ga[k], h[3], be[k]), arg, zmax);
        const float dzv = (!relu || zmax > 0.f) ? d[k] : 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) o[q][k] = ga[k] * r[k] * ((q == arg ? dzv : 0.f) - md[k] - h[q] * mx[k]);
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) st4(dx + offs[q], make_float4(o[q][0], o[q][1], o[q][2], o[q][3]));
    }
  }
}

inline int grid_for(long long total, int block, int max_blocks = 148 * 8) {
  long long b = (total + block - 1) / block;
  if (b > max_blocks) b = max_blocks;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

}  // namespace

extern "C" int ddpx_bn_finalize(const float* sum, const float* sq, float count, float* mean, float* rstd,
                                float* running_mean, float* running_var, long long* num_batches_tracked, float momentum,
                                float eps, int C, void* stream) {
  bn_finalize_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      sum, sq, count, mean, rstd, running_mean, running_var, num_batches_tracked, momentum, eps, C);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" int ddpx_bn_eval_stats(const float* running_mean, const float* running_var, float eps, float* mean,
                                  float* rstd, int C, void* stream) {
  bn_eval_stats_kernel<<<(C + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(running_mean, running_var, eps,
                                                                                     mean, rstd, C);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" int ddpx_bn_stats(const float* x, long long P, int C, float* sum, float* sq, void* stream) {
  if (C <= 0 || C % 4 || C > 1024) return DDPX_ERR_SHAPE;
  const int groups = C / 4;
  int block = 256;
  if (groups > block) block = groups;
  bn_stats_kernel<<<grid_for(P, block / groups, 148 * 4), block, 0, static_cast<cudaStream_t>(stream)>>>(x, P, C, sum, sq);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" int ddpx_bn_apply(const float* x, float* y, const float* mean, const float* rstd, const float* gamma,
                             const float* beta, int Nimg, int H, int W, int C, int relu, int pool, void* stream) {
  if (C % 4) return DDPX_ERR_SHAPE;
  if (pool && ((H | W) & 1)) return DDPX_ERR_SHAPE;
  const long long total = static_cast<long long>(Nimg) * (pool ? H / 2 : H) * (pool ? W / 2 : W) * (C / 4);
  const int block = 256;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (pool)
    bn_apply_kernel<true><<<grid_for(total, block, 148 * 16), block, 0, s>>>(x, y, mean, rstd, gamma, beta, Nimg, H, W, C, relu);
  else
    bn_apply_kernel<false><<<grid_for(total, block, 148 * 16), block, 0, s>>>(x, y, mean, rstd, gamma, beta, Nimg, H, W, C, relu);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

// BatchNorm forward from raw (sum, sum of squares): finalize + running-statistics update + normalise (+ReLU, +pool) in ONE
// launch.  training = 0 normalises with the running statistics (sum / sq may be null).  mean_rstd [2][C] is an output.
extern "C" int ddpx_bn_apply_stats(const float* x, float* y, const float* sum, const float* sq, float count, float eps,
                                   float momentum, const float* gamma, const float* beta, float* running_mean,
                                   float* running_var, long long* num_batches_tracked, float* mean_rstd, int training,
                                   int Nimg, int H, int W, int C, int relu, int pool, void* stream) {
  if (C <= 0 || C % 4) return DDPX_ERR_SHAPE;
  if (pool && ((H | W) & 1)) return DDPX_ERR_SHAPE;
  if (training ? (sum == nullptr || sq == nullptr) : (running_mean == nullptr || running_var == nullptr)) return DDPX_ERR_ARG;
  const long long total = static_cast<long long>(Nimg) * (pool ? H / 2 : H) * (pool ? W / 2 : W) * (C / 4);
  const int block = 256;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (pool)
    bn_apply_stats_kernel<true><<<grid_for(total, block, 148 * 16), block, 0, s>>>(
        x, y, sum, sq, count, eps, momentum, gamma, beta, running_mean, running_var, num_batches_tracked, mean_rstd,
        mean_rstd + C, training, Nimg, H, W, C, relu);
  else
    bn_apply_stats_kernel<false><<<grid_for(total, block, 148 * 16), block, 0, s>>>(
        x, y, sum, sq, count, eps, momentum, gamma, beta, running_mean, running_var, num_batches_tracked, mean_rstd,
        mean_rstd + C, training, Nimg, H, W, C, relu);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" int ddpx_bn_bwd_reduce(const float* x, const float* dout, const float* mean, const float* rstd,
                                  const float* gamma, const float* beta, int Nimg, int H, int W, int C, int relu, int pool,
                                  float* sum_dz, float* sum_dzx, void* stream) {
  if (C % 4 || C > 1024) return DDPX_ERR_SHAPE;
  if (pool && ((H | W) & 1)) return DDPX_ERR_SHAPE;
  const int groups = C / 4;
  const int block = 256;
  const int lanes = block / groups;
  const long long Pout = static_cast<long long>(Nimg) * (pool ? H / 2 : H) * (pool ? W / 2 : W);
  // sum_dz / sum_dzx take 16-byte vector reductions
  if ((reinterpret_cast<uintptr_t>(sum_dz) | reinterpret_cast<uintptr_t>(sum_dzx)) & 15) return DDPX_ERR_ALIGN;
  const int grid = grid_for(Pout, lanes * 4, 148 * 3);
  const size_t smem = static_cast<size_t>(block) * 8 * sizeof(float);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (pool)
    bn_bwd_reduce_kernel<true, 2><<<grid, block, smem, s>>>(x, dout, mean, rstd, gamma, beta, Nimg, H, W, C, relu, sum_dz, sum_dzx);
  else
    bn_bwd_reduce_kernel<false, 4><<<grid, block, smem, s>>>(x, dout, mean, rstd, gamma, beta, Nimg, H, W, C, relu, sum_dz, sum_dzx);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" int ddpx_bn_bwd_apply(const float* x, const float* dout, const float* mean, const float* rstd,
                                 const float* gamma, const float* beta, const float* sum_dz, const float* sum_dzx,
                                 int Nimg, int H, int W, int C, int relu, int pool, int use_batch_stats, float* dx,
                                 void* stream) {
  if (C % 4) return DDPX_ERR_SHAPE;
  const long long total = static_cast<long long>(Nimg) * (pool ? H / 2 : H) * (pool ? W / 2 : W) * (C / 4);
  const float inv_count = 1.f / (static_cast<float>(Nimg) * H * W);
  const int block = 256;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (pool)
    bn_bwd_apply_kernel<true><<<grid_for(total, block, 148 * 16), block, 0, s>>>(
        x, dout, mean, rstd, gamma, beta, sum_dz, sum_dzx, inv_count, Nimg, H, W, C, relu, use_batch_stats, dx);
  else
    bn_bwd_apply_kernel<false><<<grid_for(total, block, 148 * 16), block, 0, s>>>(
        x, dout, mean, rstd, gamma, beta, sum_dz, sum_dzx, inv_count, Nimg, H, W, C, relu, use_batch_stats, dx);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

// Training-mode BatchNorm2d forward in ONE call: batch statistics -> mean / rstd (+ running-stat update) -> normalise
// (+ReLU, +2x2 max-pool).  `workspace` is scratch for the per-channel sum / sum-of-squares accumulators: 2*C floats that
// this call clears itself; `mean_rstd` ([2][C]) is an output kept for the backward.  Used by the stand-alone
// nn.BatchNorm2d path (inputs that do not come out of the fused conv epilogue).
extern "C" int ddpx_bn_train_fwd(const float* x, float* y, const float* gamma, const float* beta, float* running_mean,
                                 float* running_var, long long* num_batches_tracked, float momentum, float eps, int Nimg,
                                 int H, int W, int C, int relu, int pool, float* mean_rstd, float* workspace, void* stream) {
  // every argument check comes before the first memset / launch: a refused call leaves the running statistics untouched
  if (C <= 0 || C % 4 || C > 1024) return DDPX_ERR_SHAPE;
  if (pool && ((H | W) & 1)) return DDPX_ERR_SHAPE;
  if (Nimg <= 0 || H <= 0 || W <= 0) return DDPX_ERR_ARG;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (cudaMemsetAsync(workspace, 0, sizeof(float) * 2 * C, s) != cudaSuccess) return DDPX_ERR_LAUNCH;
  const long long P = static_cast<long long>(Nimg) * H * W;
  int rc = ddpx_bn_stats(x, P, C, workspace, workspace + C, stream);
  if (rc != DDPX_OK) return rc;
  rc = ddpx_bn_finalize(workspace, workspace + C, static_cast<float>(P), mean_rstd, mean_rstd + C, running_mean, running_var,
                        num_batches_tracked, momentum, eps, C, stream);
  if (rc != DDPX_OK) return rc;
  return ddpx_bn_apply(x, y, mean_rstd, mean_rstd + C, gamma, beta, Nimg, H, W, C, relu, pool, stream);
}

// ------------------------------------------------------------------------------------------------
// ddpx_bn_train_fwd_static: the same training-mode forward for callers that have no memory to hand over.
//
// Who this is for.  ddpx_bn_train_fwd needs 2*C floats of workspace from its caller.  Some users of this C ABI cannot
// produce one: a host program (or a binding in another language) that was given its tensors by somebody else and owns
// no device allocator; a runtime whose device memory was planned and committed up front, so that nothing may be carved
// out afterwards; a service loop that must not call cudaMalloc because that call can stall every stream of the device
// while it runs.  For them the accumulators live in the library itself: the fixed array below, DDPX_BN_STATIC_MAX_C
// channels wide (2 * 1024 floats = 8 KiB).  Every GPU that loads this library carries those 8 KiB until the process
// exits, whether or not the entry point is ever called.
//
// What the call does and does not do.  It issues a memset of the first 2*C floats of the array and the same three
// kernels as ddpx_bn_train_fwd on `stream`; it never calls cudaMalloc, cudaFree or any routine that allocates.  Inputs
// and outputs (x, y, gamma, beta, running statistics, mean_rstd) are still the caller's.  C above
// DDPX_BN_STATIC_MAX_C is refused with DDPX_ERR_LIMIT and a message on stderr; nothing is launched in that case.
//
// First call on a device.  The address of the array is looked up once per device (cudaGetSymbolAddress, with that
// device current) and remembered; the look-up may make the CUDA runtime load this library's device code.  Make the
// first call on each device during start-up, before any stream capture begins on that device.
//
// Sharing.  All calls on a device use the same array and it keeps whatever the previous call left there, which is why
// every call clears the part it uses first.  Two calls that can run at the same time on one GPU (different streams,
// different host threads) would accumulate into each other's sums: keep calls to this entry point ordered on one
// stream, or use ddpx_bn_train_fwd with separate workspaces.
// ------------------------------------------------------------------------------------------------
#define DDPX_BN_STATIC_MAX_C 1024
static __device__ __align__(256) float g_bn_static_scratch[2 * DDPX_BN_STATIC_MAX_C];

namespace {
constexpr int kMaxDevices = 64;
std::atomic<float*> g_bn_static_ptr[kMaxDevices];        // zero-initialised; every thread would store the same value

// `g_bn_static_scratch` in host code is only a handle for the runtime, not a device address: ask the runtime for the
// address on the current device, once.
int bn_static_scratch_ptr(float** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return DDPX_ERR_DRIVER;
  if (g_bn_static_ptr[dev].load(std::memory_order_acquire) == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_bn_static_scratch);
    if (e != cudaSuccess || p == nullptr) {
      fprintf(stderr, "[ddpx] ddpx_bn_train_fwd_static: cudaGetSymbolAddress failed on device %d: %s\n", dev,
              cudaGetErrorString(e));
      return DDPX_ERR_DRIVER;
    }
    g_bn_static_ptr[dev].store(static_cast<float*>(p), std::memory_order_release);
  }
  *out = g_bn_static_ptr[dev].load(std::memory_order_acquire);
  return DDPX_OK;
}
}  // namespace

extern "C" int ddpx_bn_train_fwd_static(const float* x, float* y, const float* gamma, const float* beta,
                                        float* running_mean, float* running_var, long long* num_batches_tracked,
                                        float momentum, float eps, int Nimg, int H, int W, int C, int relu, int pool,
                                        float* mean_rstd, void* stream) {
  if (C > DDPX_BN_STATIC_MAX_C) {
    fprintf(stderr,
            "[ddpx] ddpx_bn_train_fwd_static: C = %d exceeds the built-in scratch (DDPX_BN_STATIC_MAX_C = %d channels); "
            "use ddpx_bn_train_fwd with a caller-provided workspace\n", C, DDPX_BN_STATIC_MAX_C);
    return DDPX_ERR_LIMIT;
  }
  if (C <= 0 || C % 4) return DDPX_ERR_SHAPE;
  float* scratch = nullptr;
  const int rc = bn_static_scratch_ptr(&scratch);
  if (rc != DDPX_OK) return rc;
  // ddpx_bn_train_fwd clears the 2*C floats it accumulates into, so the previous call's sums never leak in
  return ddpx_bn_train_fwd(x, y, gamma, beta, running_mean, running_var, num_batches_tracked, momentum, eps, Nimg, H, W, C,
                           relu, pool, mean_rstd, scratch, stream);
}
