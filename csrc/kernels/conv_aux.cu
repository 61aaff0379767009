// Convolution helpers that do not need tensor cores, sm_100a:
//   * weight re-packing torch [Cout][Cin][3][3] -> [tap][Cout][Cin] (fprop) and tap-flipped [tap][Cin][Cout] (dgrad)
//   * direct 3x3 convolution for tiny Cin (the RGB stem, K = 27): forward (+bias, +BN statistics) and weight gradient
//   * NCHW <-> NHWC layout transposes, stand-alone ReLU / 2x2 max-pool forward+backward, spatial mean
// Reference parity: conv0 at /root/reference/multigpu.py:53, pools :59, mean :68, DeepNN S:21-31.
#include <cuda_runtime.h>
#include <string.h>
#include <stdint.h>
#include "../../include/ddpx_kernels.h"

namespace {

inline int nblocks(long long total, int block, int cap = 148 * 16) {
  long long b = (total + block - 1) / block;
  if (b > cap) b = cap;
  if (b < 1) b = 1;
  return static_cast<int>(b);
}

// ------------------------------------------------------------------------------------------------
// weight packing
// ------------------------------------------------------------------------------------------------
__global__ void pack_conv_w_kernel(const float* __restrict__ w, float* __restrict__ wf, float* __restrict__ wd, int Cout,
                                   int Cin) {
  const long long total = static_cast<long long>(Cout) * Cin * 9;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += static_cast<long long>(gridDim.x) * blockDim.x) {
    // i indexes the fprop layout [tap][co][ci] so that writes are coalesced
    const int ci = static_cast<int>(i % Cin);
    const long long t = i / Cin;
    const int co = static_cast<int>(t % Cout);
    const int tap = static_cast<int>(t / Cout);
    const float v = w[(static_cast<long long>(co) * Cin + ci) * 9 + tap];
    if (wf != nullptr) wf[i] = v;
    if (wd != nullptr) wd[(static_cast<long long>(8 - tap) * Cin + ci) * Cout + co] = v;  // flipped tap, transposed
  }
}

// Tiled flavour for Cout, Cin multiples of 32: a block stages w[32 co][32 ci][9] (contiguous 1152-byte rows) in shared
// memory and writes both packed layouts with 128-byte rows; the simple kernel above gathers with a 36-byte stride and
// scatters the dgrad layout with a Cout-float stride (126 us per step for the 7 VGG layers, ~20 us of traffic).
__global__ void __launch_bounds__(256)
pack_conv_w_tiled_kernel(const float* __restrict__ w, float* __restrict__ wf, float* __restrict__ wd, int Cout, int Cin) {
  __shared__ float tile[32][289];
  const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 32;
  for (int idx = threadIdx.x; idx < 32 * 288; idx += 256) {
    const int co = idx / 288, j = idx - co * 288;
    tile[co][j] = w[(static_cast<long long>(co0 + co) * Cin + ci0) * 9 + j];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 9 * 1024; idx += 256) {
    const int tap = idx >> 10, a = (idx >> 5) & 31, b = idx & 31;
    if (wf != nullptr)      // a = co, b = ci
      wf[(static_cast<long long>(tap) * Cout + co0 + a) * Cin + ci0 + b] = tile[a][b * 9 + tap];
    if (wd != nullptr)      // a = ci, b = co; flipped tap, transposed
      wd[(static_cast<long long>(8 - tap) * Cin + ci0 + a) * Cout + co0 + b] = tile[b][a * 9 + tap];
  }
}

// inverse direction for the weight gradient: staging layout [9][Cout][Cin] -> torch layout [Cout][Cin][3][3]
__global__ void __launch_bounds__(256)
unpack_conv_wgrad_kernel(const float* __restrict__ wp, float* __restrict__ w, int Cout, int Cin) {
  __shared__ float tile[32][289];
  const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 32;
  for (int idx = threadIdx.x; idx < 9 * 1024; idx += 256) {
    const int tap = idx >> 10, co = (idx >> 5) & 31, ci = idx & 31;
    tile[co][ci * 9 + tap] = wp[(static_cast<long long>(tap) * Cout + co0 + co) * Cin + ci0 + ci];
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 32 * 288; idx += 256) {
    const int co = idx / 288, j = idx - co * 288;
    w[(static_cast<long long>(co0 + co) * Cin + ci0) * 9 + j] = tile[co][j];
  }
}

// All conv layers of a model in ONE launch: per-layer tiles of the kernel above, dealt to a flat grid, plus the stem's
// [Cout][27] -> [Cout][32] padding and the clearing of every layer's BatchNorm statistic accumulators.  Eight per-layer
// launches with 8..256 blocks each took 113 us per VGG step for ~110 MB of traffic; one launch keeps every SM busy.
constexpr int kMaxPackJobs = 32;
struct PackJob {
  const float* w;
  float* wf;            // fprop layout [9][Cout][Cin]   (stem: [Cout][32])
  float* wd;            // dgrad layout [9][Cin][Cout]   (may be null)
  float* zero;          // floats to clear (BatchNorm sum / sum-of-squares of this layer), may be null
  int zero_n;
  int Cout, Cin;        // Cin == 3 marks the stem
  int tiles_ci;         // Cin / 32
  int block_begin;      // first block of this job in the flat grid
};
struct PackTable { PackJob j[kMaxPackJobs]; int n; };

__global__ void __launch_bounds__(256)
pack_conv_multi_kernel(const __grid_constant__ PackTable t) {
  __shared__ float tile[32][289];
  int ji = 0;
  while (ji + 1 < t.n && static_cast<int>(blockIdx.x) >= t.j[ji + 1].block_begin) ++ji;
  const PackJob& job = t.j[ji];
  const int local = static_cast<int>(blockIdx.x) - job.block_begin;
  if (local == 0 && job.zero != nullptr)
    for (int i = threadIdx.x; i < job.zero_n; i += 256) job.zero[i] = 0.f;
  if (job.Cin == 3) {
    const int i = local * 256 + threadIdx.x;
    if (i < job.Cout * 32) {
      const int co = i >> 5, k = i & 31;
      job.wf[i] = k < 27 ? job.w[co * 27 + k] : 0.f;
    }
    return;
  }
  const int Cout = job.Cout, Cin = job.Cin;
  const int ci0 = (local % job.tiles_ci) * 32, co0 = (local / job.tiles_ci) * 32;
  // 32 rows of 288 contiguous floats (16-byte aligned: ci0 * 36 bytes * 32) -> 72 float4 per row
  for (int idx = threadIdx.x; idx < 32 * 72; idx += 256) {
    const int co = idx / 72, j4 = idx - co * 72;
    const float4 v = *reinterpret_cast<const float4*>(job.w + (static_cast<long long>(co0 + co) * Cin + ci0) * 9 + j4 * 4);
    tile[co][j4 * 4] = v.x; tile[co][j4 * 4 + 1] = v.y; tile[co][j4 * 4 + 2] = v.z; tile[co][j4 * 4 + 3] = v.w;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 9 * 1024; idx += 256) {
    const int tap = idx >> 10, a = (idx >> 5) & 31, b = idx & 31;
    job.wf[(static_cast<long long>(tap) * Cout + co0 + a) * Cin + ci0 + b] = tile[a][b * 9 + tap];
    if (job.wd != nullptr)
      job.wd[(static_cast<long long>(8 - tap) * Cin + ci0 + a) * Cout + co0 + b] = tile[b][a * 9 + tap];
  }
}

// unpack that also clears the staging tile it has read: the staging buffer of a layer is persistent, so the next step's
// weight-gradient kernel finds zeros without a memset launch (24 ddpx_zero launches, 133 us per VGG step, before)
__global__ void __launch_bounds__(256)
unpack_conv_wgrad_rezero_kernel(float* __restrict__ wp, float* __restrict__ w, int Cout, int Cin) {
  __shared__ float tile[16][289];
  const int ci0 = blockIdx.x * 32, co0 = blockIdx.y * 16;
  // all 18 loads of a thread are issued before its first zero store: loads and stores hit the same addresses, so the
  // compiler must keep their order, and load -> store -> load was 18 dependent round trips (242 us per step)
  float v[18];
#pragma unroll
  for (int k = 0; k < 18; ++k) {
    const int idx = threadIdx.x + k * 256;
    const int tap = idx >> 9, co = (idx >> 5) & 15, ci = idx & 31;
    v[k] = wp[(static_cast<long long>(tap) * Cout + co0 + co) * Cin + ci0 + ci];
  }
#pragma unroll
  for (int k = 0; k < 18; ++k) {
    const int idx = threadIdx.x + k * 256;
    const int tap = idx >> 9, co = (idx >> 5) & 15, ci = idx & 31;
    tile[co][ci * 9 + tap] = v[k];
    wp[(static_cast<long long>(tap) * Cout + co0 + co) * Cin + ci0 + ci] = 0.f;
  }
  __syncthreads();
  for (int idx = threadIdx.x; idx < 16 * 72; idx += 256) {
    const int co = idx / 72, j4 = idx - co * 72;
    *reinterpret_cast<float4*>(w + (static_cast<long long>(co0 + co) * Cin + ci0) * 9 + j4 * 4) =
        make_float4(tile[co][j4 * 4], tile[co][j4 * 4 + 1], tile[co][j4 * 4 + 2], tile[co][j4 * 4 + 3]);
  }
}

// ------------------------------------------------------------------------------------------------
// direct conv for Cin == 3 (input NCHW, output NHWC).  One thread = one output pixel, all Cout channels.
// ------------------------------------------------------------------------------------------------
constexpr int kStemPix = 128;  // pixels per block

__global__ void __launch_bounds__(kStemPix)
conv_stem_fprop_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                       float* __restrict__ y, int Nimg, int H, int W, int Cout, int act, float* __restrict__ stat_sum,
                       float* __restrict__ stat_sq) {
  extern __shared__ float sm[];
  float* sw = sm;                        // [27][Cout]  (tap-major so the per-cout read is a broadcast-free row)
  float* tile = sm + 27 * Cout;          // [kStemPix][Cout + 1]
  for (int i = threadIdx.x; i < 27 * Cout; i += blockDim.x) {
    const int co = i % Cout, k = i / Cout;   // k = ci*9 + tap
    sw[i] = w[co * 27 + k];
  }
  __syncthreads();
  const long long P = static_cast<long long>(Nimg) * H * W;
  const int ld = Cout + 1;
  for (long long p0 = static_cast<long long>(blockIdx.x) * kStemPix; p0 < P; p0 += static_cast<long long>(gridDim.x) * kStemPix) {
    const long long p = p0 + threadIdx.x;
    float in[27];
    if (p < P) {
      const int wq = static_cast<int>(p % W);
      const long long t = p / W;
      const int hq = static_cast<int>(t % H);
      const long long n = t / H;
#pragma unroll
      for (int ci = 0; ci < 3; ++ci)
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int s = 0; s < 3; ++s) {
            const int hh = hq + r - 1, ww = wq + s - 1;
            in[ci * 9 + r * 3 + s] = (hh >= 0 && hh < H && ww >= 0 && ww < W)
                                         ? __ldg(x + ((n * 3 + ci) * H + hh) * W + ww) : 0.f;
          }
    }
    for (int co = 0; co < Cout; ++co) {
      float acc = bias != nullptr ? __ldg(bias + co) : 0.f;
#pragma unroll
      for (int k = 0; k < 27; ++k) acc = fmaf(in[k], sw[k * Cout + co], acc);
      tile[threadIdx.x * ld + co] = (p < P) ? acc : 0.f;
    }
    __syncthreads();
    // coalesced store + per-channel statistics from the tile
    const long long valid = (P - p0 < kStemPix) ? (P - p0) : kStemPix;
    for (int i = threadIdx.x; i < kStemPix * Cout; i += blockDim.x) {
      const int px = i / Cout, co = i % Cout;
      if (px < valid) {
        float v = tile[px * ld + co];
        if (act == 1) v = fmaxf(v, 0.f);
        y[(p0 + px) * Cout + co] = v;
      }
    }
    if (stat_sum != nullptr) {
      for (int co = threadIdx.x; co < Cout; co += blockDim.x) {
        float s = 0.f, q = 0.f;
        for (int px = 0; px < valid; ++px) { const float v = tile[px * ld + co]; s += v; q += v * v; }
        atomicAdd(stat_sum + co, s);
        atomicAdd(stat_sq + co, q);
      }
    }
    __syncthreads();
  }
}

// weight gradient for Cin == 3: dw[co][ci][r][s] += sum_pixels dy[p][co] * x[n][ci][h+r-1][w+s-1]
// block = one image at a time (padded image staged in shared memory); thread = (co, tap-group)
__global__ void __launch_bounds__(256)
conv_stem_wgrad_kernel(const float* __restrict__ dy, const float* __restrict__ x, float* __restrict__ dw, int Nimg, int H,
                       int W, int Cout) {
  extern __shared__ float sm[];
  const int PH = H + 2, PW = W + 2;
  float* img = sm;  // [3][PH][PW]
  const int groups = blockDim.x / Cout;          // how many tap groups fit
  const int co = threadIdx.x % Cout;
  const int grp = threadIdx.x / Cout;
  const int per = (27 + groups - 1) / groups;    // taps per group
  const int k0 = grp * per;
  float acc[27];
#pragma unroll
  for (int k = 0; k < 27; ++k) acc[k] = 0.f;
  for (int n = blockIdx.x; n < Nimg; n += gridDim.x) {
    __syncthreads();
    for (int i = threadIdx.x; i < 3 * PH * PW; i += blockDim.x) {
      const int ww = i % PW - 1, hh = (i / PW) % PH - 1, ci = i / (PW * PH);
      img[i] = (hh >= 0 && hh < H && ww >= 0 && ww < W) ? __ldg(x + ((static_cast<long long>(n) * 3 + ci) * H + hh) * W + ww) : 0.f;
    }
    __syncthreads();
    if (grp < groups) {
      for (int h = 0; h < H; ++h)
        for (int wq = 0; wq < W; ++wq) {
          const float g = __ldg(dy + ((static_cast<long long>(n) * H + h) * W + wq) * Cout + co);
#pragma unroll
          for (int j = 0; j < 27; ++j) {
            if (j < per) {
              const int k = k0 + j;
              if (k < 27) {
                const int ci = k / 9, r = (k % 9) / 3, s = k % 3;
                acc[j] = fmaf(g, img[(ci * PH + h + r) * PW + wq + s], acc[j]);
              }
            }
          }
        }
    }
  }
  if (grp < groups) {
#pragma unroll
    for (int j = 0; j < 27; ++j)
      if (j < per && k0 + j < 27) atomicAdd(dw + co * 27 + k0 + j, acc[j]);
  }
}

// ------------------------------------------------------------------------------------------------
// layout transposes (per image: [C][HW] <-> [HW][C]) through a padded shared-memory tile
// ------------------------------------------------------------------------------------------------
__global__ void transpose_batched_kernel(const float* __restrict__ in, float* __restrict__ out, int rows, int cols) {
  // in: [batch][rows][cols] -> out: [batch][cols][rows]
  __shared__ float t[32][33];
  const long long b = blockIdx.z;
  const float* src = in + b * rows * cols;
  float* dst = out + b * rows * cols;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    if (r < rows && c < cols) t[j][threadIdx.x] = src[static_cast<long long>(r) * cols + c];
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = c0 + j, r = r0 + threadIdx.x;
    if (r < rows && c < cols) dst[static_cast<long long>(c) * rows + r] = t[threadIdx.x][j];
  }
}

// ------------------------------------------------------------------------------------------------
// stand-alone ReLU / max-pool (NHWC) and spatial mean
// ------------------------------------------------------------------------------------------------
__global__ void relu_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, long long n) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    y[i] = fmaxf(x[i], 0.f);
}
__global__ void relu_bwd_kernel(const float* __restrict__ y, const float* __restrict__ dy, float* __restrict__ dx, long long n) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}
__global__ void maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int Nimg, int H, int W, int C) {
  const int Ho = H / 2, Wo = W / 2;
  const long long total = static_cast<long long>(Nimg) * Ho * Wo * C;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    long long t = i / C;
    const int wo = static_cast<int>(t % Wo); t /= Wo;
    const int ho = static_cast<int>(t % Ho);
    const long long n = t / Ho;
    const float* b = x + ((n * H + 2 * ho) * W + 2 * wo) * C + c;
    y[i] = fmaxf(fmaxf(b[0], b[C]), fmaxf(b[static_cast<long long>(W) * C], b[static_cast<long long>(W) * C + C]));
  }
}
__global__ void maxpool_bwd_kernel(const float* __restrict__ x, const float* __restrict__ dy, float* __restrict__ dx, int Nimg,
                                   int H, int W, int C) {
  const int Ho = H / 2, Wo = W / 2;
  const long long total = static_cast<long long>(Nimg) * Ho * Wo * C;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    long long t = i / C;
    const int wo = static_cast<int>(t % Wo); t /= Wo;
    const int ho = static_cast<int>(t % Ho);
    const long long n = t / Ho;
    const long long o = ((n * H + 2 * ho) * W + 2 * wo) * C + c;
    const long long offs[4] = {o, o + C, o + static_cast<long long>(W) * C, o + static_cast<long long>(W) * C + C};
    int arg = 0; float m = x[offs[0]];
#pragma unroll
    for (int q = 1; q < 4; ++q) { const float v = x[offs[q]]; if (v > m) { m = v; arg = q; } }
    const float g = dy[i];
#pragma unroll
    for (int q = 0; q < 4; ++q) dx[offs[q]] = (q == arg) ? g : 0.f;
  }
}
// y[n][c] = mean over hw of x[n][hw][c]; backward broadcasts dy/HW
__global__ void spatial_mean_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int Nimg, int HW, int C) {
  const long long total = static_cast<long long>(Nimg) * C;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const long long n = i / C;
    float s = 0.f;
    for (int p = 0; p < HW; ++p) s += x[(n * HW + p) * C + c];
    y[i] = s / HW;
  }
}
__global__ void spatial_mean_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int Nimg, int HW, int C) {
  const long long total = static_cast<long long>(Nimg) * HW * C;
  const float inv = 1.f / HW;
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int c = static_cast<int>(i % C);
    const long long n = i / (static_cast<long long>(HW) * C);
    dx[i] = dy[n * C + c] * inv;
  }
}
// column sums: out[c] (+)= sum_r x[r][c]   (bias gradients)
__global__ void colsum_kernel(const float* __restrict__ x, float* __restrict__ out, long long rows, int cols, int accumulate) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (long long r = blockIdx.y; r < rows; r += gridDim.y) s += x[r * cols + c];
  if (gridDim.y == 1 && !accumulate) out[c] = s; else atomicAdd(out + c, s);
}
__global__ void scale_copy_kernel(const float* __restrict__ src, float* __restrict__ dst, long long n, float scale, int accumulate) {
  for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += static_cast<long long>(gridDim.x) * blockDim.x)
    dst[i] = accumulate ? dst[i] + src[i] * scale : src[i] * scale;
}

}  // namespace

#define LAUNCH_OK() (cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH)

extern "C" int ddpx_pack_conv_weights(const float* w, float* w_fprop, float* w_dgrad, int Cout, int Cin, void* stream) {
  const long long total = static_cast<long long>(Cout) * Cin * 9;
  if (Cout % 32 == 0 && Cin % 32 == 0 && Cout / 32 <= 65535) {
    dim3 grid(Cin / 32, Cout / 32);
    pack_conv_w_tiled_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(w, w_fprop, w_dgrad, Cout, Cin);
    return LAUNCH_OK();
  }
  pack_conv_w_kernel<<<nblocks(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(w, w_fprop, w_dgrad, Cout, Cin);
  return LAUNCH_OK();
}

// Weight packing for n conv layers in one launch (see pack_conv_multi_kernel).  Layer i: w[i] is [Cout][Cin][3][3];
// wf[i] receives the fprop layout, wd[i] (may be null) the dgrad layout; zero[i] (may be null) are zero_n[i] floats to clear.
// Cin == 3 marks the RGB stem (wf = [Cout][32] zero-padded rows).
extern "C" int ddpx_pack_conv_weights_multi(int n, const void* const* w, void* const* wf, void* const* wd, void* const* zero,
                                            const int* zero_n, const int* Cout, const int* Cin, void* stream) {
  if (n < 1 || n > kMaxPackJobs) return DDPX_ERR_LIMIT;
  PackTable t;
  memset(&t, 0, sizeof(t));
  t.n = n;
  int blocks = 0;
  for (int i = 0; i < n; ++i) {
    PackJob& j = t.j[i];
    j.w = static_cast<const float*>(w[i]); j.wf = static_cast<float*>(wf[i]);
    j.wd = wd ? static_cast<float*>(wd[i]) : nullptr;
    j.zero = zero ? static_cast<float*>(zero[i]) : nullptr;
    j.zero_n = zero_n ? zero_n[i] : 0;
    j.Cout = Cout[i]; j.Cin = Cin[i];
    j.block_begin = blocks;
    if (Cin[i] == 3) {
      blocks += (Cout[i] * 32 + 255) / 256;
    } else {
      if (Cout[i] % 32 || Cin[i] % 32) return DDPX_ERR_SHAPE;
      j.tiles_ci = Cin[i] / 32;
      blocks += (Cout[i] / 32) * (Cin[i] / 32);
    }
  }
  pack_conv_multi_kernel<<<blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(t);
  return LAUNCH_OK();
}

// unpack + clear: dw_packed is left all-zero for the next accumulation
extern "C" int ddpx_unpack_conv_wgrad_rezero(float* dw_packed, float* dw, int Cout, int Cin, void* stream) {
  if (Cout % 32 || Cin % 32 || Cout / 16 > 65535) return DDPX_ERR_SHAPE;
  dim3 grid(Cin / 32, Cout / 16);
  unpack_conv_wgrad_rezero_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(dw_packed, dw, Cout, Cin);
  return LAUNCH_OK();
}

extern "C" int ddpx_unpack_conv_wgrad(const float* dw_packed, float* dw, int Cout, int Cin, void* stream) {
  if (Cout % 32 || Cin % 32 || Cout / 32 > 65535) return DDPX_ERR_SHAPE;
  dim3 grid(Cin / 32, Cout / 32);
  unpack_conv_wgrad_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(dw_packed, dw, Cout, Cin);
  return LAUNCH_OK();
}

extern "C" int ddpx_conv_stem_fprop(const float* x_nchw, const float* w, const float* bias, float* y_nhwc, int Nimg, int H,
                                    int W, int Cout, int act, float* stat_sum, float* stat_sq, void* stream) {
  if (Cout > 128) return DDPX_ERR_SHAPE;
  const size_t smem = (27 * Cout + kStemPix * (Cout + 1)) * sizeof(float);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(conv_stem_fprop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    attr = true;
  }
  const long long P = static_cast<long long>(Nimg) * H * W;
  conv_stem_fprop_kernel<<<nblocks(P, kStemPix, 148 * 4), kStemPix, smem, static_cast<cudaStream_t>(stream)>>>(
      x_nchw, w, bias, y_nhwc, Nimg, H, W, Cout, act, stat_sum, stat_sq);
  return LAUNCH_OK();
}

extern "C" int ddpx_conv_stem_wgrad(const float* dy_nhwc, const float* x_nchw, float* dw, int Nimg, int H, int W, int Cout,
                                    void* stream) {
  if (Cout > 256 || 256 % Cout) return DDPX_ERR_SHAPE;
  const size_t smem = static_cast<size_t>(3) * (H + 2) * (W + 2) * sizeof(float);
  if (smem > 48 * 1024) return DDPX_ERR_SHAPE;
  int grid = Nimg < 148 * 2 ? Nimg : 148 * 2;
  conv_stem_wgrad_kernel<<<grid, 256, smem, static_cast<cudaStream_t>(stream)>>>(dy_nhwc, x_nchw, dw, Nimg, H, W, Cout);
  return LAUNCH_OK();
}

// [batch][rows][cols] -> [batch][cols][rows];  NCHW->NHWC: rows = C, cols = H*W;  NHWC->NCHW: rows = H*W, cols = C
extern "C" int ddpx_transpose_batched(const float* in, float* out, int batch, int rows, int cols, void* stream) {
  dim3 grid((cols + 31) / 32, (rows + 31) / 32, batch), block(32, 8);
  if (grid.y > 65535 || grid.z > 65535) return DDPX_ERR_SHAPE;
  transpose_batched_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(in, out, rows, cols);
  return LAUNCH_OK();
}

extern "C" int ddpx_relu_fwd(const float* x, float* y, long long n, void* stream) {
  relu_fwd_kernel<<<nblocks(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, n);
  return LAUNCH_OK();
}
extern "C" int ddpx_relu_bwd(const float* y, const float* dy, float* dx, long long n, void* stream) {
  relu_bwd_kernel<<<nblocks(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(y, dy, dx, n);
  return LAUNCH_OK();
}
extern "C" int ddpx_maxpool2x2_fwd(const float* x, float* y, int Nimg, int H, int W, int C, void* stream) {
  if ((H | W) & 1) return DDPX_ERR_SHAPE;
  const long long total = static_cast<long long>(Nimg) * (H / 2) * (W / 2) * C;
  maxpool_fwd_kernel<<<nblocks(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, Nimg, H, W, C);
  return LAUNCH_OK();
}
extern "C" int ddpx_maxpool2x2_bwd(const float* x, const float* dy, float* dx, int Nimg, int H, int W, int C, void* stream) {
  if ((H | W) & 1) return DDPX_ERR_SHAPE;
  const long long total = static_cast<long long>(Nimg) * (H / 2) * (W / 2) * C;
  maxpool_bwd_kernel<<<nblocks(total, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, dy, dx, Nimg, H, W, C);
  return LAUNCH_OK();
}
extern "C" int ddpx_spatial_mean_fwd(const float* x, float* y, int Nimg, int HW, int C, void* stream) {
  spatial_mean_fwd_kernel<<<nblocks(static_cast<long long>(Nimg) * C, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(x, y, Nimg, HW, C);
  return LAUNCH_OK();
}
extern "C" int ddpx_spatial_mean_bwd(const float* dy, float* dx, int Nimg, int HW, int C, void* stream) {
  spatial_mean_bwd_kernel<<<nblocks(static_cast<long long>(Nimg) * HW * C, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(dy, dx, Nimg, HW, C);
  return LAUNCH_OK();
}
extern "C" int ddpx_colsum(const float* x, float* out, long long rows, int cols, int accumulate, void* stream) {
  int gy = static_cast<int>(rows / 64);
  if (gy < 1) gy = 1;
  if (gy > 256) gy = 256;
  if (gy > 1 && !accumulate) cudaMemsetAsync(out, 0, sizeof(float) * cols, static_cast<cudaStream_t>(stream));
  dim3 grid((cols + 127) / 128, gy);
  colsum_kernel<<<grid, 128, 0, static_cast<cudaStream_t>(stream)>>>(x, out, rows, cols, accumulate);
  return LAUNCH_OK();
}
extern "C" int ddpx_scale_copy(const float* src, float* dst, long long n, float scale, int accumulate, void* stream) {
  scale_copy_kernel<<<nblocks(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(src, dst, n, scale, accumulate);
  return LAUNCH_OK();
}
extern "C" int ddpx_zero(void* p, long long bytes, void* stream) {
  return cudaMemsetAsync(p, 0, static_cast<size_t>(bytes), static_cast<cudaStream_t>(stream)) == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

// ------------------------------------------------------------------------------------------------
// small / odd-shaped fp32 GEMM on CUDA cores (classifier N = 10, toy Linear(20,1): shapes whose row pitch is not a
// multiple of 16 bytes cannot be described to TMA).  C[m][n] = act(sum_k A(m,k) * B(n,k) + bias[n]) (+ C if accumulate),
// A(m,k) = A[m*sam + k*sak], B(n,k) = B[n*sbn + k*sbk].
// ------------------------------------------------------------------------------------------------
namespace {
__global__ void gemm_small_kernel(const float* __restrict__ A, const float* __restrict__ B, float* __restrict__ C, int M, int N,
                                  int K, long long sam, long long sak, long long sbn, long long sbk, long long ldc,
                                  const float* __restrict__ bias, int act, int accumulate, int k_per_split) {
  __shared__ float As[32][33];
  __shared__ float Bs[32][33];
  const int tx = threadIdx.x, ty = threadIdx.y;           // 32 x 8
  const int m0 = blockIdx.y * 32, n0 = blockIdx.x * 32;
  // split-K over blockIdx.z (few-tile problems such as the 512 -> 10 classifier would otherwise run 16 blocks through
  // 16 dependent load-sync-compute rounds): partial sums are added atomically into a zeroed / accumulated C
  const int kbeg = blockIdx.z * k_per_split;
  const int kend = min(K, kbeg + k_per_split);
  const bool split = gridDim.z > 1;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (int k0 = kbeg; k0 < kend; k0 += 32) {
    for (int j = ty; j < 32; j += 8) {
      const int m = m0 + j, n = n0 + j, k = k0 + tx;
      As[j][tx] = (m < M && k < kend) ? A[m * sam + k * sak] : 0.f;
      Bs[j][tx] = (n < N && k < kend) ? B[n * sbn + k * sbk] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 32; ++kk) {
      const float b = Bs[tx][kk];
#pragma unroll
      for (int i = 0; i < 4; ++i) acc[i] = fmaf(As[ty + 8 * i][kk], b, acc[i]);
    }
    __syncthreads();
  }
  const int n = n0 + tx;
  if (n < N) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int m = m0 + ty + 8 * i;
      if (m < M) {
        float v = acc[i] + ((bias != nullptr && blockIdx.z == 0) ? bias[n] : 0.f);
        if (act == 1) v = fmaxf(v, 0.f);
        float* c = C + m * ldc + n;
        if (split) atomicAdd(c, v);
        else *c = accumulate ? *c + v : v;
      }
    }
  }
}
}  // namespace

extern "C" int ddpx_gemm_small(const float* A, const float* B, float* C, int M, int N, int K, long long sam, long long sak,
                               long long sbn, long long sbk, long long ldc, const float* bias, int act, int accumulate,
                               void* stream) {
  dim3 grid((N + 31) / 32, (M + 31) / 32), block(32, 8);
  if (grid.y > 65535) return DDPX_ERR_SHAPE;
  int splits = 1;
  const int tiles = static_cast<int>(grid.x * grid.y);
  if (act == 0 && tiles < 148 && K >= 128) {
    splits = (2 * 148) / tiles;
    if (splits > K / 64) splits = K / 64;
    if (splits < 1) splits = 1;
  }
  int k_per_split = ((K + splits - 1) / splits + 31) / 32 * 32;
  splits = (K + k_per_split - 1) / k_per_split;
  if (splits > 1 && !accumulate) {      // rows of C are ldc apart: clear exactly the M x N window
    cudaError_t e = cudaMemset2DAsync(C, static_cast<size_t>(ldc) * sizeof(float), 0, static_cast<size_t>(N) * sizeof(float),
                                      static_cast<size_t>(M), static_cast<cudaStream_t>(stream));
    if (e != cudaSuccess) return DDPX_ERR_LAUNCH;
  }
  grid.z = splits;
  gemm_small_kernel<<<grid, block, 0, static_cast<cudaStream_t>(stream)>>>(A, B, C, M, N, K, sam, sak, sbn, sbk, ldc, bias, act,
                                                                          accumulate, k_per_split);
  return LAUNCH_OK();
}
