// RGB stem (Cin = 3, K = 27) on tensor cores: gather the 3x3x3 patches into 32-float rows once, then the forward is a
// 1x1 tcgen05 convolution over them (ddpx_conv1x1_fprop: NHWC store + BN statistics in the epilogue) and the weight
// gradient is one split-K tcgen05 GEMM dy^T x patches (ddpx_gemm, both operands MN-major) instead of a CUDA-core loop.
// conv0 of the reference's VGG (multigpu.py:53): "K = 27 is tiny => pad K to 32, memory-bound".
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/ddpx_kernels.h"

namespace {

// patches[p][k], k = ci*9 + r*3 + s for k < 27, zero for 27..31; input NCHW
__global__ void im2col_stem_kernel(const float* __restrict__ x, float* __restrict__ patches, int Nimg, int H, int W) {
  const long long P = static_cast<long long>(Nimg) * H * W;
  for (long long p = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; p < P;
       p += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int wq = static_cast<int>(p % W);
    const long long t = p / W;
    const int hq = static_cast<int>(t % H);
    const long long n = t / H;
    float v[32];
#pragma unroll
    for (int ci = 0; ci < 3; ++ci)
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int s = 0; s < 3; ++s) {
          const int hh = hq + r - 1, ww = wq + s - 1;
          v[ci * 9 + r * 3 + s] = (hh >= 0 && hh < H && ww >= 0 && ww < W) ? __ldg(x + ((n * 3 + ci) * H + hh) * W + ww) : 0.f;
        }
#pragma unroll
    for (int k = 27; k < 32; ++k) v[k] = 0.f;
    float4* o = reinterpret_cast<float4*>(patches + p * 32);
#pragma unroll
    for (int k = 0; k < 8; ++k) o[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
  }
}

// w [Cout][27] -> wp [Cout][32] (zero padded)
__global__ void pack_stem_w_kernel(const float* __restrict__ w, float* __restrict__ wp, int Cout) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < Cout * 32) {
    const int co = i >> 5, k = i & 31;
    wp[i] = k < 27 ? w[co * 27 + k] : 0.f;
  }
}

}  // namespace

extern "C" int ddpx_im2col_stem(const float* x_nchw, float* patches, int Nimg, int H, int W, void* stream) {
  const long long P = static_cast<long long>(Nimg) * H * W;
  long long b = (P + 127) / 128;
  if (b > 148 * 32) b = 148 * 32;
  im2col_stem_kernel<<<static_cast<int>(b), 128, 0, static_cast<cudaStream_t>(stream)>>>(x_nchw, patches, Nimg, H, W);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" int ddpx_pack_stem_weights(const float* w, float* wp, int Cout, void* stream) {
  pack_stem_w_kernel<<<(Cout * 32 + 255) / 256, 256, 0, static_cast<cudaStream_t>(stream)>>>(w, wp, Cout);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}
