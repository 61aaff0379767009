// Block-level entry points: one C call per fused layer group and direction.
//
// A VGG training step used to cross the Python / ctypes boundary ~80 times (torch.empty + marshaling + launch per
// kernel, ~25 us of host time each: profiles/probe_findings.md section 9); these entry points enqueue every kernel of a
// conv-BN-ReLU(-pool) group (multigpu.py:53-59 of the reference) in one call, with every workspace handed in by the caller.
// Nothing here allocates.
#include <cuda_runtime.h>
#include "../../include/ddpx_kernels.h"

#define CHECK(call) do { int rc_ = (call); if (rc_ != DDPX_OK) return rc_; } while (0)

// Forward of conv3x3(pad 1, no bias) -> BatchNorm2d -> [ReLU] -> [MaxPool 2x2].
//   x         NHWC input [N,H,W,Cin]; for the RGB stem (Cin == 3) the NCHW-contiguous image
//   wf        packed fprop weights from ddpx_pack_conv_weights_multi ([9][Cout][Cin]; stem: [Cout][32])
//   stats     [2][Cout] sum / sum-of-squares accumulators, ZERO on entry when training (the multi-pack launch clears them)
//   y         [N,H,W,Cout] raw convolution output (kept for the backward)
//   out       [N,Ho,Wo,Cout] block output
//   mean_rstd [2][Cout] output for the backward
//   patches   stem only: [N,H,W,32] gathered 3x3x3 patches (kept for the backward)
extern "C" int ddpx_conv_bn_fwd(const float* x, const float* wf, const float* gamma, const float* beta, float* running_mean,
                                float* running_var, long long* num_batches_tracked, float momentum, float eps, int Nimg,
                                int H, int W, int Cin, int Cout, int relu, int pool, int training, float* stats, float* y,
                                float* out, float* mean_rstd, float* patches, void* stream) {
  if (Cout <= 0 || Cout % 32) return DDPX_ERR_SHAPE;
  if (pool && ((H | W) & 1)) return DDPX_ERR_SHAPE;
  float* ssum = training ? stats : nullptr;
  float* ssq = training ? stats + Cout : nullptr;
  if (Cin == 3) {
    if (patches == nullptr) return DDPX_ERR_ARG;
    CHECK(ddpx_im2col_stem(x, patches, Nimg, H, W, stream));
    CHECK(ddpx_conv1x1_fprop(patches, wf, y, Nimg, H, W, 32, Cout, nullptr, 0, ssum, ssq, stream));
  } else {
    CHECK(ddpx_conv3x3_fprop(x, wf, y, Nimg, H, W, Cin, Cout, nullptr, 0, ssum, ssq, stream));
  }
  return ddpx_bn_apply_stats(y, out, ssum, ssq, static_cast<float>(Nimg) * H * W, eps, momentum, gamma, beta, running_mean,
                             running_var, num_batches_tracked, mean_rstd, training, Nimg, H, W, Cout, relu, pool, stream);
}

// Backward of the same group.
//   y, mean_rstd   saved by the forward;  dout NHWC gradient of the block output;  xin = NHWC input (stem: the patches)
//   wd             packed dgrad weights (null when no input gradient is wanted)
//   sums           [2][Cout] output: sums[0] = dbeta, sums[1] = dgamma (cleared here)
//   dconv          [N,H,W,Cout] scratch: gradient at the convolution output
//   staging        persistent [9][Cout][Cin] accumulation buffer, all-zero on entry and left all-zero (not used by the stem)
//   dw             [Cout][Cin][3][3] output;  dx [N,H,W,Cin] output or null
extern "C" int ddpx_conv_bn_bwd(const float* y, const float* dout, const float* xin, const float* wd, const float* mean_rstd,
                                const float* gamma, const float* beta, int Nimg, int H, int W, int Cin, int Cout, int relu,
                                int pool, int training, float* sums, float* dconv, float* staging, float* dw, float* dx,
                                void* stream) {
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (Cout <= 0 || Cout % 32) return DDPX_ERR_SHAPE;
  if (cudaMemsetAsync(sums, 0, sizeof(float) * 2 * Cout, s) != cudaSuccess) return DDPX_ERR_LAUNCH;
  const float* mean = mean_rstd;
  const float* rstd = mean_rstd + Cout;
  CHECK(ddpx_bn_bwd_reduce(y, dout, mean, rstd, gamma, beta, Nimg, H, W, Cout, relu, pool, sums, sums + Cout, stream));
  CHECK(ddpx_bn_bwd_apply(y, dout, mean, rstd, gamma, beta, sums, sums + Cout, Nimg, H, W, Cout, relu, pool, training, dconv,
                          stream));
  if (Cin == 3) {
    // dw[Cout][27] = dconv^T[Cout][P] x patches[P][0:27]: both operands MN-major, two K-slices per SM, fp32 atomics
    if (cudaMemsetAsync(dw, 0, sizeof(float) * Cout * 27, s) != cudaSuccess) return DDPX_ERR_LAUNCH;
    return ddpx_gemm(dconv, xin, dw, Cout, 27, Nimg * H * W, Cout, 32, 27, 1, 1, 0, 0, nullptr, 0, 296, 1, stream);
  }
  if (staging == nullptr) return DDPX_ERR_ARG;
  CHECK(ddpx_conv3x3_wgrad_packed(dconv, xin, staging, Nimg, H, W, Cin, Cout, 0, stream));
  CHECK(ddpx_unpack_conv_wgrad_rezero(staging, dw, Cout, Cin, stream));
  if (dx != nullptr) {
    if (wd == nullptr) return DDPX_ERR_ARG;
    CHECK(ddpx_conv3x3_fprop(dconv, wd, dx, Nimg, H, W, Cout, Cin, nullptr, 0, nullptr, nullptr, stream));
  }
  return DDPX_OK;
}
