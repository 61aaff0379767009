// Block-level entry points of the transformer path: a whole GPT-2 block per direction in ONE C call.
//
// GPT-2 XL with 2048-token micro-batches was host-bound: 2033 native calls per micro-batch step, 51.9 ms of host enqueue
// time against 58.2 ms of step (profiles/raw/host_profile_xl.log) - ~25 us per call in torch.empty, ctypes marshaling and
// autograd.Function.apply.  A block here is one autograd node, two C calls (forward, backward) and a handful of
// allocations; every kernel inside is the framework's own (tcgen05 GEMMs with bias / GELU epilogues, fused causal
// attention, LayerNorm, GELU, residual add, column sums, casts).  Nothing here allocates; all buffers are the caller's.
// No counterpart in the reference (a CNN); serves the GPT-2 configurations of BASELINE.json.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>
#include "../../include/ddpx_kernels.h"

#define CHECK(call) do { int rc_ = (call); if (rc_ != DDPX_OK) return rc_; } while (0)

namespace {
typedef __nv_bfloat16 bf16;

// parameter order of a block (each bf16): LayerNorm 1, attention in / out projections, LayerNorm 2, MLP in / out
enum { LN1_W, LN1_B, ATTN_W, ATTN_B, PROJ_W, PROJ_B, LN2_W, LN2_B, FC_W, FC_B, PROJ2_W, PROJ2_B, kNumParams };
// activations a block keeps for its backward (bf16 unless noted)
enum { A_LN1, A_QKV, A_ATT, A_X1, A_LN2, A_FCPRE, A_FCACT, kNumActs };

// fp32 -> bf16 for up to 8 small vectors in one launch (bias and LayerNorm gradients of a block: eight tiny cast launches
// per block and direction before)
struct CastTable { const float* src[8]; bf16* dst[8]; int n[8]; int count; };
__global__ void cast_multi_kernel(const CastTable t) {
  const int seg = blockIdx.y;
  if (seg >= t.count) return;
  // compile-time indices only: a runtime index into the by-value table would make the compiler copy it to local memory
  const float* src = t.src[0];
  bf16* dst = t.dst[0];
  int n = t.n[0];
#pragma unroll
  for (int i = 1; i < 8; ++i)
    if (seg == i) { src = t.src[i]; dst = t.dst[i]; n = t.n[i]; }
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = __float2bfloat16(src[i]);
}

// DDPX_BLOCK_FUSE: bit mask of the elementwise passes folded into GEMM epilogues / LayerNorm backward (for A/B timing):
// 1 residual adds in the projections, 2 GELU with two outputs in the fc GEMM, 4 gelu' in the fc-activation data gradient,
// 8 residual-gradient add in LayerNorm backward.  Default: all.
int fuse_mask() {
  static int m = -1;
  if (m < 0) { const char* e = getenv("DDPX_BLOCK_FUSE"); m = e ? atoi(e) : 15; }
  return m;
}

// y[M,N] = act(x[M,K] W[N,K]^T + b)
int linear_fwd(const void* x, const void* W, const void* b, void* y, int M, int N, int K, int act, void* stream) {
  return ddpx_gemm_bias(x, W, y, M, N, K, K, K, N, 0, 0, 1, 1, b, 1, act, 1, 0, stream);
}

// Backward of y = x W^T + b for bf16 tensors: dx[M,K] = d W (skipped when dx is null), dW[N,K] = d^T x, db32[N] = colsum(d) in fp32.
// ws32: fp32 scratch of at least N*K elements that accumulates a split-K dW and must be ALL ZERO on entry (the GEMM's
// in-kernel fix-up converts the finished sums to bf16 and leaves it zero again - no memset, no cast launch); db32: N floats
// for the bias sums, OUTSIDE that zero-invariant area (layers of different sizes share ws32 from offset 0);
// cnt: the fix-up's arrival counters (all zero on entry and on exit).
// gelu_pre (optional, [M,K]): x was gelu(gelu_pre), and dx comes out multiplied by gelu'(gelu_pre) - the gradient of the
// pre-activation - in the data-gradient GEMM's epilogue.
int linear_bwd(const void* d, const void* x, const void* W, void* dx, void* dW, float* ws32, float* db32,
               unsigned int* cnt, int M, int N, int K, void* stream, const void* gelu_pre = nullptr) {
  if (dx != nullptr)                                                                            // d[M,N] @ W[N,K]
    CHECK(ddpx_gemm_fused(d, W, dx, M, K, N, N, K, K, 0, 1, nullptr, 0, 0, gelu_pre, gelu_pre != nullptr ? 2 : 0, nullptr, stream));
  // A weight gradient has few output tiles and a long reduction (the tokens): when 256x256 tiles cannot fill the 74 CTA
  // pairs, split the reduction and accumulate in fp32 (then cast) instead of shrinking the tiles.
  const int tiles = ((N + 255) / 256) * ((K + 255) / 256);
  int split = 1;
  if (!(tiles >= 60 || M < 2048)) {
    split = 74 / tiles;
    if (split > M / 1024) split = M / 1024;
    if (split < 1) split = 1;
  }
  if (split > 1 && K % 32 == 0) {
    CHECK(ddpx_gemm_splitk_bf16(d, x, dW, N, K, M, N, K, K, 1, 1, split, ws32, cnt, stream));
  } else {
    CHECK(ddpx_gemm(d, x, dW, N, K, M, N, K, K, 1, 1, 1, 1, nullptr, 0, 1, 0, stream));
  }
  // bias gradient: fp32 column sums; the block casts all its small gradients at the end.  A null db32 leaves them to the
  // caller (the GPT-2 block sums the four of them in one launch)
  return db32 != nullptr ? ddpx_colsum_bf16(d, db32, M, N, stream) : DDPX_OK;
}
}  // namespace

// Forward of one GPT-2 block on x [B*T, C] (bf16):
//   x1 = x + proj(attention(qkv(LN1(x)))),   out = x1 + proj2(gelu(fc(LN2(x1))))
// params[12]: see the enum above.  acts[7]: LN1 out [M,C], qkv [M,3C], attention out [M,C], x1 [M,C], LN2 out [M,C],
// fc pre-activation [M,4C], fc activation [M,4C] - all kept for the backward.  stats: fp32 [4*M + B*H*Tp] = LN1 mean /
// rstd, LN2 mean / rstd, attention log-sum-exp (Tp = T rounded up to 128).  tmp: bf16 [M,C] scratch (used only when the residual adds run as separate passes).
extern "C" int ddpx_gpt2_block_fwd(const void* x, void* const* params, void* const* acts, float* stats, void* tmp, void* out,
                                   int B, int T, int C, int H, float eps, void* stream) {
  if (C != H * 64 || C % 8) return DDPX_ERR_SHAPE;
  const int M = B * T;
  const long long MC = static_cast<long long>(M) * C;
  float* ln1_mean = stats; float* ln1_rstd = stats + M;
  float* ln2_mean = stats + 2 * M; float* ln2_rstd = stats + 3 * M;
  float* lse2 = stats + 4 * static_cast<size_t>(M);
  CHECK(ddpx_layernorm_fwd(x, params[LN1_W], params[LN1_B], acts[A_LN1], ln1_mean, ln1_rstd, M, C, eps, 1, stream));
  CHECK(linear_fwd(acts[A_LN1], params[ATTN_W], params[ATTN_B], acts[A_QKV], M, 3 * C, C, 0, stream));
  CHECK(ddpx_attn_fwd(acts[A_QKV], acts[A_ATT], lse2, B, T, H, 0.125f, stream));
  // the residual connections ride in the projections' epilogues
  const int fm = fuse_mask();
  if (fm & 1) {
    CHECK(ddpx_gemm_fused(acts[A_ATT], params[PROJ_W], acts[A_X1], M, C, C, C, C, C, 0, 0, params[PROJ_B], 1, 0, x, 1, nullptr,
                          stream));
  } else {
    CHECK(linear_fwd(acts[A_ATT], params[PROJ_W], params[PROJ_B], tmp, M, C, C, 0, stream));
    CHECK(ddpx_add(x, tmp, acts[A_X1], MC, 1, stream));
  }
  CHECK(ddpx_layernorm_fwd(acts[A_X1], params[LN2_W], params[LN2_B], acts[A_LN2], ln2_mean, ln2_rstd, M, C, eps, 1, stream));
  // GELU in the epilogue; the pre-activation (needed by the backward) is the kernel's second output
  if (fm & 2) {
    CHECK(ddpx_gemm_fused(acts[A_LN2], params[FC_W], acts[A_FCACT], M, 4 * C, C, C, C, 4 * C, 0, 0, params[FC_B], 1, 2, nullptr, 0,
                          acts[A_FCPRE], stream));
  } else {
    CHECK(linear_fwd(acts[A_LN2], params[FC_W], params[FC_B], acts[A_FCPRE], M, 4 * C, C, 0, stream));
    CHECK(ddpx_gelu_fwd(acts[A_FCPRE], acts[A_FCACT], 4 * MC, 1, stream));
  }
  if (fm & 1)
    return ddpx_gemm_fused(acts[A_FCACT], params[PROJ2_W], out, M, C, 4 * C, 4 * C, 4 * C, C, 0, 0, params[PROJ2_B], 1, 0,
                           acts[A_X1], 1, nullptr, stream);
  CHECK(linear_fwd(acts[A_FCACT], params[PROJ2_W], params[PROJ2_B], tmp, M, C, 4 * C, 0, stream));
  return ddpx_add(acts[A_X1], tmp, out, MC, 1, stream);
}

// Backward of ddpx_gpt2_block_fwd.  dout [M,C] is the gradient of the block output; x, acts, stats are the forward's.
// grads[12] receive the parameter gradients (bf16, same order as params), dx [M,C] the input gradient.
// wsb: bf16 scratch [M, 9C];  ws32: fp32 scratch [4C*C + 13C + B*H*Tp + ddpx_layernorm_bwd_scratch_floats(M, C)] whose first 4C*C elements must be all zero on entry
// (they are left all zero);  cnt: uint32 [16 * (4C*C/8192 + 64)] arrival counters, all zero on entry and on exit.
extern "C" int ddpx_gpt2_block_bwd(const void* x, const void* dout, void* const* params, void* const* acts, const float* stats,
                                   void* const* grads, void* dx, void* wsb, float* ws32, unsigned int* cnt, int B, int T,
                                   int C, int H, void* stream) {
  if (C != H * 64 || C % 8) return DDPX_ERR_SHAPE;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int M = B * T;
  const long long MC = static_cast<long long>(M) * C;
  const float* ln1_mean = stats; const float* ln1_rstd = stats + M;
  const float* ln2_mean = stats + 2 * M; const float* ln2_rstd = stats + 3 * M;
  const float* lse2 = stats + 4 * static_cast<size_t>(M);
  bf16* d_fc = static_cast<bf16*>(wsb);                 // [M,4C]
  bf16* d_a = d_fc + 4 * MC;                            // [M,C]
  bf16* d_b = d_a + MC;                                 // [M,C]
  bf16* dqkv = d_b + MC;                                // [M,3C]
  // small fp32 results behind the zero-invariant dW area: bias sums of the four linears, then the LayerNorm accumulators
  float* small32 = ws32 + static_cast<size_t>(4) * C * C;
  float* b_proj2 = small32, *b_fc = small32 + C, *b_proj = small32 + 5 * C, *b_attn = small32 + 6 * C;
  float* ln2acc = small32 + 9 * C;                                   // [2C] dgamma, dbeta of LayerNorm 2
  float* ln1acc = small32 + 11 * C;                                  // [2C] of LayerNorm 1
  float* delta = small32 + 13 * C;                                   // attention backward scratch [B*H*Tp]
  // per-block partial sums of the LayerNorm backward ([<= 296][2C], behind delta; lse_pad = B*H*Tp)
  const int Tp = (T + 127) / 128 * 128;
  float* ln_scratch = delta + static_cast<size_t>(B) * H * Tp;

  // ---- MLP: out = x1 + proj2(gelu(fc(LN2(x1))))
  const int fm = fuse_mask();
  CHECK(linear_bwd(dout, acts[A_FCACT], params[PROJ2_W], d_fc, grads[PROJ2_W], ws32, nullptr, cnt, M, C, 4 * C, stream,
                   (fm & 4) ? acts[A_FCPRE] : nullptr));
  if (!(fm & 4)) CHECK(ddpx_gelu_bwd(acts[A_FCPRE], d_fc, d_fc, 4 * MC, 1, stream));
  CHECK(linear_bwd(d_fc, acts[A_LN2], params[FC_W], d_a, grads[FC_W], ws32, nullptr, cnt, M, 4 * C, C, stream));
  // d_x1 = through LN2 + the residual branch (added inside the LayerNorm kernel)
  CHECK(ddpx_layernorm_bwd_res(acts[A_X1], d_a, params[LN2_W], ln2_mean, ln2_rstd, d_b, ln2acc, ln2acc + C, M, C, 1, ln_scratch,
                               (fm & 8) ? dout : nullptr, stream));
  if (!(fm & 8)) CHECK(ddpx_add(d_b, dout, d_b, MC, 1, stream));
  // ---- attention: x1 = x + proj(attention(qkv(LN1(x))))
  CHECK(linear_bwd(d_b, acts[A_ATT], params[PROJ_W], d_a, grads[PROJ_W], ws32, nullptr, cnt, M, C, C, stream));
  CHECK(ddpx_attn_bwd(acts[A_QKV], acts[A_ATT], d_a, lse2, delta, dqkv, B, T, H, 0.125f, stream));
  CHECK(linear_bwd(dqkv, acts[A_LN1], params[ATTN_W], d_a, grads[ATTN_W], ws32, nullptr, cnt, M, 3 * C, C, stream));
  CHECK(ddpx_layernorm_bwd_res(x, d_a, params[LN1_W], ln1_mean, ln1_rstd, dx, ln1acc, ln1acc + C, M, C, 1, ln_scratch,
                               (fm & 8) ? d_b : nullptr, stream));
  // ---- the four bias gradients: column sums of dout, d_fc, d_x1, dqkv (all still intact) in one launch over one
  // cleared stretch of the scratch (the first 9C floats of small32), instead of a memset + kernel pair per linear layer
  {
    if (cudaMemsetAsync(small32, 0, sizeof(float) * 9 * C, s) != cudaSuccess) return DDPX_ERR_LAUNCH;
    const void* xs[4] = {dout, d_fc, d_b, dqkv};
    float* outs[4] = {b_proj2, b_fc, b_proj, b_attn};
    const int cs[4] = {C, 4 * C, C, 3 * C};
    CHECK(ddpx_colsum_bf16_multi(4, xs, outs, M, cs, stream));
  }
  // ---- every small gradient fp32 -> bf16 in one launch
  CastTable t;
  const float* srcs[8] = {b_proj2, b_fc, b_proj, b_attn, ln2acc, ln2acc + C, ln1acc, ln1acc + C};
  const int dsts[8] = {PROJ2_B, FC_B, PROJ_B, ATTN_B, LN2_W, LN2_B, LN1_W, LN1_B};
  const int lens[8] = {C, 4 * C, C, 3 * C, C, C, C, C};
  for (int i = 0; i < 8; ++i) { t.src[i] = srcs[i]; t.dst[i] = static_cast<bf16*>(grads[dsts[i]]); t.n[i] = lens[i]; }
  t.count = 8;
  cast_multi_kernel<<<dim3((4 * C + 255) / 256, 8), 256, 0, s>>>(t);
  if (cudaGetLastError() != cudaSuccess) return DDPX_ERR_LAUNCH;
  return (fm & 8) ? DDPX_OK : ddpx_add(dx, d_b, dx, MC, 1, stream);
}
