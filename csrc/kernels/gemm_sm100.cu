// tcgen05 / TMEM / TMA tensor-core core for sm_100a.
//
// One persistent, warp-specialised kernel template serves three problem shapes:
//   MODE_GEMM   D[M,N] = op(A) * op(B)                  (Linear fwd / dgrad / wgrad; K- or MN-major operands)
//   MODE_FPROP  3x3 stride-1 pad-1 convolution as an implicit GEMM over NHWC activations
//               (also used for the data-gradient with tap-flipped, transposed weights)
//   MODE_WGRAD  3x3 convolution weight-gradient: K = pixels, both operands MN-major
//
// Reference parity: these replace the cuDNN / cuBLAS calls behind nn.Conv2d / nn.Linear in
// /root/reference/multigpu.py:53,62 (W13-W15, W20).
//
// Roles (384 threads): warp 0 = TMA producer, warp 1 = MMA issuer (one lane), warp 2 = TMEM allocator,
// warps 4-11 = epilogue (TMEM -> registers -> global; two warps per TMEM lane quarter, each draining half the columns). Accumulators are double-buffered in TMEM so the
// epilogue of tile i overlaps the MMAs of tile i+1. Operand tiles are 128-byte-swizzled in shared memory
// exactly as TMA writes them; padding of the convolution comes from TMA out-of-bounds zero fill with
// signed start coordinates (no im2col buffer is ever materialised).
#include "sm100_ptx.cuh"
#include "../../include/ddpx_kernels.h"

#include <cudaTypedefs.h>
#include <stdio.h>
#include <stdlib.h>
#include <mutex>
#include <string.h>
#include "tmap.hpp"

namespace ddpx {

enum { MODE_GEMM = 0, MODE_FPROP = 1, MODE_WGRAD = 2 };

constexpr int kBlockM = 128;          // TMEM lanes / UMMA M
constexpr int kStageABytes = 16384;   // 128 rows x 128 B
constexpr int kNumThreads = 384;     // 4 role warps (TMA, MMA, TMEM alloc, spare) + 8 epilogue warps
constexpr int kMaxStages = 8;

struct TcParams {
  int M, N;                 // logical extent of the output (rows <-> TMEM lanes, cols <-> TMEM columns)
  int m_blocks, n_blocks;
  int block_n;              // 32..256, multiple of 32
  int k_steps_total;        // GEMM: ceil(K/BK); FPROP: 9*cin_blocks; WGRAD: pixels/32
  int k_steps_per_split;
  int num_splits;
  int num_stages;
  int a_mn, b_mn;           // operand majors (1 = MN-major / "transposed")
  // batching (MODE_GEMM): problem index b = b1 + batch1 * b2; operands addressed through 4-D tensor maps
  int batch1, batch2;
  long long out_bs1, out_bs2;   // output strides (elements) of the two batch dimensions
  int causal;               // 1: skip output tiles entirely above the diagonal (S = Q K^T)
                            // 2: stop the K loop at the diagonal (O = P V with P lower-triangular)
  // convolution geometry
  int H, W, cin_blocks;
  int box_w, box_h, box_n;
  int cin_total;
  int conv_taps;            // FPROP: 9 (3x3) or 1 (1x1 over pre-gathered patches)
  int tpi;                  // filter taps per work item: WGRAD handles one filter row (3 taps) per item so the dY tile
                            // is loaded once for three accumulators; 1 everywhere else
  int acc_stages;           // TMEM accumulator buffers in flight (2 when 2*tpi*block_n <= 512 columns, else 1)
  // epilogue
  void* out;
  long long ldo;
  int out_bf16;
  int packed;               // WGRAD: output is the tap-major staging layout [9][Cout][Cin] (vector reductions)
  int atomic;               // 1 = accumulate with red.global.add.f32 (split-K / wgrad)
  float alpha;              // GEMM: D = act(alpha * A*B + bias)
  int a_big, b_big;         // GEMM: MN-major operand map carries the 128-byte chunk index as a 5th dimension (one TMA per tile)
  int big_box;              // WGRAD: operand maps have the 32-channel chunk as outermost dim (one TMA per tile)
  int st_vec;               // GEMM: output rows are 16-byte aligned -> epilogue stages 32x32 chunks through shared
                            // memory and stores whole 64/128-byte row segments (coalesced) instead of 16 B per lane
  int act;                  // 0 none, 1 relu, 2 gelu(tanh)
  const float* bias;        // GEMM: per output column; FPROP: per output channel (row)
  int bias_bf16;            // GEMM: `bias` points at bf16 values (transformer layers keep their biases in bf16)
  // split-K with an in-kernel fix-up (GEMM): partial tiles are reduced into the fp32 workspace `fix_ws` [M][N] with vector
  // reductions; every epilogue warp counts its arrivals per sub-block in `fix_cnt`, and the warp that arrives last
  // converts the finished sub-block to bf16 into `out` and leaves workspace and counter at zero for the next call
  float* fix_ws;
  unsigned int* fix_cnt;
  // GEMM, bf16 output: an elementwise pass folded into the store.  `aux` is a bf16 tensor [M][ldo]:
  //   aux_op 1  D = bf16(bf16(act(...)) + aux)              (residual connection)
  //   aux_op 2  D = bf16(bf16(...) * gelu'(aux))            (aux = the pre-activation: backward through GELU)
  // - the numerics of the separate pass without the pass.  out2 (with act = gelu): the pre-activation is stored there as
  // well, so the forward keeps both tensors the backward needs from one kernel.
  const __nv_bfloat16* aux;
  int aux_op;
  __nv_bfloat16* out2;
  float* stat_sum;          // FPROP: per-channel sum / sum-of-squares of the raw conv output
  float* stat_sq;
};

// eight bf16 values + eight bf16 values, each sum rounded once
__device__ __forceinline__ uint4 add_bf16x8(uint4 a, uint4 b) {
  uint4 r;
  const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&b);
  __nv_bfloat162* pr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = __bfloat1622float2(pa[i]), y = __bfloat1622float2(pb[i]);
    pr[i] = __floats2bfloat162_rn(x.x + y.x, x.y + y.y);
  }
  return r;
}

__device__ __forceinline__ float tanh_fast(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));     // one MUFU op; error below bf16 resolution
  return y;
}
__device__ __forceinline__ float gelu_tanh(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  float u = k0 * (x + k1 * x * x * x);
  return 0.5f * x * (1.f + tanh_fast(u));
}
__device__ __forceinline__ float gelu_tanh_grad(float x) {
  const float k0 = 0.7978845608028654f, k1 = 0.044715f;
  const float t = tanh_fast(k0 * (x + k1 * x * x * x));
  return 0.5f * (1.f + t) + 0.5f * x * (1.f - t * t) * k0 * (1.f + 3.f * k1 * x * x);
}
// eight bf16 values times gelu'(eight bf16 pre-activations)
__device__ __forceinline__ uint4 mul_gelu_grad_bf16x8(uint4 a, uint4 pre) {
  uint4 r;
  const __nv_bfloat162* pa = reinterpret_cast<const __nv_bfloat162*>(&a);
  const __nv_bfloat162* pb = reinterpret_cast<const __nv_bfloat162*>(&pre);
  __nv_bfloat162* pr = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x = __bfloat1622float2(pa[i]), y = __bfloat1622float2(pb[i]);
    pr[i] = __floats2bfloat162_rn(x.x * gelu_tanh_grad(y.x), x.y * gelu_tanh_grad(y.y));
  }
  return r;
}

// kCG = 1: one CTA per tile (128 rows).  kCG = 2: a CTA pair (cluster of 2 on one TPC) per 256-row tile with
// tcgen05 cta_group::2 - each CTA stages its own 128 rows of A and HALF of the B tile, so a k-step costs 32 KB instead
// of 48 KB of L2->SM traffic per CTA and the same shared memory holds 6 stages instead of 4 (ncu on the 1-CTA kernel:
// tensor pipe 58 % active - profiles/ncu_findings.md section 1; the larger limiter turned out to be the producer thread,
// profiles/probe_findings.md section 3).
// Register budget: the convolution flavours are capped at 96 registers (no spills) so that CTAs of other streams -
// the fused all-reduce of csrc/runtime/p2p.cu above all - find room on the same SM next to this persistent kernel;
// the GEMM flavour keeps the staged epilogue's working set in registers.
template <int kMode, bool kBF16, int kCG>
__global__ void __launch_bounds__(kNumThreads, 1) __maxnreg__(kMode == MODE_GEMM ? 168 : 96)
tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // carve shared memory: [stages x (A | B)] | barriers | tmem ptr
  const uint32_t cta_rank = (kCG == 2) ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;
  const int bn_local = p.block_n / kCG;                 // B rows staged by this CTA
  const uint32_t tap_b_bytes = static_cast<uint32_t>(bn_local) * 128u;
  const uint32_t stage_b_bytes = tap_b_bytes * static_cast<uint32_t>(p.tpi);
  const uint32_t acc_cols = static_cast<uint32_t>(p.block_n * p.tpi);
  const uint32_t stage_bytes = kStageABytes + stage_b_bytes;
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(p.num_stages) * stage_bytes);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* tfull_bar = empty_bar + kMaxStages;   // [2]
  uint64_t* tempty_bar = tfull_bar + 2;           // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);
  uint8_t* epi_stage = reinterpret_cast<uint8_t*>(full_bar) + 256;     // 8 warps x 4 KB (MODE_GEMM only)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  constexpr int kElt = kBF16 ? 2 : 4;
  constexpr int kBK = 128 / kElt;          // K elements per stage for K-major operands (= K rows for MN-major)
  constexpr int kUmmaK = 32 / kElt;        // K per MMA instruction
  constexpr int kChunk = 128 / kElt;       // MN elements per 128-byte swizzle row

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.num_stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&tfull_bar[s], 1);
      mbar_init(&tempty_bar[s], 8 * kCG);       // the leader's barrier also collects the peer CTA's epilogue warps
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    if constexpr (kCG == 2) tmem_alloc_2sm(tmem_slot, 512); else tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  if constexpr (kCG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  const int tiles_mn = p.m_blocks * p.n_blocks;
  // "tap" doubles as the batch index in MODE_GEMM
  const int taps = (kMode == MODE_WGRAD) ? (9 + p.tpi - 1) / p.tpi : (kMode == MODE_GEMM ? p.batch1 * p.batch2 : 1);
  const int num_items = tiles_mn * taps * p.num_splits;

  auto decode = [&](int item, int& m_blk, int& n_blk, int& tap, int& split) {
    m_blk = item % p.m_blocks;
    int t = item / p.m_blocks;
    n_blk = t % p.n_blocks;
    t /= p.n_blocks;
    tap = t % taps;
    split = t / taps;
  };
  constexpr int kTileM = kBlockM * kCG;       // rows per work item (p.m_blocks counts tiles of this height)
  auto ksteps_of = [&](int split, int m_blk) {
    int k0 = split * p.k_steps_per_split;
    int k1 = min(k0 + p.k_steps_per_split, p.k_steps_total);
    if (kMode == MODE_GEMM && p.causal == 2) k1 = min(k1, ((m_blk + 1) * kTileM + kBK - 1) / kBK);
    return max(k1 - k0, 0);
  };
  auto skipped = [&](int m_blk, int n_blk) {
    return kMode == MODE_GEMM && p.causal == 1 && n_blk * p.block_n >= (m_blk + 1) * kTileM;
  };
  const int first_item = blockIdx.x / kCG;    // work is dealt to clusters
  const int item_stride = gridDim.x / kCG;

  auto ld4 = [&](void* dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2, int c3) {
    if constexpr (kCG == 2) tma_load_4d_2sm(dst, m, bar, c0, c1, c2, c3); else tma_load_4d(dst, m, bar, c0, c1, c2, c3);
  };
  if (warp == 0) {
    // ======================================= TMA producer =======================================
    // (entered through elect.sync, not `lane == 0`: see elect_one() in sm100_ptx.cuh)
    if (elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int item = first_item; item < num_items; item += item_stride) {
        int m_blk, n_blk, tap, split;
        decode(item, m_blk, n_blk, tap, split);
        const int m0 = m_blk * kTileM + static_cast<int>(cta_rank) * kBlockM;   // this CTA's 128 rows of A
        const int n0 = n_blk * p.block_n + static_cast<int>(cta_rank) * bn_local; // this CTA's part of the B tile
        const int kbeg = split * p.k_steps_per_split;
        const int nk = ksteps_of(split, m_blk);
        if (skipped(m_blk, n_blk)) continue;
        const int b1 = (kMode == MODE_GEMM) ? tap % p.batch1 : 0, b2 = (kMode == MODE_GEMM) ? tap / p.batch1 : 0;
        // The producer is ONE thread: every instruction in its k-loop is latency-exposed, and with 235 of them (four
        // integer divisions among them) it could not issue a k-step per 528 tensor-pipe cycles.  All divisions are
        // done here, once per item; the loop only increments.
        int cb = 0, t9 = 0, fr = 1, fs = 1;                 // FPROP: input-channel block, tap, filter row / column
        int img = 0, h0 = 0, p0 = 0;                        // FPROP / WGRAD: image and first row of the pixel tile
        if constexpr (kMode == MODE_FPROP) {
          const int hw = p.H * p.W;
          img = n0 / hw;
          h0 = (n0 - img * hw) / p.W;
          t9 = kbeg / p.cin_blocks;
          cb = kbeg - t9 * p.cin_blocks;
          if (p.conv_taps != 1) { fr = t9 / 3; fs = t9 - fr * 3; }
        } else if constexpr (kMode == MODE_WGRAD) {
          const int hw = p.H * p.W;
          p0 = kbeg * 32;
          img = p0 / hw;
          h0 = (p0 - img * hw) / p.W;
        }
        const int tap0 = tap * p.tpi;                                   // WGRAD: first tap of the group
        const int nt = (kMode == MODE_WGRAD) ? min(p.tpi, 9 - tap0) : 1; // taps in this group (the last may be short)
        uint32_t tx_bytes = stage_bytes * kCG;                          // both CTAs' bytes land on the leader's barrier
        if constexpr (kMode == MODE_WGRAD) tx_bytes = (kStageABytes + tap_b_bytes * static_cast<uint32_t>(nt)) * kCG;
        for (int i = 0; i < nk; ++i) {
          mbar_wait(&empty_bar[stage], phase ^ 1u, 0x100u + stage);
          uint8_t* sa = smem + static_cast<size_t>(stage) * stage_bytes;
          uint8_t* sb = sa + kStageABytes;
          if (leader) mbar_arrive_expect_tx(&full_bar[stage], tx_bytes);
          if constexpr (kMode == MODE_GEMM) {
            const int k0 = (kbeg + i) * kBK;
            // MN-major operands: the tile is kBlockM/kChunk (or bn_local/kChunk) swizzle tiles of [kBK rows][128 B].  With
            // the chunk index as an extra tensor-map dimension ONE instruction fetches them all; every TMA instruction
            // costs the single producer thread ~100 cycles (profiles/probe_findings.md)
            if (!p.a_mn) {
              ld4(sa, &tmA, &full_bar[stage], k0, m0, b1, b2);
            } else if (p.a_big) {
              if constexpr (kCG == 2) tma_load_5d_2sm(sa, &tmA, &full_bar[stage], 0, k0, m0 / kChunk, b1, b2);
              else tma_load_5d(sa, &tmA, &full_bar[stage], 0, k0, m0 / kChunk, b1, b2);
            } else {
              for (int j = 0; j < kBlockM / kChunk; ++j)
                ld4(sa + j * (kBK * 128), &tmA, &full_bar[stage], m0 + j * kChunk, k0, b1, b2);
            }
            if (!p.b_mn) {
              ld4(sb, &tmB, &full_bar[stage], k0, n0, b1, b2);
            } else if (p.b_big) {
              if constexpr (kCG == 2) tma_load_5d_2sm(sb, &tmB, &full_bar[stage], 0, k0, n0 / kChunk, b1, b2);
              else tma_load_5d(sb, &tmB, &full_bar[stage], 0, k0, n0 / kChunk, b1, b2);
            } else {
              for (int j = 0; j < bn_local / kChunk; ++j)
                ld4(sb + j * (kBK * 128), &tmB, &full_bar[stage], n0 + j * kChunk, k0, b1, b2);
            }
          } else if constexpr (kMode == MODE_FPROP) {
            // K order: tap-major, then input-channel block.  A = packed weights [tap][Cout][Cin]; a single-tap problem
            // (pre-gathered patches, i.e. a 1x1 conv) reads the centre (fr = fs = 1).
            if constexpr (kCG == 2) tma_load_3d_2sm(sa, &tmA, &full_bar[stage], cb * kBK, m0, t9);
            else tma_load_3d(sa, &tmA, &full_bar[stage], cb * kBK, m0, t9);
            // B = NHWC activations; tile of block_n consecutive pixels = box (box_w, box_h, box_n) shifted by the tap
            ld4(sb, &tmB, &full_bar[stage], cb * kBK, fs - 1, h0 + fr - 1, img);
            if (++cb == p.cin_blocks) {
              cb = 0; ++t9;
              if (p.conv_taps != 1 && ++fs == 3) { fs = 0; ++fr; }
            }
          } else {
            // WGRAD: K step = 32 consecutive pixels.  A = dY [pixels][Cout] (MN-major), B = shifted X; the item covers
            // taps tap0 .. tap0+nt-1 (tpi = 3: one filter row; tpi = 2: pairs of taps; tpi = 1: a single tap)
            if (p.big_box) {
              // one TMA per operand tile: the maps carry the 32-channel chunk index as their outermost dimension, so a
              // box of (32 ch, 32 pixels, n chunks) lands as n consecutive [32 pixels][128 B] swizzle tiles
              if constexpr (kCG == 2) tma_load_3d_2sm(sa, &tmA, &full_bar[stage], 0, p0, m0 / kChunk);
              else tma_load_3d(sa, &tmA, &full_bar[stage], 0, p0, m0 / kChunk);
              for (int t = 0; t < nt; ++t) {
                const int r = (tap0 + t) / 3, s = (tap0 + t) - r * 3;
                if constexpr (kCG == 2) tma_load_5d_2sm(sb + t * tap_b_bytes, &tmB, &full_bar[stage], 0, s - 1, h0 + r - 1, img, n0 / kChunk);
                else tma_load_5d(sb + t * tap_b_bytes, &tmB, &full_bar[stage], 0, s - 1, h0 + r - 1, img, n0 / kChunk);
              }
            } else {
              for (int j = 0; j < kBlockM / kChunk; ++j)
                if constexpr (kCG == 2) tma_load_2d_2sm(sa + j * (kBK * 128), &tmA, &full_bar[stage], m0 + j * kChunk, p0);
                else tma_load_2d(sa + j * (kBK * 128), &tmA, &full_bar[stage], m0 + j * kChunk, p0);
              for (int t = 0; t < nt; ++t) {
                const int r = (tap0 + t) / 3, s = (tap0 + t) - r * 3;
                for (int j = 0; j < bn_local / kChunk; ++j)
                  ld4(sb + t * tap_b_bytes + j * (kBK * 128), &tmB, &full_bar[stage], n0 + j * kChunk, s - 1, h0 + r - 1, img);
              }
            }
            p0 += 32;                       // next 32 pixels: box_h more rows of this image, or box_n whole images
            if (p.box_n > 1) img += p.box_n;
            else { h0 += p.box_h; if (h0 >= p.H) { h0 = 0; ++img; } }
          }
          if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ======================================= MMA issuer =========================================
    if (leader && elect_one()) {
      const bool a_mn = (kMode == MODE_WGRAD) ? true : (kMode == MODE_FPROP ? false : (p.a_mn != 0));
      const bool b_mn = (kMode == MODE_WGRAD) ? true : (kMode == MODE_FPROP ? false : (p.b_mn != 0));
      const uint32_t idesc = make_idesc(kBF16, a_mn, b_mn, kTileM, static_cast<uint32_t>(p.block_n));
      // descriptor strides: K-major -> 8-row groups 1024 B apart; MN-major -> 128-B-wide chunks kBK*128 apart
      const uint32_t a_lbo = a_mn ? kBK * 128u : 16u, b_lbo = b_mn ? kBK * 128u : 16u;
      const uint32_t a_kadv = a_mn ? kUmmaK * 128u : 32u, b_kadv = b_mn ? kUmmaK * 128u : 32u;
      // MN-major tf32 must use the 32-byte-atom flavour of the 128B swizzle (4-row groups, 512 B apart)
      const uint32_t a_lay = (a_mn && !kBF16) ? 1u : 2u, b_lay = (b_mn && !kBF16) ? 1u : 2u;
      const uint32_t a_sbo = (a_mn && !kBF16) ? 512u : 1024u, b_sbo = (b_mn && !kBF16) ? 512u : 1024u;
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int item = first_item; item < num_items; item += item_stride) {
        int m_blk, n_blk, tap, split;
        decode(item, m_blk, n_blk, tap, split);
        const int nk = ksteps_of(split, m_blk);
        if (skipped(m_blk, n_blk)) continue;
        mbar_wait(&tempty_bar[acc], acc_phase ^ 1u, 0x200u + acc);
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + static_cast<uint32_t>(acc) * acc_cols;
        uint32_t idesc_item = idesc;
        if constexpr (kMode == MODE_WGRAD)
          idesc_item = make_idesc(kBF16, a_mn, b_mn, kTileM, static_cast<uint32_t>(p.block_n * min(p.tpi, 9 - tap * p.tpi)));
        for (int i = 0; i < nk; ++i) {
          mbar_wait(&full_bar[stage], phase, 0x300u + stage);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + static_cast<size_t>(stage) * stage_bytes);
          const uint32_t sb = sa + kStageABytes;
          // WGRAD with several taps per item: the taps' B tiles are consecutive 128-byte-wide chunks in the stage, so
          // ONE instruction with N = taps * block_n covers them all.  (An SS-mode MMA costs ~118 cycles whatever its N -
          // bench/mma_probe.cu - so three N=64 instructions were three times the price of one N=192.)
#pragma unroll
          for (int k = 0; k < kBK / kUmmaK; ++k) {
            const uint64_t adesc = make_smem_desc_sw128(sa + k * a_kadv, a_lbo, a_sbo, a_lay);
            const uint64_t bdesc = make_smem_desc_sw128(sb + k * b_kadv, b_lbo, b_sbo, b_lay);
            if constexpr (kCG == 2) tc_mma_2sm<kBF16>(tmem_d, adesc, bdesc, idesc_item, (i > 0 || k > 0) ? 1u : 0u);
            else tc_mma<kBF16>(tmem_d, adesc, bdesc, idesc_item, (i > 0 || k > 0) ? 1u : 0u);
          }
          if constexpr (kCG == 2) tc_commit_2sm(&empty_bar[stage]); else tc_commit(&empty_bar[stage]);   // slot free (both CTAs)
          if (++stage == p.num_stages) { stage = 0; phase ^= 1u; }
        }
        if constexpr (kCG == 2) tc_commit_2sm(&tfull_bar[acc]); else tc_commit(&tfull_bar[acc]);   // accumulator complete
        if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1u; }
      }
    }
  } else if (warp >= 4) {
    // ======================================= epilogue ===========================================
    // two epilogue warps per TMEM lane quarter: warps 4-7 drain the first half of the accumulator columns, warps 8-11
    // the second half (small-K GEMMs are paced by the epilogue, not by the MMAs)
    const int ew = (warp - 4) & 3;           // == warp % 4 -> TMEM lanes [32*ew, 32*ew+32)
    const int ehalf = (warp - 4) >> 2;
    const int cols_half = ((static_cast<int>(acc_cols) + 63) / 64) * 32;
    const int cc_begin = ehalf * cols_half;
    const int cc_end_all = min(cc_begin + cols_half, static_cast<int>(acc_cols));
    int acc = 0;
    uint32_t acc_phase = 0;
    for (int item = first_item; item < num_items; item += item_stride) {
      int m_blk, n_blk, tap, split;
      decode(item, m_blk, n_blk, tap, split);
      const int nk = ksteps_of(split, m_blk);
      if (skipped(m_blk, n_blk)) continue;
      const long long boff = (kMode == MODE_GEMM)
                                 ? static_cast<long long>(tap % p.batch1) * p.out_bs1 + static_cast<long long>(tap / p.batch1) * p.out_bs2
                                 : 0;
      const int row = m_blk * kTileM + static_cast<int>(cta_rank) * kBlockM + ew * 32 + lane;   // this thread's output row
      const int n0 = n_blk * p.block_n;     // the epilogue sees all block_n columns of this CTA's 128 rows
      mbar_wait(&tfull_bar[acc], acc_phase, 0x400u + acc);
      tc_fence_after();
      const uint32_t taddr = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + static_cast<uint32_t>(acc) * acc_cols;
      float ssum = 0.f, ssq = 0.f;
      int cc_end = cc_end_all;
      if constexpr (kMode == MODE_WGRAD) cc_end = min(cc_end_all, p.block_n * min(p.tpi, 9 - tap * p.tpi));
      for (int cc = cc_begin; cc < cc_end; cc += 32) {
        // cc walks all accumulator columns of the item; c0 is the column inside the current tap's tile
        const int tsub = cc / p.block_n;
        const int c0 = cc - tsub * p.block_n;
        uint32_t v[32];
        tmem_ld_32x32(taddr + cc, v);
        tmem_ld_wait();
        if (nk == 0) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if constexpr (kMode == MODE_GEMM) {
          if (row < p.M) {
            if (p.fix_ws != nullptr) {
              // host guarantees N % 32 == 0 and no batching: a chunk is either whole or entirely past the last column
              float* o = p.fix_ws + static_cast<long long>(row) * p.N + n0 + c0;
              if (n0 + c0 < p.N) {
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j), "f"(__uint_as_float(v[j]) * p.alpha),
                               "f"(__uint_as_float(v[j + 1]) * p.alpha), "f"(__uint_as_float(v[j + 2]) * p.alpha),
                               "f"(__uint_as_float(v[j + 3]) * p.alpha)
                               : "memory");
              }
            } else if (p.atomic) {
              float* o = reinterpret_cast<float*>(p.out) + boff + static_cast<long long>(row) * p.ldo + n0 + c0;
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (n0 + c0 + j < p.N) atomicAdd(o + j, __uint_as_float(v[j]) * p.alpha);
            }
          }
          if (!p.atomic) {
            // one bias load per lane per 32-column chunk, handed round with shuffles: a per-element __ldg put a
            // load-latency stall (and a branch) on every output element and paced the whole kernel (ncu: 25% long_sb)
            float bl = 0.f;
            if (p.bias != nullptr && n0 + c0 + lane < p.N)
              bl = p.bias_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.bias)[n0 + c0 + lane])
                               : __ldg(p.bias + n0 + c0 + lane);
            float f[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = fmaf(__uint_as_float(v[j]), p.alpha, __shfl_sync(0xffffffffu, bl, j));
            const bool full = (n0 + c0 + 32 <= p.N);
            uint8_t* stg = epi_stage + (warp - 4) * 4096;
            const int row0 = m_blk * kTileM + static_cast<int>(cta_rank) * kBlockM + ew * 32;
            // warp-uniform store of the warp's 32x32 chunk as bf16: transpose through shared memory (XOR-swizzled 16-byte
            // pieces, conflict-free both ways) so that every global store instruction writes whole row segments; the
            // elementwise pass folded into the store (aux_op) works on those 16-byte pieces
            auto store_bf16 = [&](__nv_bfloat16* base, int aux_op) {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                __nv_bfloat162 h0 = __floats2bfloat162_rn(f[8 * j], f[8 * j + 1]);
                __nv_bfloat162 h1 = __floats2bfloat162_rn(f[8 * j + 2], f[8 * j + 3]);
                __nv_bfloat162 h2 = __floats2bfloat162_rn(f[8 * j + 4], f[8 * j + 5]);
                __nv_bfloat162 h3 = __floats2bfloat162_rn(f[8 * j + 6], f[8 * j + 7]);
                uint4 u;
                u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
                u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
                *reinterpret_cast<uint4*>(stg + lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4)) = u;
              }
              __syncwarp();
              __nv_bfloat16* ob = base + boff + n0 + c0;
#pragma unroll
              for (int it = 0; it < 4; ++it) {
                const int r = it * 8 + (lane >> 2), q = lane & 3;
                uint4 u = *reinterpret_cast<const uint4*>(stg + r * 64 + ((q ^ ((r >> 1) & 3)) << 4));
                if (row0 + r < p.M) {
                  const long long off = static_cast<long long>(row0 + r) * p.ldo + q * 8;
                  if (aux_op == 1) u = add_bf16x8(u, __ldg(reinterpret_cast<const uint4*>(p.aux + n0 + c0 + off)));
                  else if (aux_op == 2) u = mul_gelu_grad_bf16x8(u, __ldg(reinterpret_cast<const uint4*>(p.aux + n0 + c0 + off)));
                  *reinterpret_cast<uint4*>(ob + off) = u;
                }
              }
              __syncwarp();
            };
            const bool staged = p.st_vec && full;
            if (p.out2 != nullptr) {
              // the pre-activation goes out as well (bf16; the host accepts out2 only with whole, 16-byte-aligned rows)
              if (staged) {
                store_bf16(p.out2, 0);
              } else if (row < p.M) {
                __nv_bfloat16* o2 = p.out2 + static_cast<long long>(row) * p.ldo + n0 + c0;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (n0 + c0 + j < p.N) o2[j] = __float2bfloat16(f[j]);
              }
            }
            if (p.act == 1) {
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = fmaxf(f[j], 0.f);
            } else if (p.act == 2) {
#pragma unroll
              for (int j = 0; j < 32; ++j) f[j] = gelu_tanh(f[j]);
            }
            if (staged) {
              if (p.out_bf16) {
                store_bf16(reinterpret_cast<__nv_bfloat16*>(p.out), p.aux_op);
              } else {
#pragma unroll
                for (int j = 0; j < 8; ++j)
                  *reinterpret_cast<float4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                      make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                __syncwarp();
                float* of = reinterpret_cast<float*>(p.out) + boff + n0 + c0;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                  const int r = it * 4 + (lane >> 3), q = lane & 7;
                  const float4 u = *reinterpret_cast<const float4*>(stg + r * 128 + ((q ^ (r & 7)) << 4));
                  if (row0 + r < p.M) *reinterpret_cast<float4*>(of + static_cast<long long>(row0 + r) * p.ldo + q * 4) = u;
                }
                __syncwarp();
              }
            } else if (row < p.M) {
              if (p.out_bf16) {
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + boff + static_cast<long long>(row) * p.ldo + n0 + c0;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (n0 + c0 + j < p.N) {
                    __nv_bfloat16 hv = __float2bfloat16(f[j]);
                    if (p.aux_op != 0) {
                      const float a = __bfloat162float(p.aux[static_cast<long long>(row) * p.ldo + n0 + c0 + j]);
                      hv = __float2bfloat16(p.aux_op == 1 ? __bfloat162float(hv) + a : __bfloat162float(hv) * gelu_tanh_grad(a));
                    }
                    o[j] = hv;
                  }
              } else {
                float* o = reinterpret_cast<float*>(p.out) + boff + static_cast<long long>(row) * p.ldo + n0 + c0;
#pragma unroll
                for (int j = 0; j < 32; ++j)
                  if (n0 + c0 + j < p.N) o[j] = f[j];
              }
            }
          }
        } else if constexpr (kMode == MODE_FPROP) {
          // row = output channel, column = pixel; NHWC store is coalesced across the warp's 32 channels
          if (row < p.M) {
            float* o = reinterpret_cast<float*>(p.out) + static_cast<long long>(n0 + c0) * p.ldo + row;
            const float b = (p.bias != nullptr) ? __ldg(p.bias + row) : 0.f;
            // the host guarantees that pixel tiles are whole (P % block_n == 0): no per-element bound check
            if (p.act == 1) {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float x = __uint_as_float(v[j]) + b;
                ssum += x;
                ssq += x * x;
                o[static_cast<long long>(j) * p.ldo] = fmaxf(x, 0.f);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const float x = __uint_as_float(v[j]) + b;
                ssum += x;
                ssq += x * x;
                o[static_cast<long long>(j) * p.ldo] = x;
              }
            }
          }
        } else {
          // WGRAD: row = Cout, column = Cin; scatter-accumulate into the torch-layout gradient [Cout][Cin][3][3]
          if (row < p.M && n0 + c0 + 32 <= p.N) {        // Cin is a multiple of 32: whole chunks only
            if (p.packed) {
              // tap-major staging layout [9][Cout][Cin]: the thread's 32 columns are 128 contiguous bytes, so eight
              // 16-byte vector reductions replace 32 scalar ones 36 bytes apart (the SM issues ~1.3 cycles per scalar RED
              // lane, and with one wave of work items the epilogue is not hidden behind a next item)
              float* o = reinterpret_cast<float*>(p.out) +
                         (static_cast<long long>(tap * p.tpi + tsub) * p.M + row) * p.cin_total + n0 + c0;
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(o + j), "r"(v[j]), "r"(v[j + 1]),
                             "r"(v[j + 2]), "r"(v[j + 3])
                             : "memory");
            } else {
              float* o = reinterpret_cast<float*>(p.out) + (static_cast<long long>(row) * p.cin_total + n0 + c0) * 9 +
                         tap * p.tpi + tsub;
#pragma unroll
              for (int j = 0; j < 32; ++j) atomicAdd(o + j * 9, __uint_as_float(v[j]));
            }
          }
        }
      }
      if constexpr (kMode == MODE_FPROP) {
        if (p.stat_sum != nullptr && row < p.M) {
          atomicAdd(p.stat_sum + row, ssum);
          atomicAdd(p.stat_sq + row, ssq);
        }
      }
      if constexpr (kMode == MODE_GEMM) {
        if (p.fix_ws != nullptr) {
          // this warp's share of the item is in the workspace; the last of the num_splits arrivals for this sub-block
          // (same warp slot of the same tile in every split) owns the finished sums
          __threadfence();
          __syncwarp();
          unsigned int old = 0;
          const int slot = ((m_blk * p.n_blocks + n_blk) * kCG + static_cast<int>(cta_rank)) * 8 + (warp - 4);
          if (lane == 0) old = atomicAdd(p.fix_cnt + slot, 1u);
          old = __shfl_sync(0xffffffffu, old, 0);
          if (old == static_cast<unsigned int>(p.num_splits - 1)) {
            __threadfence();
            if (row < p.M) {
              for (int cc = cc_begin; cc < cc_end && n0 + cc < p.N; cc += 32) {
                float* w = p.fix_ws + static_cast<long long>(row) * p.N + n0 + cc;
                __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(p.out) + static_cast<long long>(row) * p.ldo + n0 + cc;
                float4 q[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) q[j] = __ldcg(reinterpret_cast<const float4*>(w) + j);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  __nv_bfloat162 h0 = __floats2bfloat162_rn(q[2 * j].x, q[2 * j].y), h1 = __floats2bfloat162_rn(q[2 * j].z, q[2 * j].w);
                  __nv_bfloat162 h2 = __floats2bfloat162_rn(q[2 * j + 1].x, q[2 * j + 1].y), h3 = __floats2bfloat162_rn(q[2 * j + 1].z, q[2 * j + 1].w);
                  uint4 u;
                  u.x = *reinterpret_cast<uint32_t*>(&h0); u.y = *reinterpret_cast<uint32_t*>(&h1);
                  u.z = *reinterpret_cast<uint32_t*>(&h2); u.w = *reinterpret_cast<uint32_t*>(&h3);
                  reinterpret_cast<uint4*>(o)[j] = u;
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) reinterpret_cast<float4*>(w)[j] = make_float4(0.f, 0.f, 0.f, 0.f);
              }
            }
            __syncwarp();
            if (lane == 0) p.fix_cnt[slot] = 0u;
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kCG == 2 && !leader) mbar_arrive_cluster(&tempty_bar[acc], 0); else mbar_arrive(&tempty_bar[acc]);
      }
      if (++acc == p.acc_stages) { acc = 0; acc_phase ^= 1u; }
    }
  }

  tc_fence_before();
  if constexpr (kCG == 2) cluster_sync_all(); else __syncthreads();   // the peer may still read this CTA's smem / barriers
  tc_fence_after();
  if (warp == 2) {
    if constexpr (kCG == 2) tmem_dealloc_2sm(tmem_base, 512); else tmem_dealloc(tmem_base, 512);
  }
}

// =================================================================================================
// host side
// =================================================================================================
static PFN_cuTensorMapEncodeTiled_v12000 get_encode_fn() {
  static PFN_cuTensorMapEncodeTiled_v12000 fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(ptr);
  });
  return fn;
}

// rank-N tiled map with 128B swizzle; dims/strides innermost first; strides in bytes for dims 1..rank-1.
// Encoded maps are cached (key: base pointer, type, rank, extents, strides, box, swizzle flavour): a training step asks
// for the same few hundred maps every iteration, and cuTensorMapEncodeTiled was a visible part of the per-call host time
// (profiles/probe_findings.md section 9).  A map depends only on its key, so a stale entry cannot exist.
struct TmapKey {
  uint64_t base; uint64_t dims[5]; uint64_t strides[4]; uint32_t box[5]; int rank; int flags;
  bool operator==(const TmapKey& o) const { return memcmp(this, &o, sizeof(TmapKey)) == 0; }
};
struct TmapEntry { TmapKey key; CUtensorMap map; bool valid; };
constexpr int kTmapCacheSize = 8192;     // direct-mapped
static TmapEntry* g_tmap_cache = nullptr;
static std::mutex g_tmap_mutex;
static unsigned long long g_tmap_hits = 0, g_tmap_misses = 0;

int make_tmap(CUtensorMap* out, const void* base, bool bf16, int rank, const uint64_t* dims,
              const uint64_t* strides_bytes, const uint32_t* box, bool atom32) {
  TmapKey key;
  memset(&key, 0, sizeof(key));
  key.base = reinterpret_cast<uint64_t>(base); key.rank = rank; key.flags = (bf16 ? 1 : 0) | (atom32 ? 2 : 0);
  for (int i = 0; i < rank; ++i) { key.dims[i] = dims[i]; key.box[i] = box[i]; }
  for (int i = 0; i < rank - 1; ++i) key.strides[i] = strides_bytes[i];
  uint64_t h = 1469598103934665603ull;
  const uint64_t* w = reinterpret_cast<const uint64_t*>(&key);
  for (size_t i = 0; i < sizeof(TmapKey) / 8; ++i) { h ^= w[i]; h *= 1099511628211ull; h ^= h >> 29; }
  const int slot = static_cast<int>(h % kTmapCacheSize);
  {
    std::lock_guard<std::mutex> g(g_tmap_mutex);
    if (!g_tmap_cache) g_tmap_cache = static_cast<TmapEntry*>(calloc(kTmapCacheSize, sizeof(TmapEntry)));
    TmapEntry& e = g_tmap_cache[slot];
    if (e.valid && e.key == key) { *out = e.map; ++g_tmap_hits; return DDPX_OK; }
  }
  auto fn = get_encode_fn();
  if (!fn) return DDPX_ERR_DRIVER;
  cuuint64_t gdim[5];
  cuuint64_t gstr[4];
  cuuint32_t bx[5], es[5];
  for (int i = 0; i < rank; ++i) { gdim[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
  for (int i = 0; i < rank - 1; ++i) gstr[i] = strides_bytes[i];
  // cuTensorMapEncodeTiled is a driver-API call: it needs a context current on THIS thread.  A thread that has not yet made
  // a runtime call of its own - autograd's backward thread whose first native operation builds a tensor map - has none
  // (CUDA_ERROR_INVALID_CONTEXT), so bind the runtime's primary context once per thread, and once more if the driver
  // still says so.
  static thread_local bool ctx_bound = false;
  if (!ctx_bound) { cudaFree(nullptr); ctx_bound = true; }
  auto encode = [&]() {
    return fn(out, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, rank,
              const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
              atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
              CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  };
  CUresult r = encode();
  if (r == CUDA_ERROR_INVALID_CONTEXT || r == CUDA_ERROR_NOT_INITIALIZED) {
    cudaFree(nullptr);
    r = encode();
  }
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "[ddpx] cuTensorMapEncodeTiled failed: %d (rank %d)\n", static_cast<int>(r), rank);
    return DDPX_ERR_TMAP;
  }
  {
    std::lock_guard<std::mutex> g(g_tmap_mutex);
    TmapEntry& e = g_tmap_cache[slot];
    e.key = key; e.map = *out; e.valid = true; ++g_tmap_misses;
  }
  return DDPX_OK;
}

static int g_num_sms = 0;
int num_sms() {
  if (g_num_sms == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
  }
  return g_num_sms;
}

constexpr int kEpiStageBytes = 8 * 4096;   // MODE_GEMM epilogue staging: 4 KB per epilogue warp

static int pick_stages(int block_n, int tpi, int reserve) {
  const int stage = kStageABytes + block_n * 128 * tpi;
  int s = (224 * 1024 - reserve) / stage;
  if (s > kMaxStages) s = kMaxStages;
  static int cap = -1;       // DDPX_MAX_STAGES: pipeline-depth experiments (latency- vs bandwidth-bound)
  if (cap < 0) { const char* e = getenv("DDPX_MAX_STAGES"); cap = e ? atoi(e) : 0; }
  if (cap >= 2 && s > cap) s = cap;
  if (s < 2) s = 2;
  return s;
}

template <int kMode, bool kBF16, int kCG>
static int launch_tc_cg(const CUtensorMap& a, const CUtensorMap& b, TcParams& p, cudaStream_t stream) {
  if (p.tpi < 1) p.tpi = 1;
  p.acc_stages = (2 * p.tpi * p.block_n <= 512) ? 2 : 1;
  const int reserve = (kMode == MODE_GEMM) ? kEpiStageBytes : 0;
  const int bn_local = p.block_n / kCG;
  p.num_stages = pick_stages(bn_local, p.tpi, reserve);
  const size_t smem = static_cast<size_t>(p.num_stages) * (kStageABytes + bn_local * 128 * p.tpi) + 1024 /*align*/ + 256 + reserve;
  static bool attr_set = false;
  if (!attr_set) {
    cudaError_t e = cudaFuncSetAttribute(tc_kernel<kMode, kBF16, kCG>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return DDPX_ERR_LAUNCH;
    attr_set = true;
  }
  const int taps = (kMode == MODE_WGRAD) ? (9 + p.tpi - 1) / p.tpi : (kMode == MODE_GEMM ? p.batch1 * p.batch2 : 1);
  const int items = p.m_blocks * p.n_blocks * taps * p.num_splits;
  const int max_clusters = num_sms() / kCG;
  int clusters = items < max_clusters ? items : max_clusters;
  if (clusters < 1) clusters = 1;
  if (kCG == 1) {
    tc_kernel<kMode, kBF16, kCG><<<clusters, kNumThreads, smem, stream>>>(a, b, p);
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(clusters * kCG);
    cfg.blockDim = dim3(kNumThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kCG; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    if (cudaLaunchKernelEx(&cfg, tc_kernel<kMode, kBF16, kCG>, a, b, p) != cudaSuccess) return DDPX_ERR_LAUNCH;
  }
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

// DDPX_CTA_PAIRS=0 forces single-CTA tiles everywhere (A/B comparisons, fallback)
static bool pairs_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DDPX_CTA_PAIRS"); v = (e && e[0] == '0') ? 0 : 1; }
  return v == 1;
}

// cg = 2 asks for CTA pairs: the caller has laid the problem out in 256-row tiles (p.m_blocks) and built the B tensor
// map with a box of block_n / 2 rows.
template <int kMode, bool kBF16>
static int launch_tc(const CUtensorMap& a, const CUtensorMap& b, TcParams& p, cudaStream_t stream, int cg = 1) {
  if (cg == 2) return launch_tc_cg<kMode, kBF16, 2>(a, b, p, stream);
  return launch_tc_cg<kMode, kBF16, 1>(a, b, p, stream);
}

}  // namespace ddpx

using namespace ddpx;

// Batched GEMM: problem (b1, b2) reads A + b1*a_bs1 + b2*a_bs2 (elements) etc.  All strides in elements.
static int gemm_impl(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                     long long ldd, int a_mn, int b_mn, int bf16, int out_bf16, const void* bias_, int bias_bf16, int act,
                     float alpha, int split_k, int accumulate, int batch1, int batch2, long long a_bs1,
                     long long a_bs2, long long b_bs1, long long b_bs2, long long d_bs1, long long d_bs2,
                     int causal, void* stream_, float* fix_ws = nullptr, unsigned int* fix_cnt = nullptr,
                     const void* aux = nullptr, int aux_op = 0, void* out2 = nullptr) {
  const float* bias = static_cast<const float*>(bias_);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int elt = bf16 ? 2 : 4;
  const int bk = 128 / elt, chunk = 128 / elt;
  if (M <= 0 || N <= 0 || K <= 0 || batch1 < 1 || batch2 < 1) return DDPX_ERR_ARG;
  if ((lda * elt) % 16 || (ldb * elt) % 16) return DDPX_ERR_ALIGN;
  if ((a_bs1 * elt) % 16 || (a_bs2 * elt) % 16 || (b_bs1 * elt) % 16 || (b_bs2 * elt) % 16) return DDPX_ERR_ALIGN;
  if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) return DDPX_ERR_ALIGN;
  TcParams p{};
  p.M = M; p.N = N;
  p.batch1 = batch1; p.batch2 = batch2; p.out_bs1 = d_bs1; p.out_bs2 = d_bs2; p.causal = causal; p.alpha = alpha;
  // N tile: the largest of {256,128,64} that still gives every SM a tile; narrow outputs get a narrow tile
  const int mb = (M + kBlockM - 1) / kBlockM;
  const int nb_total = batch1 * batch2;
  const int min_bn = (b_mn && chunk > 32) ? chunk : 32;
  int bn = 256;
  const int want_splits = split_k > 1 ? split_k : 1;      // split-K multiplies the number of work items
  while (bn > 64 && static_cast<long long>(mb) * ((N + bn - 1) / bn) * nb_total * want_splits < num_sms()) bn /= 2;
  while (bn > min_bn && bn / 2 >= N) bn /= 2;
  // Compute-bound bf16 problems (>= 1024 rows, long K) that the rule above would cut into 128- or 64-column tiles to get a
  // tile per SM: an SS-mode MMA costs ~118 cycles whatever its N (profiles/probe_findings.md section 1), so a 128-column
  // tile runs the tensor pipe at half rate.  Compare modelled times (waves x k-steps x MMA cost + epilogue per item) of the
  // narrow choice and of 256-column tiles and keep the cheaper: GPT-2 XL's 2048 x 1600 outputs went from 208 tiles of
  // 128 x 128 (two waves at half rate) to 56 CTA-pair tiles of 256 x 256 (one wave at full rate).
  if (bf16 && bn < 256 && N >= 256 && M >= 1024 && K >= 512 && nb_total == 1 && causal == 0 && min_bn <= 256) {
    auto model = [&](int tile_m, int tile_n, int units) {
      const long long items = static_cast<long long>((M + tile_m - 1) / tile_m) * ((N + tile_n - 1) / tile_n) * want_splits;
      const long long waves = (items + units - 1) / units;
      const long long ksteps = ((K + bk - 1) / bk + want_splits - 1) / want_splits;
      const long long mma = 4LL * (tile_n / 2 > 118 ? tile_n / 2 : 118);      // 4 MMAs per 64-element k-step
      return waves * (ksteps * mma + 1500);
    };
    const long long narrow = model(kBlockM, bn, num_sms());
    const long long wide1 = model(kBlockM, 256, num_sms());
    const long long wide2 = pairs_enabled() && M >= 256 ? model(256, 256, num_sms() / 2) : (1LL << 62);
    if (wide1 < narrow || wide2 < narrow) bn = 256;
  }
  if (causal == 1 && bn > kBlockM) bn = kBlockM;      // square tiles so whole tiles fall above the diagonal
  if (bn < min_bn) bn = min_bn;
  // CTA pairs (256-row tiles) whenever the problem is tall and wide enough for 256x256 tiles to fill the machine
  const int cg = (pairs_enabled() && bn == 256 && M >= 256 && causal == 0 &&
                  static_cast<long long>((M + 255) / 256) * ((N + 255) / 256) * nb_total * want_splits >= num_sms() / 4) ? 2 : 1;
  p.block_n = bn;
  p.m_blocks = cg == 2 ? (M + 255) / 256 : mb;
  p.n_blocks = (N + bn - 1) / bn;
  p.k_steps_total = (K + bk - 1) / bk;
  int splits = split_k > 0 ? split_k : 1;
  if (splits > p.k_steps_total) splits = p.k_steps_total;
  p.k_steps_per_split = (p.k_steps_total + splits - 1) / splits;
  p.num_splits = (p.k_steps_total + p.k_steps_per_split - 1) / p.k_steps_per_split;
  p.a_mn = a_mn; p.b_mn = b_mn;
  p.out = D; p.ldo = ldd; p.out_bf16 = out_bf16;
  p.atomic = (accumulate || p.num_splits > 1) ? 1 : 0;
  if (fix_ws != nullptr) {
    // split-K with the in-kernel fix-up: bf16 result, fp32 partial sums in the caller's (all-zero) workspace
    if (!out_bf16 || N % 32 || batch1 * batch2 != 1 || accumulate || bias != nullptr || act != 0 || (ldd % 8) ||
        (reinterpret_cast<uintptr_t>(D) & 15) || (reinterpret_cast<uintptr_t>(fix_ws) & 15) || fix_cnt == nullptr)
      return DDPX_ERR_ARG;
    p.fix_ws = fix_ws; p.fix_cnt = fix_cnt; p.atomic = 1;
  } else if (p.atomic && out_bf16) {
    return DDPX_ERR_ARG;
  }
  p.act = act; p.bias = bias; p.bias_bf16 = bias_bf16;
  if (aux_op != 0 || out2 != nullptr) {
    if (!out_bf16 || p.atomic || batch1 * batch2 != 1 || aux_op < 0 || aux_op > 2 || (aux_op != 0 && aux == nullptr) ||
        (reinterpret_cast<uintptr_t>(aux) & 15) || (reinterpret_cast<uintptr_t>(out2) & 15) ||
        (out2 != nullptr && (act != 2 || (ldd % 8))))
      return DDPX_ERR_ARG;
    p.aux = static_cast<const __nv_bfloat16*>(aux); p.aux_op = aux_op; p.out2 = static_cast<__nv_bfloat16*>(out2);
  }
  {
    const int oe = out_bf16 ? 2 : 4;
    p.st_vec = ((ldd * oe) % 16 == 0 && (d_bs1 * oe) % 16 == 0 && (d_bs2 * oe) % 16 == 0 &&
                (reinterpret_cast<uintptr_t>(D) & 15) == 0) ? 1 : 0;
  }
  CUtensorMap ta, tb;
  int rc;
  {
    uint64_t dims[4], str[3]; uint32_t box[4];
    dims[2] = batch1; dims[3] = batch2; box[2] = 1; box[3] = 1;
    static int gemm_big = -1;      // DDPX_GEMM_BIGBOX=0: one TMA per 128-byte chunk (the older path), for A/B runs
    if (gemm_big < 0) { const char* e = getenv("DDPX_GEMM_BIGBOX"); gemm_big = (e && e[0] == '0') ? 0 : 1; }
    // chunked 5-D view of an MN-major operand: (chunk elements, K rows, chunks, batch1, batch2); whole chunks only
    auto big_map = [&](CUtensorMap* tm, const void* base, int MN, long long ld, long long bs1, long long bs2, int tile_rows) {
      uint64_t d5[5] = {static_cast<uint64_t>(chunk), static_cast<uint64_t>(K), static_cast<uint64_t>(MN / chunk),
                        static_cast<uint64_t>(batch1), static_cast<uint64_t>(batch2)};
      uint64_t s5[4] = {static_cast<uint64_t>(ld) * elt, 128,
                        batch1 > 1 ? static_cast<uint64_t>(bs1) * elt : static_cast<uint64_t>(ld) * elt * K,
                        batch2 > 1 ? static_cast<uint64_t>(bs2) * elt : static_cast<uint64_t>(ld) * elt * K * batch1};
      uint32_t b5[5] = {static_cast<uint32_t>(chunk), static_cast<uint32_t>(bk), static_cast<uint32_t>(tile_rows / chunk), 1, 1};
      return make_tmap(tm, base, bf16, 5, d5, s5, b5, !bf16);
    };
    p.a_big = (gemm_big && a_mn && M % chunk == 0) ? 1 : 0;
    p.b_big = (gemm_big && b_mn && N % chunk == 0 && (bn / cg) % chunk == 0) ? 1 : 0;
    if (!a_mn) { dims[0] = K; dims[1] = M; box[0] = bk; box[1] = kBlockM; }
    else       { dims[0] = M; dims[1] = K; box[0] = chunk; box[1] = bk; }
    str[0] = static_cast<uint64_t>(lda) * elt;
    str[1] = batch1 > 1 ? static_cast<uint64_t>(a_bs1) * elt : str[0] * dims[1];
    str[2] = batch2 > 1 ? static_cast<uint64_t>(a_bs2) * elt : str[1] * dims[2];
    if (p.a_big && big_map(&ta, A, M, lda, a_bs1, a_bs2, kBlockM) != DDPX_OK) p.a_big = 0;
    if (!p.a_big && (rc = make_tmap(&ta, A, bf16, 4, dims, str, box, a_mn && !bf16)) != DDPX_OK) return rc;
    if (!b_mn) { dims[0] = K; dims[1] = N; box[0] = bk; box[1] = bn / cg; }
    else       { dims[0] = N; dims[1] = K; box[0] = chunk; box[1] = bk; }
    str[0] = static_cast<uint64_t>(ldb) * elt;
    str[1] = batch1 > 1 ? static_cast<uint64_t>(b_bs1) * elt : str[0] * dims[1];
    str[2] = batch2 > 1 ? static_cast<uint64_t>(b_bs2) * elt : str[1] * dims[2];
    if (p.b_big && big_map(&tb, B, N, ldb, b_bs1, b_bs2, bn / cg) != DDPX_OK) p.b_big = 0;
    if (!p.b_big && (rc = make_tmap(&tb, B, bf16, 4, dims, str, box, b_mn && !bf16)) != DDPX_OK) return rc;
  }
  return bf16 ? launch_tc<MODE_GEMM, true>(ta, tb, p, stream, cg) : launch_tc<MODE_GEMM, false>(ta, tb, p, stream, cg);
}

extern "C" int ddpx_gemm_batched(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                                 long long ldd, int a_mn, int b_mn, int bf16, int out_bf16, const float* bias, int act,
                                 float alpha, int split_k, int accumulate, int batch1, int batch2, long long a_bs1,
                                 long long a_bs2, long long b_bs1, long long b_bs2, long long d_bs1, long long d_bs2,
                                 int causal, void* stream_) {
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, bf16, out_bf16, bias, 0, act, alpha, split_k, accumulate,
                   batch1, batch2, a_bs1, a_bs2, b_bs1, b_bs2, d_bs1, d_bs2, causal, stream_);
}

// Split-K GEMM with a bf16 result and no separate memset / cast launches: D[M,N] (bf16) = op(A) op(B), K split `split_k`
// ways.  ws = fp32 [M*N] and cnt = uint32 [(ceil(M/128) * ceil(N/64) + 1) * 16] must be ALL ZERO on entry and are left all
// zero (persistent scratch: zero them once).  N % 32 == 0, ldd % 8 == 0.  Calls that share ws / cnt must be stream-ordered.
extern "C" int ddpx_gemm_splitk_bf16(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                                     long long ldd, int a_mn, int b_mn, int split_k, float* ws, unsigned int* cnt,
                                     void* stream_) {
  if (split_k < 1) return DDPX_ERR_ARG;
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, 1, 1, nullptr, 0, 0, 1.0f, split_k, 0, 1, 1, 0, 0, 0, 0, 0, 0,
                   0, stream_, ws, cnt);
}

// ddpx_gemm with the bias given in bf16 (bias_bf16 = 1) or fp32 (0)
extern "C" int ddpx_gemm_bias(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                              long long ldd, int a_mn, int b_mn, int bf16, int out_bf16, const void* bias, int bias_bf16,
                              int act, int split_k, int accumulate, void* stream_) {
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, bf16, out_bf16, bias, bias_bf16, act, 1.0f, split_k,
                   accumulate, 1, 1, 0, 0, 0, 0, 0, 0, 0, stream_);
}

// bf16 GEMM with an elementwise pass folded into the epilogue's store (see TcParams::aux):
//   aux_op 1: D = act(A B + bias) + aux (residual);  aux_op 2: D = (A B) * gelu'(aux);  out2 (act = 2 only): also store the
//   pre-activation A B + bias there.  aux / out2 are bf16 [M][ldd].
extern "C" int ddpx_gemm_fused(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                               long long ldd, int a_mn, int b_mn, const void* bias, int bias_bf16, int act, const void* aux,
                               int aux_op, void* out2, void* stream_) {
  return gemm_impl(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, 1, 1, bias, bias_bf16, act, 1.0f, 1, 0, 1, 1, 0, 0, 0, 0, 0, 0,
                   0, stream_, nullptr, nullptr, aux, aux_op, out2);
}

extern "C" int ddpx_gemm(const void* A, const void* B, void* D, int M, int N, int K, long long lda, long long ldb,
                         long long ldd, int a_mn, int b_mn, int bf16, int out_bf16, const float* bias, int act,
                         int split_k, int accumulate, void* stream_) {
  return ddpx_gemm_batched(A, B, D, M, N, K, lda, ldb, ldd, a_mn, b_mn, bf16, out_bf16, bias, act, 1.0f, split_k,
                           accumulate, 1, 1, 0, 0, 0, 0, 0, 0, 0, stream_);
}

// pixel-tile geometry shared by fprop / wgrad: `pixels` consecutive NHWC pixels as a (w,h,n) box
static int conv_box(int pixels, int H, int W, int Nimg, int* bw, int* bh, int* bn) {
  if (W > pixels || pixels % W) return DDPX_ERR_SHAPE;
  int rows = pixels / W;
  if (rows <= H) {
    if (H % rows) return DDPX_ERR_SHAPE;
    *bw = W; *bh = rows; *bn = 1;
  } else {
    if (rows % H) return DDPX_ERR_SHAPE;
    *bw = W; *bh = H; *bn = rows / H;
    if (Nimg % *bn) return DDPX_ERR_SHAPE;
  }
  if (*bw > 256 || *bh > 256 || *bn > 256) return DDPX_ERR_SHAPE;
  return DDPX_OK;
}

static int conv_fprop_impl(const float* x, const float* wpacked, float* y, int Nimg, int H, int W, int Cin, int Cout,
                           const float* bias, int act, float* stat_sum, float* stat_sq, int taps, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (Cin % 32 || Cout % 32) return DDPX_ERR_SHAPE;
  const long long P = static_cast<long long>(Nimg) * H * W;
  int bn = 256;
  while (bn > 32 && (P % bn)) bn /= 2;
  int bw, bh, bnn;
  while (bn >= 32 && conv_box(bn, H, W, Nimg, &bw, &bh, &bnn) != DDPX_OK) bn /= 2;
  if (bn < 32 || P % bn) return DDPX_ERR_SHAPE;
  int cg = 1;
  if (pairs_enabled() && bn == 256 && Cout % 256 == 0) {
    int bw2, bh2, bn2;       // each CTA of a pair stages 128 of the tile's 256 pixels
    if (conv_box(128, H, W, Nimg, &bw2, &bh2, &bn2) == DDPX_OK) { cg = 2; bw = bw2; bh = bh2; bnn = bn2; }
  }
  TcParams p{};
  p.M = Cout; p.N = static_cast<int>(P);
  p.block_n = bn;
  p.m_blocks = (Cout + kBlockM * cg - 1) / (kBlockM * cg);
  p.n_blocks = static_cast<int>(P / bn);
  p.cin_blocks = Cin / 32;
  p.k_steps_total = taps * p.cin_blocks;
  p.k_steps_per_split = p.k_steps_total;
  p.num_splits = 1;
  p.conv_taps = taps;
  p.H = H; p.W = W; p.box_w = bw; p.box_h = bh; p.box_n = bnn; p.cin_total = Cin;
  p.out = y; p.ldo = Cout; p.act = act; p.bias = bias; p.stat_sum = stat_sum; p.stat_sq = stat_sq;
  CUtensorMap ta, tb;
  int rc;
  {
    uint64_t dims[3] = {static_cast<uint64_t>(Cin), static_cast<uint64_t>(Cout), static_cast<uint64_t>(taps)};
    uint64_t str[2] = {static_cast<uint64_t>(Cin) * 4, static_cast<uint64_t>(Cin) * Cout * 4};
    uint32_t box[3] = {32, kBlockM, 1};
    if ((rc = make_tmap(&ta, wpacked, false, 3, dims, str, box)) != DDPX_OK) return rc;
  }
  {
    uint64_t dims[4] = {static_cast<uint64_t>(Cin), static_cast<uint64_t>(W), static_cast<uint64_t>(H),
                        static_cast<uint64_t>(Nimg)};
    uint64_t str[3] = {static_cast<uint64_t>(Cin) * 4, static_cast<uint64_t>(Cin) * W * 4,
                       static_cast<uint64_t>(Cin) * W * H * 4};
    uint32_t box[4] = {32, static_cast<uint32_t>(bw), static_cast<uint32_t>(bh), static_cast<uint32_t>(bnn)};
    if ((rc = make_tmap(&tb, x, false, 4, dims, str, box)) != DDPX_OK) return rc;
  }
  return launch_tc<MODE_FPROP, false>(ta, tb, p, stream, cg);
}

extern "C" int ddpx_conv3x3_fprop(const float* x, const float* wpacked, float* y, int Nimg, int H, int W, int Cin,
                                  int Cout, const float* bias, int act, float* stat_sum, float* stat_sq,
                                  void* stream_) {
  return conv_fprop_impl(x, wpacked, y, Nimg, H, W, Cin, Cout, bias, act, stat_sum, stat_sq, 9, stream_);
}

// 1x1 convolution == per-pixel GEMM with the fprop epilogue (NHWC store, bias, ReLU, BN statistics).  Used for the RGB
// stem after ddpx_im2col_stem gathered the 3x3x3 patches into 32-wide rows (wpacked = [Cout][32]).
extern "C" int ddpx_conv1x1_fprop(const float* x, const float* wpacked, float* y, int Nimg, int H, int W, int Cin,
                                  int Cout, const float* bias, int act, float* stat_sum, float* stat_sq,
                                  void* stream_) {
  return conv_fprop_impl(x, wpacked, y, Nimg, H, W, Cin, Cout, bias, act, stat_sum, stat_sq, 1, stream_);
}

static int conv_wgrad_impl(const float* dy, const float* x, float* dw, int Nimg, int H, int W, int Cin, int Cout,
                           int split_k, int packed, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (Cin % 32 || Cout % 32) return DDPX_ERR_SHAPE;
  const long long P = static_cast<long long>(Nimg) * H * W;
  if (P % 32) return DDPX_ERR_SHAPE;
  int bw, bh, bnn, rc;
  if ((rc = conv_box(32, H, W, Nimg, &bw, &bh, &bnn)) != DDPX_OK) return rc;
  TcParams p{};
  p.M = Cout; p.N = Cin;
  // taps per work item.  Narrow layers (Cin <= 64) are bound by re-reading dY once per tap: handle one filter row
  // (3 taps, three accumulators fed by one dY tile) per item - measured 803 -> 391 us on 64->128 @ 32x32.  Wide layers
  // keep one tap per item with a 256-column tile and double-buffered accumulators (3 taps there measured slower).
  // DDPX_WGRAD_TPI=1|3 forces either.
  static int tpi_cfg = -1;
  if (tpi_cfg < 0) {
    const char* e = getenv("DDPX_WGRAD_TPI");
    tpi_cfg = (e && e[0] >= '1' && e[0] <= '4') ? (e[0] - '0') : 0;
  }
  // taps per item so that one MMA is as close to N = 256 as the layer allows: Cin <= 64 -> a filter row (3 taps,
  // N = 192), Cin = 128 -> 2 taps (N = 256; groups 2,2,2,2,1), wider -> 1 tap with a 256-column tile
  p.tpi = tpi_cfg ? tpi_cfg : (Cin <= 64 ? 3 : (Cin == 128 ? 2 : 1));
  const int max_bn = 256 / p.tpi / 32 * 32;    // all taps of a group form ONE MMA: taps * block_n <= 256
  int bn = Cin >= max_bn ? max_bn : Cin;       // Cin is a multiple of 32
  while (Cin % bn) bn -= 32;
  if (bn < 32 || bn % 32) return DDPX_ERR_SHAPE;
  const int cg = (pairs_enabled() && p.tpi == 1 && bn == 256 && Cout % 256 == 0) ? 2 : 1;
  p.block_n = bn;
  p.m_blocks = (Cout + kBlockM * cg - 1) / (kBlockM * cg);
  p.n_blocks = (Cin + bn - 1) / bn;
  p.k_steps_total = static_cast<int>(P / 32);
  const int tiles = p.m_blocks * p.n_blocks * ((9 + p.tpi - 1) / p.tpi);
  int splits = split_k;
  if (splits <= 0) {
    // ONE wave of work items: every extra split re-adds a whole [Cout][Cin][9] tile with scattered fp32 atomics
    // (512->512 @ 4x4 spent 3/4 of its time there with 9 splits), so split K only until every SM / CTA pair has an item.
    // DDPX_WGRAD_WAVES=n asks for n waves instead.
    static int waves = -1;
    if (waves < 0) { const char* e = getenv("DDPX_WGRAD_WAVES"); waves = e ? atoi(e) : 1; if (waves < 1) waves = 1; }
    const int units = num_sms() / cg;
    splits = waves * units / tiles;
    if (splits < 1) splits = 1;
  }
  if (splits > p.k_steps_total) splits = p.k_steps_total;
  p.k_steps_per_split = (p.k_steps_total + splits - 1) / splits;
  p.num_splits = (p.k_steps_total + p.k_steps_per_split - 1) / p.k_steps_per_split;
  p.H = H; p.W = W; p.box_w = bw; p.box_h = bh; p.box_n = bnn; p.cin_total = Cin; p.cin_blocks = Cin / 32;
  p.out = dw; p.atomic = 1; p.packed = packed;
  CUtensorMap ta, tb;
  static int big_cfg = -1;
  if (big_cfg < 0) { const char* e = getenv("DDPX_WGRAD_BIGBOX"); big_cfg = (e && e[0] == '0') ? 0 : 1; }
  p.big_box = (big_cfg && Cout % kBlockM == 0) ? 1 : 0;
  if (p.big_box) {
    // chunked views: (32 ch, pixels, Cout/32) and (32 ch, W, H, N, Cin/32); the chunk dimension has a 128-byte stride
    {
      uint64_t dims[3] = {32, static_cast<uint64_t>(P), static_cast<uint64_t>(Cout / 32)};
      uint64_t str[2] = {static_cast<uint64_t>(Cout) * 4, 128};
      uint32_t box[3] = {32, 32, static_cast<uint32_t>(kBlockM / 32)};
      if ((rc = make_tmap(&ta, dy, false, 3, dims, str, box, true)) != DDPX_OK) { big_cfg = 0; p.big_box = 0; }
    }
    if (p.big_box) {
      uint64_t dims[5] = {32, static_cast<uint64_t>(W), static_cast<uint64_t>(H), static_cast<uint64_t>(Nimg),
                          static_cast<uint64_t>(Cin / 32)};
      uint64_t str[4] = {static_cast<uint64_t>(Cin) * 4, static_cast<uint64_t>(Cin) * W * 4,
                         static_cast<uint64_t>(Cin) * W * H * 4, 128};
      uint32_t box[5] = {32, static_cast<uint32_t>(bw), static_cast<uint32_t>(bh), static_cast<uint32_t>(bnn),
                         static_cast<uint32_t>(bn / cg / 32)};
      if ((rc = make_tmap(&tb, x, false, 5, dims, str, box, true)) != DDPX_OK) { big_cfg = 0; p.big_box = 0; }
    }
    if (p.big_box) return launch_tc<MODE_WGRAD, false>(ta, tb, p, stream, cg);
  }
  {
    uint64_t dims[2] = {static_cast<uint64_t>(Cout), static_cast<uint64_t>(P)};
    uint64_t str[1] = {static_cast<uint64_t>(Cout) * 4};
    uint32_t box[2] = {32, 32};
    if ((rc = make_tmap(&ta, dy, false, 2, dims, str, box, true)) != DDPX_OK) return rc;
  }
  {
    uint64_t dims[4] = {static_cast<uint64_t>(Cin), static_cast<uint64_t>(W), static_cast<uint64_t>(H),
                        static_cast<uint64_t>(Nimg)};
    uint64_t str[3] = {static_cast<uint64_t>(Cin) * 4, static_cast<uint64_t>(Cin) * W * 4,
                       static_cast<uint64_t>(Cin) * W * H * 4};
    uint32_t box[4] = {32, static_cast<uint32_t>(bw), static_cast<uint32_t>(bh), static_cast<uint32_t>(bnn)};
    if ((rc = make_tmap(&tb, x, false, 4, dims, str, box, true)) != DDPX_OK) return rc;
  }
  return launch_tc<MODE_WGRAD, false>(ta, tb, p, stream, cg);
}

extern "C" int ddpx_conv3x3_wgrad(const float* dy, const float* x, float* dw, int Nimg, int H, int W, int Cin,
                                  int Cout, int split_k, void* stream_) {
  return conv_wgrad_impl(dy, x, dw, Nimg, H, W, Cin, Cout, split_k, 0, stream_);
}

// Same gradient accumulated into the tap-major staging layout dw_packed[9][Cout][Cin] (zeroed by the caller) with
// 16-byte vector reductions; ddpx_unpack_conv_wgrad turns it into the torch layout [Cout][Cin][3][3].
extern "C" int ddpx_conv3x3_wgrad_packed(const float* dy, const float* x, float* dw_packed, int Nimg, int H, int W, int Cin,
                                         int Cout, int split_k, void* stream_) {
  return conv_wgrad_impl(dy, x, dw_packed, Nimg, H, W, Cin, Cout, split_k, 1, stream_);
}

extern "C" unsigned int ddpx_read_error_word(void) {
  unsigned int v = 0;
  cudaMemcpyFromSymbol(&v, g_ddpx_error_word, sizeof(v));
  return v;
}
