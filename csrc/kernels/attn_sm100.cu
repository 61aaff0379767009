// Fused causal attention (flash style) on tcgen05 / TMEM / TMA for sm_100a, bf16, head dimension 64.
//
// No [B,H,T,T] tensor exists in either direction:
//   forward   work item = (batch, head, 128-query tile):  S = Q K^T lands in tensor memory, 128 threads (one per query
//             row = TMEM lane) run the online softmax, P goes back into tensor memory as bf16 and is the A operand of
//             O += P V (TS-mode tcgen05.mma: no shared-memory round trip for P); output + per-row log-sum-exp.
//   backward  probabilities are recomputed from the saved log-sum-exp.  Kernel "dkv" (work item = 128-key tile, K and V
//             resident in tensor memory) walks the query tiles and accumulates dK, dV; kernel "dq" (work item =
//             128-query tile, Q and dO resident in tensor memory) walks the key tiles and accumulates dQ.  Every
//             contraction whose left operand lives in tensor memory is a TS-mode MMA.
// All three kernels are persistent: a fixed number of CTAs (two per SM forward, one per SM backward) walk the items,
// heaviest first, and the TMA producer runs ahead into the next item.
//
// The reference is a CNN (/root/reference/multigpu.py:36-71) and has no attention; this serves the GPT-2 configurations
// of BASELINE.json, where torch's SDPA/flash kernels are the bar.
#include "attn_common.cuh"

#include <stdio.h>

namespace ddpx {

// =====================================================================================================================
// forward
// =====================================================================================================================
// 192 threads: warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2-5 = softmax (row = 32 * (warp % 4) + lane).
// Tensor memory (256 columns, so two CTAs share an SM and one CTA's softmax overlaps the other's MMAs):
//   [0,128) S fp32   [128,192) P bf16 (two per column)   [192,256) O fp32
// PERSISTENT: two CTAs per SM walk their shares of the (batch, head, query tile) items (heaviest first, dealt in a
// snake); barriers and tensor memory are set up once, the K / V stages run on across items and the next item's Q tile is
// fetched (second Q buffer) while the current item is computed - with one CTA per item every item paid launch,
// tcgen05.alloc, barrier set-up and a cold round trip for Q and the first K tile before its first MMA.
constexpr int kFwdThreads = 192;
constexpr uint32_t kColS = 0, kColP = 128, kColO = 192;

__global__ void __launch_bounds__(kFwdThreads, 2)
attn_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                        // 2 buffers (item n uses n & 1)
  uint8_t* sK = smem + 2 * kTileBytes;       // 2 stages
  uint8_t* sV = smem + 4 * kTileBytes;       // 2 stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * kTileBytes);
  uint64_t* q_full = bars;         // [2]
  uint64_t* q_empty = bars + 2;    // [2]
  uint64_t* k_full = bars + 4;     // [2]
  uint64_t* k_empty = bars + 6;    // [2]
  uint64_t* v_full = bars + 8;     // [2]
  uint64_t* v_empty = bars + 10;   // [2]
  uint64_t* s_full = bars + 12;
  uint64_t* p_full = bars + 13;
  uint64_t* o_done = bars + 14;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 15);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int bh_count = p.B * p.H;
  const int nitems = p.nq * bh_count;
  // item -> (query tile, batch, head); heaviest query tiles (most key blocks under the causal mask) first
  auto decode = [&](int idx, int& qt, int& b, int& h) {
    qt = p.nq - 1 - idx / bh_count;
    const int bh = idx % bh_count;
    b = bh / p.H; h = bh - b * p.H;
  };

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmQKV);
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < 2; ++s) {
        mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1);
        mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1);
        mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1);
      }
      mbar_init(s_full, 1);
      mbar_init(p_full, 4);          // one arrival per softmax warp
      mbar_init(o_done, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 256);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================================================ TMA producer ================================================
    if (elect_one()) {
      auto load_q = [&](int idx, int n) {
        int qt, b, h;
        decode(idx, qt, b, h);
        const int qb = n & 1;
        mbar_wait(&q_empty[qb], ((n >> 1) & 1) ^ 1u, 0x580u + qb);
        mbar_arrive_expect_tx(&q_full[qb], kTileBytes);
        tma_load_3d(sQ + qb * kTileBytes, &tmQKV, &q_full[qb], h * kHD, qt * kTile, b);
      };
      int g = 0;                                       // key blocks loaded so far (stage = g & 1)
      int idx = snake_item(0, nitems);
      if (idx >= 0) load_q(idx, 0);
      for (int n = 0; idx >= 0; ++n) {
        int qt, b, h;
        decode(idx, qt, b, h);
        const int nkv = qt + 1;
        const int next = snake_item(n + 1, nitems);
        for (int j = 0; j < nkv; ++j, ++g) {
          const int s = g & 1;
          const uint32_t ph = (g >> 1) & 1;
          mbar_wait(&k_empty[s], ph ^ 1u, 0x500u + s);
          mbar_arrive_expect_tx(&k_full[s], kTileBytes);
          tma_load_3d(sK + s * kTileBytes, &tmQKV, &k_full[s], p.C + h * kHD, j * kTile, b);
          mbar_wait(&v_empty[s], ph ^ 1u, 0x510u + s);
          mbar_arrive_expect_tx(&v_full[s], kTileBytes);
          tma_load_3d(sV + s * kTileBytes, &tmQKV, &v_full[s], 2 * p.C + h * kHD, j * kTile, b);
          if (j == 0 && next >= 0) load_q(next, n + 1);      // the other Q buffer: free since item n - 1's last Q K^T
        }
        idx = next;
      }
    }
  } else if (warp == 1) {
    // ================================================ MMA issuer ==================================================
    if (elect_one()) {
      const uint32_t idesc_qk = make_idesc(true, false, false, kTile, kTile);   // S[128 q][128 kv]
      const uint32_t idesc_pv = make_idesc(true, false, true, kTile, kHD);      // O[128 q][64], V is MN-major
      int g = 0;                 // key blocks before this item
      int idx = snake_item(0, nitems);
      for (int n = 0; idx >= 0; ++n, idx = snake_item(n, nitems)) {
        int qt, b, h;
        decode(idx, qt, b, h);
        const int nkv = qt + 1;
        const int qb = n & 1;
        const uint32_t aQ = smem_u32(sQ + qb * kTileBytes);
        mbar_wait(&q_full[qb], (n >> 1) & 1, 0x520u + qb);
        for (int j = 0; j <= nkv; ++j) {
          const int gj = g + j;
          if (j < nkv) mbar_wait(&k_full[gj & 1], (gj >> 1) & 1, 0x530u + (gj & 1));
          if (j > 0) {
            // O += P(j-1) V(j-1): the softmax warps have finished reading S and writing P
            const int gp = gj - 1;
            mbar_wait(p_full, gp & 1, 0x540u);
            mbar_wait(&v_full[gp & 1], (gp >> 1) & 1, 0x550u + (gp & 1));
            tc_fence_after();
            const uint32_t aV = smem_u32(sV + (gp & 1) * kTileBytes);
#pragma unroll
            for (int kk = 0; kk < kTile / 16; ++kk)
              tc_mma_ts_bf16(tmem + kColO, tmem + kColP + kk * 8, desc_mnmajor(aV, kk), idesc_pv, (j > 1 || kk > 0) ? 1u : 0u);
            tc_commit(&v_empty[gp & 1]);
          }
          if (j < nkv) {
            tc_fence_after();
            const uint32_t aK = smem_u32(sK + (gj & 1) * kTileBytes);
#pragma unroll
            for (int kk = 0; kk < kHD / 16; ++kk)
              tc_mma<true>(tmem + kColS, desc_kmajor(aQ, kk), desc_kmajor(aK, kk), idesc_qk, kk > 0 ? 1u : 0u);
            tc_commit(&k_empty[gj & 1]);
            if (j == nkv - 1) tc_commit(&q_empty[qb]);     // the item's last read of its Q tile
            tc_commit(s_full);            // also covers the P V MMAs issued just before: O is stable when S is ready
          } else {
            tc_commit(o_done);
          }
        }
        g += nkv;
      }
    }
  } else {
    // ================================================ softmax =====================================================
    const int quarter = warp & 3;
    const int r = quarter * 32 + lane;                  // row inside the tile == TMEM lane
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t tS = tmem + lane_base + kColS, tP = tmem + lane_base + kColP, tO = tmem + lane_base + kColO;
    const float sl2 = p.scale_log2;
    int g = 0;
    int idx = snake_item(0, nitems);
    for (int n = 0; idx >= 0; ++n, idx = snake_item(n, nitems)) {
      int qt, b, h;
      decode(idx, qt, b, h);
      const int nkv = qt + 1;
      float m = -INFINITY, l = 0.f;                       // running maximum (raw scores) and sum
      for (int j = 0; j < nkv; ++j) {
        mbar_wait(s_full, (g + j) & 1, 0x560u);
        tc_fence_after();
        const bool diag = (j == qt);                      // only the last block crosses the diagonal
        // Both passes over the 128 scores are software-pipelined: the tcgen05.ld of chunk c+1 is in flight while chunk c
        // is processed; maxima and sums run as independent chains.
        uint32_t va[32], vb[32];
        float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
        tmem_ld_32x32(tS, va);
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t (&v)[32] = (c & 1) ? vb : va;
          tmem_ld_wait();
          if (c < 3) tmem_ld_32x32(tS + (c + 1) * 32, (c & 1) ? va : vb);
          if (diag) {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              if (c * 32 + i <= r) mx0 = fmaxf(mx0, __uint_as_float(v[i]));
              if (c * 32 + i + 1 <= r) mx1 = fmaxf(mx1, __uint_as_float(v[i + 1]));
              if (c * 32 + i + 2 <= r) mx2 = fmaxf(mx2, __uint_as_float(v[i + 2]));
              if (c * 32 + i + 3 <= r) mx3 = fmaxf(mx3, __uint_as_float(v[i + 3]));
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 4) {
              mx0 = fmaxf(mx0, __uint_as_float(v[i]));
              mx1 = fmaxf(mx1, __uint_as_float(v[i + 1]));
              mx2 = fmaxf(mx2, __uint_as_float(v[i + 2]));
              mx3 = fmaxf(mx3, __uint_as_float(v[i + 3]));
            }
          }
        }
        // the second pass starts with chunk 0 again: issue its load before the bookkeeping
        tmem_ld_32x32(tS, va);
        const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
        const float m_new = fmaxf(m, mx);
        if (j > 0) {
          // the previous blocks were accumulated relative to the old maximum: rescale the sum and (only if some row of
          // this warp moved) the output accumulator
          const float alpha = ex2((m - m_new) * sl2);
          l *= alpha;
          if (__any_sync(0xffffffffu, m_new > m)) {
            uint32_t o[32];
            tmem_ld_wait();                       // chunk 0 of the scores (in va) and nothing else is outstanding
            tmem_ld_32x32(tO, o);
            tmem_ld_32x32(tO + 32, vb);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
            tmem_st_32x32(tO, o);
#pragma unroll
            for (int i = 0; i < 32; ++i) vb[i] = __float_as_uint(__uint_as_float(vb[i]) * alpha);
            tmem_st_32x32(tO + 32, vb);
          }
        }
        m = m_new;
        const float moff = m * sl2;
        float l0 = 0.f, l1 = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          uint32_t (&v)[32] = (c & 1) ? vb : va;
          tmem_ld_wait();
          if (c < 3) tmem_ld_32x32(tS + (c + 1) * 32, (c & 1) ? va : vb);
          uint32_t pk[16];
          if (!diag) {          // the diagonal block's masking is its own code path: no compare / select on the other blocks
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const float p0 = ex2(fmaf(__uint_as_float(v[i]), sl2, -moff));
              const float p1 = ex2(fmaf(__uint_as_float(v[i + 1]), sl2, -moff));
              l0 += p0; l1 += p1;
              pk[i >> 1] = pack_bf16(p0, p1);
            }
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              float p0 = ex2(fmaf(__uint_as_float(v[i]), sl2, -moff));
              float p1 = ex2(fmaf(__uint_as_float(v[i + 1]), sl2, -moff));
              if (c * 32 + i > r) p0 = 0.f;
              if (c * 32 + i + 1 > r) p1 = 0.f;
              l0 += p0; l1 += p1;
              pk[i >> 1] = pack_bf16(p0, p1);
            }
          }
          tmem_st_32x16(tP + c * 16, pk);
        }
        l += l0 + l1;
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(p_full);
      }
      // ---- epilogue: O / l -> bf16, log-sum-exp
      mbar_wait(o_done, n & 1, 0x570u);
      tc_fence_after();
      const int q = qt * kTile + r;
      const float inv = 1.f / l;
      uint4 ov[8];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t o[32];
        tmem_ld_32x32(tO + c * 32, o);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          ov[c * 4 + i].x = pack_bf16(__uint_as_float(o[8 * i]) * inv, __uint_as_float(o[8 * i + 1]) * inv);
          ov[c * 4 + i].y = pack_bf16(__uint_as_float(o[8 * i + 2]) * inv, __uint_as_float(o[8 * i + 3]) * inv);
          ov[c * 4 + i].z = pack_bf16(__uint_as_float(o[8 * i + 4]) * inv, __uint_as_float(o[8 * i + 5]) * inv);
          ov[c * 4 + i].w = pack_bf16(__uint_as_float(o[8 * i + 6]) * inv, __uint_as_float(o[8 * i + 7]) * inv);
        }
      }
      if (q < p.T) {
        uint4* dst = reinterpret_cast<uint4*>(p.out + (static_cast<long long>(b) * p.T + q) * p.C + h * kHD);
#pragma unroll
        for (int i = 0; i < 8; ++i) dst[i] = ov[i];
        p.lse2[(static_cast<long long>(b) * p.H + h) * p.Tp + q] = m * sl2 + log2f(l);
      }
      g += nkv;
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem, 256);
}

}  // namespace ddpx

// =====================================================================================================================
// host side
// =====================================================================================================================
using namespace ddpx;

static int attn_check(const void* qkv, int B, int T, int H) {
  if (B < 1 || T < 1 || H < 1) return DDPX_ERR_ARG;
  if (reinterpret_cast<uintptr_t>(qkv) & 15) return DDPX_ERR_ALIGN;
  return DDPX_OK;
}

extern "C" int ddpx_attn_fwd(const void* qkv, void* out, float* lse2, int B, int T, int H, float scale, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  int rc = attn_check(qkv, B, T, H);
  if (rc != DDPX_OK) return rc;
  if (reinterpret_cast<uintptr_t>(out) & 15) return DDPX_ERR_ALIGN;
  AttnParams p{};
  p.B = B; p.T = T; p.H = H; p.C = H * kHD; p.nq = (T + kTile - 1) / kTile; p.Tp = p.nq * kTile;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  p.out = static_cast<__nv_bfloat16*>(out); p.lse2 = lse2; p.qkv = static_cast<const __nv_bfloat16*>(qkv);
  CUtensorMap tm;
  if ((rc = attn_map(&tm, qkv, B, T, 3 * p.C)) != DDPX_OK) return rc;
  const size_t smem = 6 * kTileBytes + 1024 + 128;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(attn_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)) != cudaSuccess)
      return DDPX_ERR_LAUNCH;
    attr_set = true;
  }
  const long long nitems = static_cast<long long>(p.nq) * B * H;
  if (nitems > 0x7fffffffLL) return DDPX_ERR_SHAPE;
  // persistent: two CTAs per SM (256 tensor-memory columns each), every CTA walks its share of the items
  const long long slots = 2LL * num_sms();
  attn_fwd_kernel<<<static_cast<unsigned>(nitems < slots ? nitems : slots), kFwdThreads, smem, stream>>>(tm, p);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

// bounded waits of this translation unit report here (the error word is per translation unit)
extern "C" unsigned int ddpx_attn_read_error_word(void) {
  unsigned int v = 0;
  cudaMemcpyFromSymbol(&v, g_ddpx_error_word, sizeof(v));
  return v;
}
