// Backward of the fused causal attention (see attn_sm100.cu for the forward and the overall scheme).
//
// Two kernels, both recompute P = exp2(S * scale_log2 - lse2) from the saved log-sum-exp, neither uses atomics:
//   attn_bwd_dq_kernel   work item = (batch, head, 128-query tile).  Q_i and dO_i are written ONCE into tensor memory
//                        (bf16, K-major) by the threads that own the rows, so S = Q K^T and dP = dO V^T are TS-mode MMAs
//                        against the K / V tiles TMA streams in; dS goes back into tensor memory and dQ += dS K is a third
//                        TS-mode MMA.  The same threads compute delta = rowsum(dO * O) on the way and publish it.
//   attn_bwd_dkv_kernel  work item = (batch, head, 128-key tile).  K_j and V_j live in tensor memory; S^T = K Q^T and
//                        dP^T = V dO^T put keys on the TMEM lanes, so P^T and dS^T are already the A operands of
//                        dV += P^T dO and dK += dS^T Q (Q / dO tiles are read K-major for the first pair and MN-major for
//                        the second - the same shared-memory bytes).
// Both kernels are PERSISTENT: one CTA per SM walks its share of the work items (heaviest first, dealt in a snake so the
// shares are even), tensor memory and barriers are set up once, and the TMA producer runs ahead into the next item - its
// resident rows (Q, dO, O / K, V) arrive in shared memory while the current item is still being computed.  With one CTA
// per item every item paid launch, tcgen05.alloc, barrier set-up and a cold global-memory round trip for its rows before
// the first MMA (ncu: the compute warps spent 17 % of their samples on those loads and 26 % waiting for the first scores).
// 320 threads: warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2-9 = compute (two warps per TMEM lane
// quarter, each working on half of a step's 64 columns).
#include "attn_common.cuh"

namespace ddpx {

constexpr int kBwdThreads = 320;

// row r (64 bf16 = 128 B) of a K-major tile as TMA wrote it with the 128-byte swizzle: 16-byte piece c sits at c ^ (r & 7)
__device__ __forceinline__ void read_row_sw128(const uint8_t* tile, int r, uint32_t (&v)[32]) {
  const uint8_t* row = tile + r * 128;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const uint4 u = *reinterpret_cast<const uint4*>(row + ((c ^ (r & 7)) << 4));
    v[4 * c] = u.x; v[4 * c + 1] = u.y; v[4 * c + 2] = u.z; v[4 * c + 3] = u.w;
  }
}
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }

// 1-D bulk copy global -> shared, completion on an mbarrier (used for the 512-byte lse2 / delta rows of a query tile)
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(reinterpret_cast<uint64_t>(src)), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

// =====================================================================================================================
// dQ
// =====================================================================================================================
// tensor memory (384 of the 512 columns): [0,32) Q  [32,64) dO  [64,192) S, two buffers of 64 columns  [192,320) dP, two
// buffers of 64 columns (dS overlays the dP columns a warp has consumed)  [320,384) dQ
// A 128-key tile in shared memory is consumed as two halves of 64 keys ("steps").  S / dP of step t+1 are issued before
// the MMA thread waits for dS of step t, so the compute warps find their next scores ready: with one buffer the tensor
// pipe and the warps took turns and every step paid two barrier round trips.
namespace dq {
constexpr uint32_t kColQ = 0, kColDO = 32, kColS = 64, kColDP = 192, kColDQ = 320;
}
constexpr int kSub = 64;                     // keys (dq kernel) / queries (dkv kernel) per step
constexpr uint32_t kSubBytes = kSub * 128;   // the step's rows inside a K-major 128-byte-swizzled tile

__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO,
                   const __grid_constant__ CUtensorMap tmO, const AttnParams p) {
  using namespace dq;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;                        // 2 stages
  uint8_t* sV = smem + 2 * kTileBytes;       // 2 stages
  uint8_t* sRowQ = smem + 4 * kTileBytes;    // the item's resident rows: Q, dO, O tiles and the lse2 row
  uint8_t* sRowDO = smem + 5 * kTileBytes;
  uint8_t* sRowO = smem + 6 * kTileBytes;
  float* sLse = reinterpret_cast<float*>(smem + 7 * kTileBytes);      // [128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sLse + kTile);
  uint64_t* kv_full = bars;        // [2]
  uint64_t* kv_empty = bars + 2;   // [2]
  uint64_t* rows_full = bars + 4;  // the item's rows are in shared memory (TMA)
  uint64_t* rows_empty = bars + 5; // ... and have been read (8 compute warps)
  uint64_t* rows_ready = bars + 6; // Q / dO are in tensor memory (8 compute warps)
  uint64_t* sdp_full = bars + 7;   // [2] S and dP of a step complete
  uint64_t* ds_full = bars + 9;    // [2] dS of a step written (8 compute warps)
  uint64_t* acc_done = bars + 11;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int bh_count = p.B * p.H;
  const int nitems = p.nq * bh_count;
  // item -> (query tile, batch, head); heaviest (most key tiles) first
  auto decode = [&](int idx, int& qt, int& b, int& h) {
    qt = p.nq - 1 - idx / bh_count;
    const int bh = idx % bh_count;
    b = bh / p.H; h = bh - b * p.H;
  };

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmDO); tma_prefetch_desc(&tmO); }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < 2; ++s) {
        mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1);
        mbar_init(&sdp_full[s], 1); mbar_init(&ds_full[s], 8);
      }
      mbar_init(rows_full, 1);
      mbar_init(rows_empty, 8);
      mbar_init(rows_ready, 8);
      mbar_init(acc_done, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================================================ TMA producer ================================================
    if (elect_one()) {
      // the rows of the n-th item of this CTA (one buffer: refilled as soon as the compute warps have read it)
      auto load_rows = [&](int idx, int n) {
        int qt, b, h;
        decode(idx, qt, b, h);
        mbar_wait(rows_empty, (n & 1) ^ 1u, 0x660u);
        mbar_arrive_expect_tx(rows_full, 3 * kTileBytes + kTile * 4);
        tma_load_3d(sRowQ, &tmQKV, rows_full, h * kHD, qt * kTile, b);
        tma_load_3d(sRowDO, &tmDO, rows_full, h * kHD, qt * kTile, b);
        tma_load_3d(sRowO, &tmO, rows_full, h * kHD, qt * kTile, b);
        bulk_load(sLse, p.lse2 + (static_cast<long long>(b) * p.H + h) * p.Tp + qt * kTile, kTile * 4, rows_full);
      };
      int g = 0;                                       // key tiles loaded so far (stage = g & 1)
      int idx = snake_item(0, nitems);
      if (idx >= 0) load_rows(idx, 0);
      for (int n = 0; idx >= 0; ++n) {
        int qt, b, h;
        decode(idx, qt, b, h);
        const int nkv = qt + 1;
        const int next = snake_item(n + 1, nitems);
        for (int j = 0; j < nkv; ++j, ++g) {
          const int s = g & 1;
          mbar_wait(&kv_empty[s], ((g >> 1) & 1) ^ 1u, 0x600u + s);
          mbar_arrive_expect_tx(&kv_full[s], 2 * kTileBytes);
          tma_load_3d(sK + s * kTileBytes, &tmQKV, &kv_full[s], p.C + h * kHD, j * kTile, b);
          tma_load_3d(sV + s * kTileBytes, &tmQKV, &kv_full[s], 2 * p.C + h * kHD, j * kTile, b);
          // once this item's first tiles are on their way: the next item's rows (their buffer is free as soon as the
          // compute warps have copied this item's rows out, right at the start of the item)
          if (j == (nkv > 1 ? 1 : 0) && next >= 0) load_rows(next, n + 1);
        }
        idx = next;
      }
    }
  } else if (warp == 1) {
    // ================================================ MMA issuer ==================================================
    if (elect_one()) {
      const uint32_t idesc_s = make_idesc(true, false, false, kTile, kSub);     // [128 q][64 kv]
      const uint32_t idesc_dq = make_idesc(true, false, true, kTile, kHD);      // [128 q][64], K tile read MN-major
      int g = 0;                 // key tiles consumed before this item
      uint32_t T = 0;            // steps issued before this item (buffer = step & 1; always even here)
      int idx = snake_item(0, nitems);
      for (int n = 0; idx >= 0; ++n, idx = snake_item(n, nitems)) {
        int qt, b, h;
        decode(idx, qt, b, h);
        const int nkv = qt + 1, nsub = 2 * nkv;
        mbar_wait(rows_ready, n & 1, 0x610u);
        tc_fence_after();
        // S = Q K^T and dP = dO V^T of step t into buffer (T + t) & 1
        auto issue_sdp = [&](int t) {
          const int gj = g + (t >> 1), hh = t & 1, s = gj & 1;
          const uint32_t st = T + static_cast<uint32_t>(t);
          const uint32_t buf = (st & 1u) * kSub;
          if (hh == 0) {
            mbar_wait(&kv_full[s], (gj >> 1) & 1, 0x620u + s);
            tc_fence_after();
          }
          const uint32_t aK = smem_u32(sK + s * kTileBytes) + hh * kSubBytes, aV = smem_u32(sV + s * kTileBytes) + hh * kSubBytes;
#pragma unroll
          for (int kk = 0; kk < kHD / 16; ++kk)
            tc_mma_ts_bf16(tmem + kColS + buf, tmem + kColQ + kk * 8, desc_kmajor(aK, kk), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < kHD / 16; ++kk)
            tc_mma_ts_bf16(tmem + kColDP + buf, tmem + kColDO + kk * 8, desc_kmajor(aV, kk), idesc_s, kk > 0 ? 1u : 0u);
          tc_commit(&sdp_full[st & 1u]);
        };
        issue_sdp(0);
        for (int t = 0; t < nsub; ++t) {
          // buffer (t + 1) & 1 was last used by step t - 1: its dS was waited for one iteration ago, and the dQ MMAs that
          // read it are ahead of these in the (in-order) tensor pipe
          if (t + 1 < nsub) issue_sdp(t + 1);
          const int gj = g + (t >> 1), hh = t & 1, s = gj & 1;
          const uint32_t st = T + static_cast<uint32_t>(t);
          const uint32_t buf = (st & 1u) * kSub;
          mbar_wait(&ds_full[st & 1u], (st >> 1) & 1u, 0x630u + (st & 1u));
          tc_fence_after();
          const uint32_t aK = smem_u32(sK + s * kTileBytes);
          // dS columns (bf16 pairs): keys 0-31 of the step sit at [0,16), keys 32-63 at [32,48) of the buffer (each compute
          // warp overlays the columns it has read); the step's K rows are the 16-row slabs hh * 4 .. hh * 4 + 3 of the tile
#pragma unroll
          for (int kk = 0; kk < kSub / 16; ++kk)
            tc_mma_ts_bf16(tmem + kColDQ, tmem + kColDP + buf + (kk >> 1) * 32 + (kk & 1) * 8, desc_mnmajor(aK, hh * 4 + kk),
                           idesc_dq, (t > 0 || kk > 0) ? 1u : 0u);
          if (hh == 1) tc_commit(&kv_empty[s]);
        }
        tc_commit(acc_done);
        g += nkv;
        T += static_cast<uint32_t>(nsub);
      }
    }
  } else {
    // ================================================ compute warps ===============================================
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;                   // which 32 key columns of a step this warp handles
    const int r = quarter * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    const float sl2 = p.scale_log2;
    const uint32_t tS = tmem + lane_base + kColS + half * 32;
    const uint32_t tDP = tmem + lane_base + kColDP + half * 32;
    uint32_t T = 0;
    int idx = snake_item(0, nitems);
    for (int n = 0; idx >= 0; ++n, idx = snake_item(n, nitems)) {
      int qt, b, h;
      decode(idx, qt, b, h);
      const int nsub = 2 * (qt + 1);
      const int q = qt * kTile + r;
      const bool valid = q < p.T;
      // ---- the item's rows: delta = rowsum(dO * O) (both warps of the row need it; one publishes it for the dkv kernel),
      // half 0 warps bring Q into tensor memory, half 1 warps dO.  The previous item's MMAs are complete (acc_done).
      float delta = 0.f, lse2;
      {
        uint32_t rowv[32], o[32];
        mbar_wait(rows_full, n & 1, 0x670u);
        read_row_sw128(sRowDO, r, rowv);
        read_row_sw128(sRowO, r, o);
#pragma unroll
        for (int i = 0; i < 32; ++i) delta += bf_lo(rowv[i]) * bf_lo(o[i]) + bf_hi(rowv[i]) * bf_hi(o[i]);
        if (half == 0) read_row_sw128(sRowQ, r, rowv);
        lse2 = valid ? sLse[r] : 0.f;
        __syncwarp();
        if (lane == 0) mbar_arrive(rows_empty);
        // (published pre-multiplied by the softmax scale: the dkv kernel only ever uses delta * scale)
        if (half == 1) p.delta[(static_cast<long long>(b) * p.H + h) * p.Tp + q] = delta * p.scale;      // q < Tp always
        tmem_st_32x32(tmem + lane_base + (half == 0 ? kColQ : kColDO), rowv);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(rows_ready);
      }
      for (int t = 0; t < nsub; ++t) {
        const uint32_t st = T + static_cast<uint32_t>(t);
        const uint32_t buf = (st & 1u) * kSub;
        mbar_wait(&sdp_full[st & 1u], (st >> 1) & 1u, 0x640u + (st & 1u));
        tc_fence_after();
        const bool diag = ((t >> 1) == qt);
        const int col0 = (t & 1) * kSub + half * 32;        // first key of this warp's columns inside the 128-key tile
        uint32_t sv[32], dv[32];
        tmem_ld_32x32(tS + buf, sv);
        tmem_ld_32x32(tDP + buf, dv);
        tmem_ld_wait();
        // dS = P * (dP - delta) * scale = P * fma(dP, scale, -delta * scale).  The diagonal tile's masking is its own code
        // path (warp-uniform branch): folded into one loop it cost a compare and a select per element on every tile.
        uint32_t pk[16];
        const float nds = -delta * p.scale;
        if (!diag) {
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float p0 = ex2(fmaf(__uint_as_float(sv[i]), sl2, -lse2));
            const float p1 = ex2(fmaf(__uint_as_float(sv[i + 1]), sl2, -lse2));
            pk[i >> 1] = pack_bf16(p0 * fmaf(__uint_as_float(dv[i]), p.scale, nds), p1 * fmaf(__uint_as_float(dv[i + 1]), p.scale, nds));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            const float p0 = ex2(fmaf(__uint_as_float(sv[i]), sl2, -lse2));
            const float p1 = ex2(fmaf(__uint_as_float(sv[i + 1]), sl2, -lse2));
            float d0 = p0 * fmaf(__uint_as_float(dv[i]), p.scale, nds);
            float d1 = p1 * fmaf(__uint_as_float(dv[i + 1]), p.scale, nds);
            if (col0 + i > r) d0 = 0.f;
            if (col0 + i + 1 > r) d1 = 0.f;
            pk[i >> 1] = pack_bf16(d0, d1);
          }
        }
        tmem_st_32x16(tDP + buf, pk);           // overlays dP columns this warp has already consumed
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&ds_full[st & 1u]);
      }
      T += static_cast<uint32_t>(nsub);
      // ---- epilogue: each warp stores 32 of the row's 64 dQ columns
      mbar_wait(acc_done, n & 1, 0x650u);
      tc_fence_after();
      uint32_t a[32];
      tmem_ld_32x32(tmem + lane_base + kColDQ + half * 32, a);
      tmem_ld_wait();
      if (valid) {
        uint4* dst = reinterpret_cast<uint4*>(p.dqkv + (static_cast<long long>(b) * p.T + q) * 3 * p.C + h * kHD + half * 32);
#pragma unroll
        for (int i = 0; i < 4; ++i)
          dst[i] = make_uint4(pack_bf16(__uint_as_float(a[8 * i]), __uint_as_float(a[8 * i + 1])),
                              pack_bf16(__uint_as_float(a[8 * i + 2]), __uint_as_float(a[8 * i + 3])),
                              pack_bf16(__uint_as_float(a[8 * i + 4]), __uint_as_float(a[8 * i + 5])),
                              pack_bf16(__uint_as_float(a[8 * i + 6]), __uint_as_float(a[8 * i + 7])));
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// =====================================================================================================================
// dK, dV
// =====================================================================================================================
// tensor memory (448 of the 512 columns): [0,32) K  [32,64) V  [64,192) S^T, two buffers of 64 columns (P^T overlays)
//   [192,320) dP^T, two buffers of 64 columns (dS^T overlays)  [320,384) dV  [384,448) dK
// A 128-query tile in shared memory is consumed as two steps of 64 queries, double-buffered like the dq kernel.
namespace dkv {
constexpr uint32_t kColK = 0, kColV = 32, kColS = 64, kColDP = 192, kColDV = 320, kColDK = 384;
}

__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQKV, const __grid_constant__ CUtensorMap tmDO, const AttnParams p) {
  using namespace dkv;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                        // 2 stages
  uint8_t* sDO = smem + 2 * kTileBytes;      // 2 stages
  uint8_t* sRowK = smem + 4 * kTileBytes;    // the item's resident rows: K and V tiles
  uint8_t* sRowV = smem + 5 * kTileBytes;
  float* sLse = reinterpret_cast<float*>(smem + 6 * kTileBytes);      // [2][128]
  float* sDelta = sLse + 2 * kTile;                                   // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDelta + 2 * kTile);
  uint64_t* q_full = bars;         // [2]
  uint64_t* q_empty = bars + 2;    // [2]
  uint64_t* rows_full = bars + 4;
  uint64_t* rows_empty = bars + 5;
  uint64_t* rows_ready = bars + 6;
  uint64_t* sdp_full = bars + 7;   // [2]
  uint64_t* pds_full = bars + 9;   // [2]
  uint64_t* acc_done = bars + 11;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 12);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int bh_count = p.B * p.H;
  const int nitems = p.nq * bh_count;
  // item -> (key tile, batch, head); key tile 0 meets every query tile: heaviest first
  auto decode = [&](int idx, int& kt, int& b, int& h) {
    kt = idx / bh_count;
    const int bh = idx % bh_count;
    b = bh / p.H; h = bh - b * p.H;
  };

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmDO); }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < 2; ++s) {
        mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1);
        mbar_init(&sdp_full[s], 1); mbar_init(&pds_full[s], 8);
      }
      mbar_init(rows_full, 1);
      mbar_init(rows_empty, 8);
      mbar_init(rows_ready, 8);
      mbar_init(acc_done, 1);
      fence_barrier_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot;

  if (warp == 0) {
    // ================================================ TMA producer ================================================
    if (elect_one()) {
      auto load_rows = [&](int idx, int n) {
        int kt, b, h;
        decode(idx, kt, b, h);
        mbar_wait(rows_empty, (n & 1) ^ 1u, 0x770u);
        mbar_arrive_expect_tx(rows_full, 2 * kTileBytes);
        tma_load_3d(sRowK, &tmQKV, rows_full, p.C + h * kHD, kt * kTile, b);
        tma_load_3d(sRowV, &tmQKV, rows_full, 2 * p.C + h * kHD, kt * kTile, b);
      };
      int g = 0;                                       // query tiles loaded so far (stage = g & 1)
      int idx = snake_item(0, nitems);
      if (idx >= 0) load_rows(idx, 0);
      for (int n = 0; idx >= 0; ++n) {
        int kt, b, h;
        decode(idx, kt, b, h);
        const int nqt = p.nq - kt;                     // query tiles kt .. nq-1
        const int next = snake_item(n + 1, nitems);
        const long long stat_row = (static_cast<long long>(b) * p.H + h) * p.Tp;
        for (int it = 0; it < nqt; ++it, ++g) {
          const int s = g & 1, i = kt + it;
          mbar_wait(&q_empty[s], ((g >> 1) & 1) ^ 1u, 0x700u + s);
          mbar_arrive_expect_tx(&q_full[s], 2 * kTileBytes + 2 * kTile * 4);
          tma_load_3d(sQ + s * kTileBytes, &tmQKV, &q_full[s], h * kHD, i * kTile, b);
          tma_load_3d(sDO + s * kTileBytes, &tmDO, &q_full[s], h * kHD, i * kTile, b);
          bulk_load(sLse + s * kTile, p.lse2 + stat_row + i * kTile, kTile * 4, &q_full[s]);
          bulk_load(sDelta + s * kTile, p.delta + stat_row + i * kTile, kTile * 4, &q_full[s]);
          if (it == (nqt > 1 ? 1 : 0) && next >= 0) load_rows(next, n + 1);
        }
        idx = next;
      }
    }
  } else if (warp == 1) {
    // ================================================ MMA issuer ==================================================
    if (elect_one()) {
      const uint32_t idesc_s = make_idesc(true, false, false, kTile, kSub);     // [128 kv][64 q]
      const uint32_t idesc_acc = make_idesc(true, false, true, kTile, kHD);     // [128 kv][64], Q / dO tiles read MN-major
      int g = 0;
      uint32_t T = 0;
      int idx = snake_item(0, nitems);
      for (int n = 0; idx >= 0; ++n, idx = snake_item(n, nitems)) {
        int kt, b, h;
        decode(idx, kt, b, h);
        const int nqt = p.nq - kt, nsub = 2 * nqt;
        mbar_wait(rows_ready, n & 1, 0x710u);
        tc_fence_after();
        // S^T = K Q^T and dP^T = V dO^T of step t (queries hh * 64 .. of the tile) into buffer (T + t) & 1
        auto issue_sdp = [&](int t) {
          const int gj = g + (t >> 1), hh = t & 1, s = gj & 1;
          const uint32_t st = T + static_cast<uint32_t>(t);
          const uint32_t buf = (st & 1u) * kSub;
          if (hh == 0) {
            mbar_wait(&q_full[s], (gj >> 1) & 1, 0x720u + s);
            tc_fence_after();
          }
          const uint32_t aQ = smem_u32(sQ + s * kTileBytes) + hh * kSubBytes, aDO = smem_u32(sDO + s * kTileBytes) + hh * kSubBytes;
#pragma unroll
          for (int kk = 0; kk < kHD / 16; ++kk)
            tc_mma_ts_bf16(tmem + kColS + buf, tmem + kColK + kk * 8, desc_kmajor(aQ, kk), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < kHD / 16; ++kk)
            tc_mma_ts_bf16(tmem + kColDP + buf, tmem + kColV + kk * 8, desc_kmajor(aDO, kk), idesc_s, kk > 0 ? 1u : 0u);
          tc_commit(&sdp_full[st & 1u]);
        };
        issue_sdp(0);
        for (int t = 0; t < nsub; ++t) {
          if (t + 1 < nsub) issue_sdp(t + 1);        // its buffer was released by step t - 1 (see the dq kernel)
          const int gj = g + (t >> 1), hh = t & 1, s = gj & 1;
          const uint32_t st = T + static_cast<uint32_t>(t);
          const uint32_t buf = (st & 1u) * kSub;
          mbar_wait(&pds_full[st & 1u], (st >> 1) & 1u, 0x730u + (st & 1u));
          tc_fence_after();
          const uint32_t aQ = smem_u32(sQ + s * kTileBytes), aDO = smem_u32(sDO + s * kTileBytes);
          // P^T / dS^T columns (bf16 pairs): queries 0-31 of the step at [0,16), 32-63 at [32,48) of the buffer
#pragma unroll
          for (int kk = 0; kk < kSub / 16; ++kk)
            tc_mma_ts_bf16(tmem + kColDV, tmem + kColS + buf + (kk >> 1) * 32 + (kk & 1) * 8, desc_mnmajor(aDO, hh * 4 + kk),
                           idesc_acc, (t > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
          for (int kk = 0; kk < kSub / 16; ++kk)
            tc_mma_ts_bf16(tmem + kColDK, tmem + kColDP + buf + (kk >> 1) * 32 + (kk & 1) * 8, desc_mnmajor(aQ, hh * 4 + kk),
                           idesc_acc, (t > 0 || kk > 0) ? 1u : 0u);
          if (hh == 1) tc_commit(&q_empty[s]);
        }
        tc_commit(acc_done);
        g += nqt;
        T += static_cast<uint32_t>(nsub);
      }
    }
  } else {
    // ================================================ compute warps ===============================================
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;                   // which 32 query columns of a step this warp handles
    const int r = quarter * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    const float sl2 = p.scale_log2;
    const uint32_t tS = tmem + lane_base + kColS + half * 32;
    const uint32_t tDP = tmem + lane_base + kColDP + half * 32;
    int g = 0;
    uint32_t T = 0;
    int idx = snake_item(0, nitems);
    for (int n = 0; idx >= 0; ++n, idx = snake_item(n, nitems)) {
      int kt, b, h;
      decode(idx, kt, b, h);
      const int nqt = p.nq - kt, nsub = 2 * nqt;
      const int kv = kt * kTile + r;
      const bool valid = kv < p.T;
      {
        // resident operands: half 0 warps bring the K rows, half 1 warps the V rows (rows past T arrive as zeros)
        uint32_t rowv[32];
        mbar_wait(rows_full, n & 1, 0x780u);
        read_row_sw128(half == 0 ? sRowK : sRowV, r, rowv);
        __syncwarp();
        if (lane == 0) mbar_arrive(rows_empty);
        tmem_st_32x32(tmem + lane_base + (half == 0 ? kColK : kColV), rowv);
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(rows_ready);
      }
      for (int t = 0; t < nsub; ++t) {
        const int it = t >> 1, gj = g + it, s = gj & 1, i = kt + it;
        const uint32_t st = T + static_cast<uint32_t>(t);
        const uint32_t buf = (st & 1u) * kSub;
        if ((t & 1) == 0) mbar_wait(&q_full[s], (gj >> 1) & 1, 0x740u + s);   // lse2 / delta of this query tile are in shared memory
        mbar_wait(&sdp_full[st & 1u], (st >> 1) & 1u, 0x750u + (st & 1u));
        tc_fence_after();
        // keys above the query (diagonal tile) and queries past the end of the sequence contribute nothing
        const bool masked = (it == 0) || ((i + 1) * kTile > p.T);
        const int col0 = (t & 1) * kSub + half * 32;          // first query row (inside the tile) of this warp's columns
        const float4* lse4 = reinterpret_cast<const float4*>(sLse + s * kTile + col0);
        const float4* del4 = reinterpret_cast<const float4*>(sDelta + s * kTile + col0);
        uint32_t sv[32], dv[32];
        tmem_ld_32x32(tS + buf, sv);
        tmem_ld_32x32(tDP + buf, dv);
        tmem_ld_wait();
        // sDelta holds delta * scale (the dq kernel publishes it that way): dS^T = P^T * fma(dP^T, scale, -delta * scale).
        // Masked tiles (the diagonal one, and the last one of a sequence that is not a multiple of 128) take their own path.
        uint32_t pk[16], dk[16];
        if (!masked) {
#pragma unroll
          for (int gq = 0; gq < 8; ++gq) {
            const float4 L = lse4[gq], D = del4[gq];
            const float ls[4] = {L.x, L.y, L.z, L.w}, ds[4] = {D.x, D.y, D.z, D.w};
            float pv[4], dsv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int x = gq * 4 + e;
              pv[e] = ex2(fmaf(__uint_as_float(sv[x]), sl2, -ls[e]));
              dsv[e] = pv[e] * fmaf(__uint_as_float(dv[x]), p.scale, -ds[e]);
            }
            pk[gq * 2] = pack_bf16(pv[0], pv[1]); pk[gq * 2 + 1] = pack_bf16(pv[2], pv[3]);
            dk[gq * 2] = pack_bf16(dsv[0], dsv[1]); dk[gq * 2 + 1] = pack_bf16(dsv[2], dsv[3]);
          }
        } else {
#pragma unroll
          for (int gq = 0; gq < 8; ++gq) {
            const float4 L = lse4[gq], D = del4[gq];
            const float ls[4] = {L.x, L.y, L.z, L.w}, ds[4] = {D.x, D.y, D.z, D.w};
            float pv[4], dsv[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int x = gq * 4 + e;
              pv[e] = ex2(fmaf(__uint_as_float(sv[x]), sl2, -ls[e]));
              dsv[e] = pv[e] * fmaf(__uint_as_float(dv[x]), p.scale, -ds[e]);
              const int col = col0 + x;                               // query row inside the tile
              if ((it == 0 && col < r) || i * kTile + col >= p.T) { pv[e] = 0.f; dsv[e] = 0.f; }
            }
            pk[gq * 2] = pack_bf16(pv[0], pv[1]); pk[gq * 2 + 1] = pack_bf16(pv[2], pv[3]);
            dk[gq * 2] = pack_bf16(dsv[0], dsv[1]); dk[gq * 2 + 1] = pack_bf16(dsv[2], dsv[3]);
          }
        }
        tmem_st_32x16(tS + buf, pk);        // P^T over the S^T columns this warp has consumed
        tmem_st_32x16(tDP + buf, dk);       // dS^T over the dP^T columns
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&pds_full[st & 1u]);
      }
      g += nqt;
      T += static_cast<uint32_t>(nsub);
      // ---- epilogue: half 0 warps store the dV row, half 1 warps the dK row
      mbar_wait(acc_done, n & 1, 0x760u);
      tc_fence_after();
      __nv_bfloat16* dst_row = p.dqkv + (static_cast<long long>(b) * p.T + (valid ? kv : 0)) * 3 * p.C + (half == 0 ? 2 : 1) * p.C + h * kHD;
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t a[32];
        tmem_ld_32x32(tmem + lane_base + (half == 0 ? kColDV : kColDK) + c * 32, a);
        tmem_ld_wait();
        if (valid) {
          uint4* dst = reinterpret_cast<uint4*>(dst_row + c * 32);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            dst[i] = make_uint4(pack_bf16(__uint_as_float(a[8 * i]), __uint_as_float(a[8 * i + 1])),
                                pack_bf16(__uint_as_float(a[8 * i + 2]), __uint_as_float(a[8 * i + 3])),
                                pack_bf16(__uint_as_float(a[8 * i + 4]), __uint_as_float(a[8 * i + 5])),
                                pack_bf16(__uint_as_float(a[8 * i + 6]), __uint_as_float(a[8 * i + 7])));
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

}  // namespace ddpx

// =====================================================================================================================
// host side
// =====================================================================================================================
using namespace ddpx;

// dqkv[B,T,3C] = gradient of the packed q|k|v tensor.  `delta` [B,H,Tp] (Tp = T rounded up to 128) is scratch the dq
// kernel fills and the dkv kernel reads; lse2 is the forward's.
extern "C" int ddpx_attn_bwd(const void* qkv, const void* out, const void* dout, const float* lse2, float* delta, void* dqkv,
                             int B, int T, int H, float scale, void* stream_) {
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (B < 1 || T < 1 || H < 1) return DDPX_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(dout) |
       reinterpret_cast<uintptr_t>(dqkv) | reinterpret_cast<uintptr_t>(lse2) | reinterpret_cast<uintptr_t>(delta)) & 15)
    return DDPX_ERR_ALIGN;
  AttnParams p{};
  p.B = B; p.T = T; p.H = H; p.C = H * kHD; p.nq = (T + kTile - 1) / kTile; p.Tp = p.nq * kTile;
  p.scale = scale; p.scale_log2 = scale * 1.4426950408889634f;
  p.qkv = static_cast<const __nv_bfloat16*>(qkv); p.o = static_cast<const __nv_bfloat16*>(out);
  p.dout = static_cast<const __nv_bfloat16*>(dout); p.lse2 = const_cast<float*>(lse2); p.delta = delta;
  p.dqkv = static_cast<__nv_bfloat16*>(dqkv);
  CUtensorMap tq, td, to;
  int rc;
  if ((rc = attn_map(&tq, qkv, B, T, 3 * p.C)) != DDPX_OK) return rc;
  if ((rc = attn_map(&td, dout, B, T, p.C)) != DDPX_OK) return rc;
  if ((rc = attn_map(&to, out, B, T, p.C)) != DDPX_OK) return rc;
  const size_t smem_dq = 7 * kTileBytes + kTile * 4 + 1024 + 128;
  const size_t smem_dkv = 6 * kTileBytes + 4 * kTile * 4 + 1024 + 128;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_dq)) != cudaSuccess ||
        cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_dkv)) != cudaSuccess)
      return DDPX_ERR_LAUNCH;
    attr_set = true;
  }
  const long long nitems = static_cast<long long>(p.nq) * B * H;
  if (nitems > 0x7fffffffLL) return DDPX_ERR_SHAPE;
  // persistent: one CTA per SM (each needs all 512 tensor-memory columns), walking its share of the items
  const unsigned grid = static_cast<unsigned>(nitems < num_sms() ? nitems : num_sms());
  attn_bwd_dq_kernel<<<grid, kBwdThreads, smem_dq, stream>>>(tq, td, to, p);
  if (cudaGetLastError() != cudaSuccess) return DDPX_ERR_LAUNCH;
  attn_bwd_dkv_kernel<<<grid, kBwdThreads, smem_dkv, stream>>>(tq, td, p);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" unsigned int ddpx_attn_bwd_read_error_word(void) {
  unsigned int v = 0;
  cudaMemcpyFromSymbol(&v, g_ddpx_error_word, sizeof(v));
  return v;
}
