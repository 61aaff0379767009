// Backward of the fused causal attention (see attn_sm100.cu for the forward and the overall scheme).
//
// Two kernels, both recompute P = exp2(S * scale_log2 - lse2) from the saved log-sum-exp, neither uses atomics:
//   attn_bwd_dq_kernel   one CTA per (batch, head, 128-query tile).  Q_i and dO_i are written ONCE into tensor memory
//                        (bf16, K-major) by the threads that own the rows, so S = Q K^T and dP = dO V^T are TS-mode MMAs
//                        against the K / V tiles TMA streams in; dS goes back into tensor memory and dQ += dS K is a third
//                        TS-mode MMA.  The same threads compute delta = rowsum(dO * O) on the way and publish it.
//   attn_bwd_dkv_kernel  one CTA per (batch, head, 128-key tile).  K_j and V_j live in tensor memory; S^T = K Q^T and
//                        dP^T = V dO^T put keys on the TMEM lanes, so P^T and dS^T are already the A operands of
//                        dV += P^T dO and dK += dS^T Q (Q / dO tiles are read K-major for the first pair and MN-major for
//                        the second - the same shared-memory bytes).
// 320 threads: warp 0 = TMA producer, warp 1 = MMA issuer + TMEM owner, warps 2-9 = compute (two warps per TMEM lane
// quarter, each working on half of the 128 columns).
#include "attn_common.cuh"

namespace ddpx {

constexpr int kBwdThreads = 320;

// the thread's row (64 bf16 = 128 B) from global memory; rows past the end read as zeros
__device__ __forceinline__ void load_row(const __nv_bfloat16* base, bool valid, uint32_t (&v)[32]) {
  const uint4* src = reinterpret_cast<const uint4*>(base);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    uint4 u = valid ? __ldg(src + i) : make_uint4(0u, 0u, 0u, 0u);
    v[4 * i] = u.x; v[4 * i + 1] = u.y; v[4 * i + 2] = u.z; v[4 * i + 3] = u.w;
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
// tensor memory (512 columns): [0,32) Q  [32,64) dO  [64,192) S  [192,320) dP (dS overlays it)  [320,384) dQ
namespace dq {
constexpr uint32_t kColQ = 0, kColDO = 32, kColS = 64, kColDP = 192, kColDQ = 320;
}

__global__ void __launch_bounds__(kBwdThreads, 1)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams p) {
  using namespace dq;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;                        // 2 stages
  uint8_t* sV = smem + 2 * kTileBytes;       // 2 stages
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 4 * kTileBytes);
  uint64_t* kv_full = bars;        // [2]
  uint64_t* kv_empty = bars + 2;   // [2]
  uint64_t* rows_ready = bars + 4; // Q / dO are in tensor memory (8 compute warps)
  uint64_t* sdp_full = bars + 5;   // S and dP complete
  uint64_t* ds_full = bars + 6;    // dS written (8 compute warps)
  uint64_t* acc_done = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int bh_count = p.B * p.H;
  const int qt = p.nq - 1 - static_cast<int>(blockIdx.x) / bh_count;     // heaviest tiles first
  const int bh = static_cast<int>(blockIdx.x) % bh_count;
  const int b = bh / p.H, h = bh - b * p.H;
  const int nkv = qt + 1;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmQKV);
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < 2; ++s) { mbar_init(&kv_full[s], 1); mbar_init(&kv_empty[s], 1); }
      mbar_init(rows_ready, 8);
      mbar_init(sdp_full, 1);
      mbar_init(ds_full, 8);
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
    if (lane == 0) {
      for (int j = 0; j < nkv; ++j) {
        const int s = j & 1;
        mbar_wait(&kv_empty[s], ((j >> 1) & 1) ^ 1u, 0x600u + s);
        mbar_arrive_expect_tx(&kv_full[s], 2 * kTileBytes);
        tma_load_3d(sK + s * kTileBytes, &tmQKV, &kv_full[s], p.C + h * kHD, j * kTile, b);
        tma_load_3d(sV + s * kTileBytes, &tmQKV, &kv_full[s], 2 * p.C + h * kHD, j * kTile, b);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = make_idesc(true, false, false, kTile, kTile);    // [128 q][128 kv]
      const uint32_t idesc_dq = make_idesc(true, false, true, kTile, kHD);      // [128 q][64], K tile read MN-major
      mbar_wait(rows_ready, 0, 0x610u);
      tc_fence_after();
      for (int j = 0; j < nkv; ++j) {
        const int s = j & 1;
        mbar_wait(&kv_full[s], (j >> 1) & 1, 0x620u + s);
        tc_fence_after();
        const uint32_t aK = smem_u32(sK + s * kTileBytes), aV = smem_u32(sV + s * kTileBytes);
#pragma unroll
        for (int kk = 0; kk < kHD / 16; ++kk)
          tc_mma_ts_bf16(tmem + kColS, tmem + kColQ + kk * 8, desc_kmajor(aK, kk), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < kHD / 16; ++kk)
          tc_mma_ts_bf16(tmem + kColDP, tmem + kColDO + kk * 8, desc_kmajor(aV, kk), idesc_s, kk > 0 ? 1u : 0u);
        tc_commit(sdp_full);
        mbar_wait(ds_full, j & 1, 0x630u);
        tc_fence_after();
        // dS columns: half 0 of the key block sits at [0,32), half 1 at [64,96) of the dP region (see the compute warps)
#pragma unroll
        for (int kk = 0; kk < kTile / 16; ++kk)
          tc_mma_ts_bf16(tmem + kColDQ, tmem + kColDP + (kk >> 2) * 64 + (kk & 3) * 8, desc_mnmajor(aK, kk), idesc_dq,
                         (j > 0 || kk > 0) ? 1u : 0u);
        tc_commit(&kv_empty[s]);
      }
      tc_commit(acc_done);
    }
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;                   // which 64 key columns of a block this warp handles
    const int r = quarter * 32 + lane;
    const int q = qt * kTile + r;
    const bool valid = q < p.T;
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    const float sl2 = p.scale_log2;
    // ---- resident operands: half 0 warps bring Q, half 1 warps bring dO
    {
      uint32_t v[32];
      if (half == 0) {
        load_row(p.qkv + (static_cast<long long>(b) * p.T + (valid ? q : 0)) * 3 * p.C + h * kHD, valid, v);
        tmem_st_32x32(tmem + lane_base + kColQ, v);
      } else {
        const long long off = (static_cast<long long>(b) * p.T + (valid ? q : 0)) * p.C + h * kHD;
        load_row(p.dout + off, valid, v);
        tmem_st_32x32(tmem + lane_base + kColDO, v);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(rows_ready);
    }
    // per-row scalars: delta = rowsum(dO * O), needed by both warps of the row; one of them publishes it for the dkv kernel
    float delta;
    {
      const long long off = (static_cast<long long>(b) * p.T + (valid ? q : 0)) * p.C + h * kHD;
      uint32_t v[32], o[32];
      load_row(p.dout + off, valid, v);
      load_row(p.o + off, valid, o);
      float d = 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) d += bf_lo(v[i]) * bf_lo(o[i]) + bf_hi(v[i]) * bf_hi(o[i]);
      delta = d;
      if (half == 1) p.delta[(static_cast<long long>(b) * p.H + h) * p.Tp + q] = d;      // q < Tp always
    }
    const float lse2 = valid ? p.lse2[(static_cast<long long>(b) * p.H + h) * p.Tp + q] : 0.f;
    const uint32_t tS = tmem + lane_base + kColS + half * 64;
    const uint32_t tDP = tmem + lane_base + kColDP + half * 64;
    for (int j = 0; j < nkv; ++j) {
      mbar_wait(sdp_full, j & 1, 0x640u);
      tc_fence_after();
      const bool diag = (j == qt);
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t sv[32], dv[32];
        tmem_ld_32x32(tS + c * 32, sv);
        tmem_ld_32x32(tDP + c * 32, dv);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = ex2(fmaf(__uint_as_float(sv[i]), sl2, -lse2));
          float p1 = ex2(fmaf(__uint_as_float(sv[i + 1]), sl2, -lse2));
          float d0 = p0 * (__uint_as_float(dv[i]) - delta) * p.scale;
          float d1 = p1 * (__uint_as_float(dv[i + 1]) - delta) * p.scale;
          if (diag) {
            const int col = half * 64 + c * 32 + i;
            if (col > r) d0 = 0.f;
            if (col + 1 > r) d1 = 0.f;
          }
          pk[i >> 1] = pack_bf16(d0, d1);
        }
        tmem_st_32x16(tDP + c * 16, pk);      // overlays dP columns this warp has already consumed
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(ds_full);
    }
    // ---- epilogue: each warp stores 32 of the row's 64 dQ columns
    mbar_wait(acc_done, 0, 0x650u);
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

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

// =====================================================================================================================
// dK, dV
// =====================================================================================================================
// tensor memory (512 columns): [0,32) K  [32,64) V  [64,192) S^T (P^T overlays it)  [192,320) dP^T (dS^T overlays it)
//                              [320,384) dV  [384,448) dK
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
  float* sLse = reinterpret_cast<float*>(smem + 4 * kTileBytes);      // [2][128]
  float* sDelta = sLse + 2 * kTile;                                   // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sDelta + 2 * kTile);
  uint64_t* q_full = bars;         // [2]
  uint64_t* q_empty = bars + 2;    // [2]
  uint64_t* rows_ready = bars + 4;
  uint64_t* sdp_full = bars + 5;
  uint64_t* pds_full = bars + 6;
  uint64_t* acc_done = bars + 7;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int bh_count = p.B * p.H;
  const int kt = static_cast<int>(blockIdx.x) / bh_count;                // key tile 0 meets every query tile: heaviest first
  const int bh = static_cast<int>(blockIdx.x) % bh_count;
  const int b = bh / p.H, h = bh - b * p.H;
  const int nqt = p.nq - kt;                                             // query tiles kt .. nq-1

  if (warp == 0 && lane == 0) { tma_prefetch_desc(&tmQKV); tma_prefetch_desc(&tmDO); }
  if (warp == 1) {
    if (lane == 0) {
      for (int s = 0; s < 2; ++s) { mbar_init(&q_full[s], 1); mbar_init(&q_empty[s], 1); }
      mbar_init(rows_ready, 8);
      mbar_init(sdp_full, 1);
      mbar_init(pds_full, 8);
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
  const long long stat_row = (static_cast<long long>(b) * p.H + h) * p.Tp;

  if (warp == 0) {
    if (lane == 0) {
      for (int it = 0; it < nqt; ++it) {
        const int s = it & 1, i = kt + it;
        mbar_wait(&q_empty[s], ((it >> 1) & 1) ^ 1u, 0x700u + s);
        mbar_arrive_expect_tx(&q_full[s], 2 * kTileBytes + 2 * kTile * 4);
        tma_load_3d(sQ + s * kTileBytes, &tmQKV, &q_full[s], h * kHD, i * kTile, b);
        tma_load_3d(sDO + s * kTileBytes, &tmDO, &q_full[s], h * kHD, i * kTile, b);
        bulk_load(sLse + s * kTile, p.lse2 + stat_row + i * kTile, kTile * 4, &q_full[s]);
        bulk_load(sDelta + s * kTile, p.delta + stat_row + i * kTile, kTile * 4, &q_full[s]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc_s = make_idesc(true, false, false, kTile, kTile);    // [128 kv][128 q]
      const uint32_t idesc_acc = make_idesc(true, false, true, kTile, kHD);     // [128 kv][64], Q / dO tiles read MN-major
      mbar_wait(rows_ready, 0, 0x710u);
      tc_fence_after();
      for (int it = 0; it < nqt; ++it) {
        const int s = it & 1;
        mbar_wait(&q_full[s], (it >> 1) & 1, 0x720u + s);
        tc_fence_after();
        const uint32_t aQ = smem_u32(sQ + s * kTileBytes), aDO = smem_u32(sDO + s * kTileBytes);
#pragma unroll
        for (int kk = 0; kk < kHD / 16; ++kk)
          tc_mma_ts_bf16(tmem + kColS, tmem + kColK + kk * 8, desc_kmajor(aQ, kk), idesc_s, kk > 0 ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < kHD / 16; ++kk)
          tc_mma_ts_bf16(tmem + kColDP, tmem + kColV + kk * 8, desc_kmajor(aDO, kk), idesc_s, kk > 0 ? 1u : 0u);
        tc_commit(sdp_full);
        mbar_wait(pds_full, it & 1, 0x730u);
        tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < kTile / 16; ++kk)
          tc_mma_ts_bf16(tmem + kColDV, tmem + kColS + (kk >> 2) * 64 + (kk & 3) * 8, desc_mnmajor(aDO, kk), idesc_acc,
                         (it > 0 || kk > 0) ? 1u : 0u);
#pragma unroll
        for (int kk = 0; kk < kTile / 16; ++kk)
          tc_mma_ts_bf16(tmem + kColDK, tmem + kColDP + (kk >> 2) * 64 + (kk & 3) * 8, desc_mnmajor(aQ, kk), idesc_acc,
                         (it > 0 || kk > 0) ? 1u : 0u);
        tc_commit(&q_empty[s]);
      }
      tc_commit(acc_done);
    }
  } else {
    const int quarter = warp & 3;
    const int half = (warp - 2) >> 2;                   // which 64 query columns of a tile this warp handles
    const int r = quarter * 32 + lane;
    const int kv = kt * kTile + r;
    const bool valid = kv < p.T;
    const uint32_t lane_base = static_cast<uint32_t>(quarter * 32) << 16;
    const float sl2 = p.scale_log2;
    {
      // resident operands: half 0 warps bring the K rows, half 1 warps the V rows
      uint32_t v[32];
      load_row(p.qkv + (static_cast<long long>(b) * p.T + (valid ? kv : 0)) * 3 * p.C + (half + 1) * p.C + h * kHD, valid, v);
      tmem_st_32x32(tmem + lane_base + (half == 0 ? kColK : kColV), v);
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(rows_ready);
    }
    const uint32_t tS = tmem + lane_base + kColS + half * 64;
    const uint32_t tDP = tmem + lane_base + kColDP + half * 64;
    for (int it = 0; it < nqt; ++it) {
      const int s = it & 1, i = kt + it;
      mbar_wait(&q_full[s], (it >> 1) & 1, 0x740u + s);       // lse2 / delta of this query tile are in shared memory
      mbar_wait(sdp_full, it & 1, 0x750u);
      tc_fence_after();
      // keys above the query (diagonal tile) and queries past the end of the sequence contribute nothing
      const bool masked = (it == 0) || ((i + 1) * kTile > p.T);
      const float4* lse4 = reinterpret_cast<const float4*>(sLse + s * kTile + half * 64);
      const float4* del4 = reinterpret_cast<const float4*>(sDelta + s * kTile + half * 64);
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {
        uint32_t sv[32], dv[32];
        tmem_ld_32x32(tS + c * 32, sv);
        tmem_ld_32x32(tDP + c * 32, dv);
        tmem_ld_wait();
        uint32_t pk[16], dk[16];
#pragma unroll
        for (int g = 0; g < 8; ++g) {
          const float4 L = lse4[c * 8 + g], D = del4[c * 8 + g];
          const float ls[4] = {L.x, L.y, L.z, L.w}, ds[4] = {D.x, D.y, D.z, D.w};
          float pv[4], dsv[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int x = g * 4 + e;
            pv[e] = ex2(fmaf(__uint_as_float(sv[x]), sl2, -ls[e]));
            dsv[e] = pv[e] * (__uint_as_float(dv[x]) - ds[e]) * p.scale;
            if (masked) {
              const int col = half * 64 + c * 32 + x;                 // query row inside the tile
              if ((it == 0 && col < r) || i * kTile + col >= p.T) { pv[e] = 0.f; dsv[e] = 0.f; }
            }
          }
          pk[g * 2] = pack_bf16(pv[0], pv[1]); pk[g * 2 + 1] = pack_bf16(pv[2], pv[3]);
          dk[g * 2] = pack_bf16(dsv[0], dsv[1]); dk[g * 2 + 1] = pack_bf16(dsv[2], dsv[3]);
        }
        tmem_st_32x16(tS + c * 16, pk);       // P^T over the S^T columns this warp has consumed
        tmem_st_32x16(tDP + c * 16, dk);      // dS^T over the dP^T columns
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(pds_full);
    }
    // ---- epilogue: half 0 warps store the dV row, half 1 warps the dK row
    mbar_wait(acc_done, 0, 0x760u);
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
  CUtensorMap tq, td;
  int rc;
  if ((rc = attn_map(&tq, qkv, B, T, 3 * p.C)) != DDPX_OK) return rc;
  if ((rc = attn_map(&td, dout, B, T, p.C)) != DDPX_OK) return rc;
  const size_t smem_dq = 4 * kTileBytes + 1024 + 128;
  const size_t smem_dkv = 4 * kTileBytes + 4 * kTile * 4 + 1024 + 128;
  static bool attr_set = false;
  if (!attr_set) {
    if (cudaFuncSetAttribute(attn_bwd_dq_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_dq)) != cudaSuccess ||
        cudaFuncSetAttribute(attn_bwd_dkv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem_dkv)) != cudaSuccess)
      return DDPX_ERR_LAUNCH;
    attr_set = true;
  }
  const long long grid = static_cast<long long>(p.nq) * B * H;
  if (grid > 0x7fffffffLL) return DDPX_ERR_SHAPE;
  attn_bwd_dq_kernel<<<static_cast<unsigned>(grid), kBwdThreads, smem_dq, stream>>>(tq, p);
  if (cudaGetLastError() != cudaSuccess) return DDPX_ERR_LAUNCH;
  attn_bwd_dkv_kernel<<<static_cast<unsigned>(grid), kBwdThreads, smem_dkv, stream>>>(tq, td, p);
  return cudaGetLastError() == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH;
}

extern "C" unsigned int ddpx_attn_bwd_read_error_word(void) {
  unsigned int v = 0;
  cudaMemcpyFromSymbol(&v, g_ddpx_error_word, sizeof(v));
  return v;
}
