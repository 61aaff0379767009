// Tensor-pipe probe: how many cycles does one tcgen05.mma (M=128, N, K=32 bytes) take when issued back to back from
// shared-memory operands, (a) into one accumulator, (b) alternating between two accumulators, and (c) while TMA keeps
// writing operand-sized tiles into OTHER shared-memory buffers at the pace of a real pipeline (no data dependence)?
// Separates "tensor pipe / dependent-accumulate" limits from "operand delivery" limits for the kernels in
// csrc/kernels/gemm_sm100.cu.   Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I csrc/kernels
//                                      bench/mma_probe.cu -o build/mma_probe -lcuda
#include <cuda.h>
#include <cuda_runtime.h>
#include <cudaTypedefs.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>

#include "sm100_ptx.cuh"

using namespace ddpx;

struct ProbeParams {
  int iters;        // k-steps; each k-step = 4 MMAs + 1 commit
  int N;            // MMA N
  int nacc;         // 1 or 2 accumulators
  int bf16;         // 1 = kind::f16 (bf16), 0 = kind::tf32
  int a_mn, b_mn;   // operand majorness (descriptor fields only - the data is noise)
  int tma;          // 1 = a producer warp streams tiles with TMA concurrently
  int stages;       // TMA ring depth (buffers alias the MMA operand region: the data is noise anyway)
  int pattern;      // 0: CTAs read overlapping tiles; 1: A shared by all CTAs, B distinct per CTA, every B tile read 9
                    // times in a row (conv taps); 2: A and B distinct, 9x reuse; 3: B distinct, never re-read
  int spin;         // 1: warps 4-11 (all lanes) poll a barrier that completes at the very end, like epilogue warps
                    // waiting for their accumulator; 2: only lane 0 of each polls
  int real_ops;     // 1: the MMAs read the tiles TMA just delivered (random data => realistic switching power)
  int b_rows;       // rows of the streamed B tile (256 = single-CTA traffic, 128 = what a CTA of a pair loads)
  long long* cycles;
};

constexpr int kStages = 8;       // barrier slots; p.stages of them are used
constexpr int kOperandBytes = 16384 + 32768;

template <bool kBF16>
__global__ void __launch_bounds__(128, 1) probe_kernel(const __grid_constant__ CUtensorMap tmA,
                                                        const __grid_constant__ CUtensorMap tmB, const ProbeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + 4 * kOperandBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* done_bar = empty_bar + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  // operand region: small finite values
  for (int i = threadIdx.x; i < kOperandBytes / 4; i += blockDim.x) {
    if (kBF16) reinterpret_cast<uint32_t*>(smem)[i] = 0x3C003C00u + ((i * 2654435761u) & 0x007F007Fu);
    else reinterpret_cast<float*>(smem)[i] = 0.01f * static_cast<float>((i * 2654435761u) >> 24);
  }
  fence_proxy_async_smem();
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0 && p.tma) {
    int stage = 0; uint32_t phase = 0;
    const uint32_t bytes = 16384u + static_cast<uint32_t>(p.b_rows) * 128u;
    for (int it = 0; it < p.iters; ++it) {
      mbar_wait(&empty_bar[stage], phase ^ 1u, 0x100u);
      uint8_t* sa = smem + stage * bytes;
      mbar_arrive_expect_tx(&full_bar[stage], bytes);
      int ka = (it * 32) & 4095, ra = ((blockIdx.x * 7 + it) & 31) * 256, kb = ka, rb = ra;
      if (p.pattern >= 1) {
        // the buffer is 32 row blocks (256 rows) x 128 column blocks (32 floats): 4096 tiles; CTA b's n-th fresh tile
        const int reuse = (p.pattern == 3) ? 1 : 9;
        const int t = (static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) * (it / reuse)) & 4095;
        kb = (t & 127) * 32; rb = (t >> 7) * 256;
        if (p.pattern == 2) { const int u = (t + 2048) & 4095; ka = (u & 127) * 32; ra = (u >> 7) * 256; }
        else { ka = (it * 32) & 4095; ra = 0; }
      }
      tma_load_2d(sa, &tmA, &full_bar[stage], ka, ra);
      tma_load_2d(sa + 16384, &tmB, &full_bar[stage], kb, rb);
      if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1 && lane == 0) {
    constexpr int kElt = kBF16 ? 2 : 4;
    constexpr uint32_t kBK = 128 / kElt, kUmmaK = 32 / kElt;
    const bool a_mn = p.a_mn, b_mn = p.b_mn;
    const uint32_t idesc = make_idesc(kBF16, a_mn, b_mn, 128, static_cast<uint32_t>(p.N));
    const uint32_t a_lbo = a_mn ? kBK * 128u : 16u, b_lbo = b_mn ? kBK * 128u : 16u;
    const uint32_t a_kadv = a_mn ? kUmmaK * 128u : 32u, b_kadv = b_mn ? kUmmaK * 128u : 32u;
    const uint32_t a_lay = (a_mn && !kBF16) ? 1u : 2u, b_lay = (b_mn && !kBF16) ? 1u : 2u;
    const uint32_t a_sbo = (a_mn && !kBF16) ? 512u : 1024u, b_sbo = (b_mn && !kBF16) ? 512u : 1024u;
    const uint32_t sa = smem_u32(smem), sb = sa + 16384u;
    int stage = 0;
    uint32_t phase = 0;
    const long long t0 = clock64();
    for (int it = 0; it < p.iters; ++it) {
      if (p.tma) {                       // paced like the real pipeline: the tile must have landed (operands are
        mbar_wait(&full_bar[stage], phase, 0x300u);   // still read from the fixed region, so there is no data hazard)
        tc_fence_after();
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint64_t adesc = make_smem_desc_sw128(sa + k * a_kadv, a_lbo, a_sbo, a_lay);
        const uint64_t bdesc = make_smem_desc_sw128(sb + k * b_kadv, b_lbo, b_sbo, b_lay);
        const uint32_t d = tmem_base + ((p.nacc == 2 && (k & 1)) ? 256u : 0u);
        tc_mma<kBF16>(d, adesc, bdesc, idesc, (it > 0 || k > 1) ? 1u : 0u);
      }
      tc_commit(&empty_bar[stage]);
      if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }
    tc_commit(done_bar);
    mbar_wait(done_bar, 0, 0x600u);
    const long long t1 = clock64();
    p.cycles[blockIdx.x] = t1 - t0;
  }
  __syncthreads();
  tc_fence_before();
  __syncthreads();
  if (warp == 2) tmem_dealloc(tmem_base, 512);
}

// CTA-pair flavour (cta_group::2): M = 256 across the pair, each CTA streams A (128 rows) + half of B (128 rows) =
// 32 KB per k-step; completions land on the leader's barrier, the leader's commits reach both CTAs.
template <bool kBF16>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(384, 1)
probe_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const ProbeParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + 4 * kOperandBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* done_bar = empty_bar + kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done_bar + 1);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  for (int i = threadIdx.x; i < kOperandBytes / 4; i += blockDim.x) {
    if (kBF16) reinterpret_cast<uint32_t*>(smem)[i] = 0x3C003C00u + ((i * 2654435761u) & 0x007F007Fu);
    else reinterpret_cast<float*>(smem)[i] = 0.01f * static_cast<float>((i * 2654435761u) >> 24);
  }
  fence_proxy_async_smem();
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(done_bar, 1);
    fence_barrier_init();
  }
  if (warp == 2) tmem_alloc_2sm(tmem_slot, 512);
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t bytes = 32768u;
  if (warp == 0 && lane == 0 && p.tma) {
    int stage = 0; uint32_t phase = 0;
    for (int it = 0; it < p.iters; ++it) {
      mbar_wait(&empty_bar[stage], phase ^ 1u, 0x100u);
      uint8_t* sa = smem + stage * bytes;
      if (leader) mbar_arrive_expect_tx(&full_bar[stage], bytes * 2);
      const int pairid = blockIdx.x >> 1;
      const int t = (pairid * 2 + static_cast<int>(rank) + static_cast<int>(gridDim.x) * (it / 9)) & 4095;
      const int kb = (t & 127) * 32, rb = (t >> 7) * 256;
      tma_load_2d_2sm(sa, &tmA, &full_bar[stage], (it * 32) & 4095, static_cast<int>(rank) * 128);
      tma_load_2d_2sm(sa + 16384, &tmB, &full_bar[stage], kb, rb);
      if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1 && lane == 0 && leader) {
    constexpr int kElt = kBF16 ? 2 : 4;
    const uint32_t idesc = make_idesc(kBF16, false, false, 256, static_cast<uint32_t>(p.N));
    const uint32_t sa = smem_u32(smem), sb = sa + 16384u;
    int stage = 0;
    uint32_t phase = 0;
    const long long t0 = clock64();
    for (int it = 0; it < p.iters; ++it) {
      if (p.tma) { mbar_wait(&full_bar[stage], phase, 0x300u); tc_fence_after(); }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t so = p.real_ops ? static_cast<uint32_t>(stage) * bytes : 0u;
        const uint64_t adesc = make_smem_desc_sw128(sa + so + k * 32u, 16u, 1024u, 2u);
        const uint64_t bdesc = make_smem_desc_sw128(sb + so + k * 32u, 16u, 1024u, 2u);
        tc_mma_2sm<kBF16>(tmem_base, adesc, bdesc, idesc, (it > 0 || k > 0) ? 1u : 0u);
      }
      tc_commit_2sm(&empty_bar[stage]);
      if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }
    tc_commit_2sm(done_bar);
    mbar_wait(done_bar, 0, 0x600u);
    const long long t1 = clock64();
    p.cycles[blockIdx.x >> 1] = t1 - t0;
    (void)kElt;
  } else if (warp == 1 && lane == 0) {
    mbar_wait(done_bar, 0, 0x601u);       // the peer's smem and barriers are in use until the leader's last commit
  } else if (warp >= 4 && p.spin) {
    if (p.spin == 1 || lane == 0) mbar_wait(done_bar, 0, 0x602u);
    __syncwarp();
  }
  __syncthreads();
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  if (warp == 2) tmem_dealloc_2sm(tmem_base, 512);
}

// normal-ish random fp32 values; read as bf16 pairs the upper halves are still ordinary numbers
__global__ void fill_random(float* p, size_t n) {
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    uint32_t h = static_cast<uint32_t>(i) * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
    p[i] = (static_cast<float>(h & 0xFFFF) - 32768.f) * (1.f / 16384.f);
  }
}

static PFN_cuTensorMapEncodeTiled_v12000 encode_fn() {
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q);
  return reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(ptr);
}

int main() {
  const size_t rows = 8192, cols = 4096;       // 128 MiB of fp32: about the size of L2
  float* buf = nullptr;
  cudaMalloc(&buf, rows * cols * 4);
  fill_random<<<1024, 256>>>(buf, rows * cols);
  cudaDeviceSynchronize();
  long long* d_cycles = nullptr;
  cudaMalloc(&d_cycles, 148 * sizeof(long long));
  auto enc = encode_fn();
  auto mk = [&](CUtensorMap* m, uint32_t box_rows) {
    cuuint64_t gd[2] = {cols, rows};
    cuuint64_t gs[1] = {cols * 4};
    cuuint32_t bx[2] = {32, box_rows}, es[2] = {1, 1};
    CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, buf, gd, gs, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("tensor map encode failed %d\n", (int)r); exit(1); }
  };
  const size_t smem = 4 * kOperandBytes + 1024 + 256;
  cudaFuncSetAttribute(probe_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(probe_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(probe_pair_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaFuncSetAttribute(probe_pair_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  struct Case { const char* name; int bf16, N, nacc, a_mn, b_mn, tma, b_rows, stages, pattern, pair, spin, real_ops; };
  std::vector<Case> cases = {
      {"tf32 PAIR +tma32K x6 fixed operands", 0, 256, 1, 0, 0, 1, 128, 6, 1, 1, 0, 0},
      {"tf32 PAIR +tma32K x6 real operands (random data)", 0, 256, 1, 0, 0, 1, 128, 6, 1, 1, 0, 1},
      {"bf16 PAIR +tma32K x6 fixed operands", 1, 256, 1, 0, 0, 1, 128, 6, 1, 1, 0, 0},
      {"bf16 PAIR +tma32K x6 real operands (random data)", 1, 256, 1, 0, 0, 1, 128, 6, 1, 1, 0, 1},
  };
  for (const auto& c : cases) {
    CUtensorMap ta, tb;
    mk(&ta, 128);
    mk(&tb, (uint32_t)c.b_rows);
    ProbeParams p{};
    p.iters = 16384; p.N = c.N; p.nacc = c.nacc; p.bf16 = c.bf16; p.a_mn = c.a_mn; p.b_mn = c.b_mn; p.tma = c.tma;
    p.b_rows = c.b_rows; p.stages = c.stages; p.pattern = c.pattern; p.spin = c.spin; p.real_ops = c.real_ops; p.cycles = d_cycles;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEventRecord(e0);
      if (c.pair) {
        if (c.bf16) probe_pair_kernel<true><<<sms, 384, smem>>>(ta, tb, p);
        else probe_pair_kernel<false><<<sms, 384, smem>>>(ta, tb, p);
      } else if (c.bf16) probe_kernel<true><<<sms, 128, smem>>>(ta, tb, p);
      else probe_kernel<false><<<sms, 128, smem>>>(ta, tb, p);
      cudaEventRecord(e1);
      cudaError_t e = cudaEventSynchronize(e1);
      if (e != cudaSuccess) { printf("%s: CUDA error %s\n", c.name, cudaGetErrorString(e)); return 1; }
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      best = std::min(best, ms);
    }
    const int nent = c.pair ? sms / 2 : sms;
    std::vector<long long> h(nent);
    cudaMemcpy(h.data(), d_cycles, nent * sizeof(long long), cudaMemcpyDeviceToHost);
    std::sort(h.begin(), h.end());
    const double n_mma = 4.0 * p.iters;
    const double flop = 2.0 * 128 * c.N * (c.bf16 ? 16 : 8) * n_mma * sms;
    printf("%-52s cyc/MMA median %.1f (min %.1f max %.1f) | kernel %.3f ms -> %.0f TFLOP/s | eff. clock %.0f MHz | TMA %.2f TB/s\n", c.name,
           h[nent / 2] / n_mma, h[0] / n_mma, h[nent - 1] / n_mma, best, flop / best * 1e-9, h[nent / 2] / best * 1e-3,
           c.tma ? (c.pair ? 32768.0 : 16384.0 + c.b_rows * 128.0) * p.iters * sms / best * 1e-9 : 0.0);
    fflush(stdout);
  }
  return 0;
}
