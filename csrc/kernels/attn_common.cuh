// Shared pieces of the fused-attention kernels (attn_sm100.cu forward, attn_bwd_sm100.cu backward).
#pragma once
#include "sm100_ptx.cuh"
#include "tmap.hpp"
#include "../../include/ddpx_kernels.h"

namespace ddpx {

constexpr int kHD = 64;            // head dimension (all GPT-2 sizes)
constexpr int kTile = 128;         // query rows per tile = key rows per block = TMEM lanes
constexpr int kTileBytes = kTile * kHD * 2;    // 16 KB: 128 rows x 128 B, one 128-byte-swizzled operand tile

struct AttnParams {
  int B, T, H, C;                  // C = H * 64; qkv is [B, T, 3C] (q | k | v), out / dout are [B, T, C]
  int nq;                          // ceil(T / 128)
  int Tp;                          // nq * 128: row pitch of lse2 / delta (whole tiles can be fetched with bulk copies)
  float scale_log2;                // (1/sqrt(hd)) * log2(e): probabilities are exp2(s * scale_log2 - lse2)
  float scale;
  __nv_bfloat16* out;              // forward: O
  float* lse2;                     // [B, H, Tp] row log-sum-exp in base 2 (of the scaled scores)
  const __nv_bfloat16* qkv;
  const __nv_bfloat16* dout;       // backward
  float* delta;                    // [B, H, Tp] rowsum(dO * O) * scale: written by the dq kernel, read by the dkv kernel
  const __nv_bfloat16* o;          // backward: the forward output
  __nv_bfloat16* dqkv;             // [B, T, 3C]
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// K-major operand tile [128 rows][64 bf16] (what TMA writes with the 128-byte swizzle): k-step kk covers 16 elements
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t saddr, int kk) {
  return make_smem_desc_sw128(saddr + kk * 32, 16, 1024, 2);
}
// the same bytes read as an MN-major operand: [K = 128 rows][N = 64 contiguous]; k-step kk covers 16 rows
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t saddr, int kk) {
  return make_smem_desc_sw128(saddr + kk * 2048, 8192, 1024, 2);
}

// Work item of round k for this CTA: items are sorted heaviest first; even rounds deal them left to right, odd rounds right
// to left, so that every CTA's share has nearly the same weight.  Returns -1 when the CTA has no item in that round.
__device__ __forceinline__ int snake_item(int k, int nitems) {
  const int G = static_cast<int>(gridDim.x), c = static_cast<int>(blockIdx.x);
  const long long idx = static_cast<long long>(k) * G + ((k & 1) ? (G - 1 - c) : c);
  return idx < nitems ? static_cast<int>(idx) : -1;
}

// [B][T][row_elems] bf16 viewed as (row_elems, T, B); a box is 128 rows x 64 columns (one head's slice of 128 tokens).
// Rows past T are out of bounds in dimension 1 and arrive as zeros.
inline int attn_map(CUtensorMap* tm, const void* base, int B, int T, int row_elems) {
  uint64_t dims[3] = {static_cast<uint64_t>(row_elems), static_cast<uint64_t>(T), static_cast<uint64_t>(B)};
  uint64_t str[2] = {static_cast<uint64_t>(row_elems) * 2, static_cast<uint64_t>(row_elems) * 2 * T};
  uint32_t box[3] = {kHD, kTile, 1};
  return make_tmap(tm, base, true, 3, dims, str, box);
}

}  // namespace ddpx
