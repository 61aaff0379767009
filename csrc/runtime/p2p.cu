// Peer-to-peer communication backend over NVLink 5 / NVSwitch, sm_100a.
//
// Replaces ProcessGroupNCCL (ncclAllReduce / ncclBroadcast) that torch DDP drives for the reference at
// /root/reference/multigpu.py:32,89 (W6-W9):
//   * symmetric buffers: every rank cudaMalloc's the same sizes in the same order and maps every peer's allocation
//     into its own address space through CUDA IPC handles (exchanged over the control-plane store by the Python side);
//   * a signal pad per rank (monotonic generation counters written with st.release.sys, polled with ld.acquire.sys);
//   * ONE kernel per gradient bucket that does all-reduce (two-shot: reduce-scatter by peer loads + all-gather by peer
//     stores; one-shot for small buckets) AND the optimizer update (SGD-momentum or AdamW) of the parameters the bucket
//     covers - no NCCL call and no separate elementwise kernel on that path;
//   * a pull-based broadcast (rank 0 packs tensors into its staging buffer, peers copy them out over NVLink) used for
//     the initial parameter sync and the per-forward BatchNorm-buffer sync.
// Every spin is bounded: after the configured timeout (ddpx_p2p_set_timeout) a rank writes an error code into its pad and
// traps instead of hanging the box.
#include <cuda_runtime.h>
#include <cstdlib>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string.h>
#include "../../include/ddpx_kernels.h"

namespace {

constexpr int kMaxRanks = 8;
constexpr int kMaxSlots = 128;
// signal-pad layout (uint32 words)
constexpr int PAD_ARRIVE = 0;     // [8]  peer r's bucket is filled for generation g
constexpr int PAD_DONE = 8;       // [8]  peer r finished reducing + distributing its shard for generation g
constexpr int PAD_EXIT = 16;      // [8]  peer r finished reading everybody's bucket (one-shot)
constexpr int PAD_BCAST = 24;     // [1]  source finished packing broadcast generation g
constexpr int PAD_ACK = 32;       // [8]  peer r consumed broadcast generation g   (lives on the source's pad)
constexpr int PAD_ERROR = 48;     // [1]  non-zero after a timed-out wait
constexpr int PAD_COUNTER = 56;   // [4]  local CTA-arrival counters (never touched by peers)
// How long a device-side flag wait may last before the rank gives up (writes PAD_ERROR, traps).  Set per device by
// ddpx_p2p_set_timeout (config.wait_timeout_s, default 600 s like torch's process-group timeout): inter-rank skew is
// legitimate - rank 0 writing a checkpoint while the peers already wait in the next step's collective (the reference saves
// without a barrier, multigpu.py:118), a stalled data loader, a debugger.
__device__ unsigned long long g_timeout_ns = 600ull * 1000ull * 1000ull * 1000ull;

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// monotonic publish: the flag only ever grows, whichever of two writers (two successive broadcast sources) lands last
__device__ __forceinline__ void red_max_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("red.release.sys.global.max.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// wait until *flag >= gen (wrap-safe signed compare); bounded
__device__ __forceinline__ void wait_flag(uint32_t* flag, uint32_t gen, uint32_t* pad, uint32_t code) {
  if (static_cast<int32_t>(ld_acquire_sys(flag) - gen) >= 0) return;
  const unsigned long long t0 = globaltimer();
  while (static_cast<int32_t>(ld_acquire_sys(flag) - gen) < 0) {
    __nanosleep(64);
    if (globaltimer() - t0 > g_timeout_ns) {
      pad[PAD_ERROR] = code;
      __threadfence_system();
      asm volatile("trap;");
    }
  }
}

struct Peers {
  void* buf[kMaxRanks];        // every rank's copy of the symmetric buffer, indexed by rank (buf[rank] is local)
  uint32_t* pad[kMaxRanks];    // every rank's signal pad
  int rank, world;
  void* mc;                    // NVLS: multicast mapping of the same symmetric buffer (nullptr = none)
};

struct Slot {
  void* param;
  void* state0;     // SGD momentum buffer / AdamW exp_avg
  void* state1;     // AdamW exp_avg_sq
  long long offset; // element offset inside the bucket
  long long numel;
};
struct SlotTable {
  Slot s[kMaxSlots];
  int n;
};
struct Hyper {
  int opt;          // 0 none (just all-reduce), 1 SGD, 2 AdamW
  float lr, momentum, dampening, wd;
  int nesterov, first_step;
  float beta1, beta2, eps, bc1, bc2_sqrt;
};

template <typename T> struct Vec;
template <> struct Vec<float> { static constexpr int N = 4; };
template <> struct Vec<__nv_bfloat16> { static constexpr int N = 8; };

template <typename T> __device__ __forceinline__ void unpack(const uint4& u, float* f);
template <> __device__ __forceinline__ void unpack<float>(const uint4& u, float* f) {
  f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
}
template <> __device__ __forceinline__ void unpack<__nv_bfloat16>(const uint4& u, float* f) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <typename T> __device__ __forceinline__ uint4 pack(const float* f);
template <> __device__ __forceinline__ uint4 pack<float>(const float* f) {
  return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
}
template <> __device__ __forceinline__ uint4 pack<__nv_bfloat16>(const float* f) {
  uint32_t w[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    w[i] = *reinterpret_cast<uint32_t*>(&h);
  }
  return make_uint4(w[0], w[1], w[2], w[3]);
}
template <typename T> __device__ __forceinline__ float ldT(const T* p);
template <> __device__ __forceinline__ float ldT<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldT<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T> __device__ __forceinline__ void stT(T* p, float v);
template <> __device__ __forceinline__ void stT<float>(float* p, float v) { *p = v; }
template <> __device__ __forceinline__ void stT<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16(v); }

// grid-wide "everybody finished the previous phase": the last CTA to arrive runs `then` (one thread)
template <typename F>
__device__ __forceinline__ void last_cta_does(uint32_t* counter, F then) {
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int old = atomicAdd(counter, 1u);
    if (old == gridDim.x - 1) {
      *counter = 0;
      __threadfence_system();
      then();
    }
  }
}

// optimizer update of one element (torch SGD / AdamW semantics, see loss_optim.cu)
template <typename T>
__device__ __forceinline__ void apply_update(const Hyper& h, const Slot& s, long long j, float g) {
  T* p = static_cast<T*>(s.param) + j;
  float pv = ldT(p);
  if (h.opt == 1) {
    float gv = g + h.wd * pv;
    if (h.momentum != 0.f) {
      T* b = static_cast<T*>(s.state0) + j;
      const float bv = h.first_step ? gv : h.momentum * ldT(b) + (1.f - h.dampening) * gv;
      stT(b, bv);
      gv = h.nesterov ? gv + h.momentum * bv : bv;
    }
    stT(p, pv - h.lr * gv);
  } else {
    T* m = static_cast<T*>(s.state0) + j;
    T* v = static_cast<T*>(s.state1) + j;
    pv *= (1.f - h.lr * h.wd);
    const float mv = h.beta1 * ldT(m) + (1.f - h.beta1) * g;
    const float vv = h.beta2 * ldT(v) + (1.f - h.beta2) * g * g;
    stT(m, mv);
    stT(v, vv);
    stT(p, pv - (h.lr / h.bc1) * mv / (sqrtf(vv) / h.bc2_sqrt + h.eps));
  }
}

// the same update on registers: g, p, s0 (momentum / exp_avg), s1 (exp_avg_sq) hold one 16-byte vector each
template <int V>
__device__ __forceinline__ void update_regs(const Hyper& h, const float* g, float* p, float* s0, float* s1) {
#pragma unroll
  for (int k = 0; k < V; ++k) {
    if (h.opt == 1) {
      float gv = g[k] + h.wd * p[k];
      if (h.momentum != 0.f) {
        const float bv = h.first_step ? gv : h.momentum * s0[k] + (1.f - h.dampening) * gv;
        s0[k] = bv;
        gv = h.nesterov ? gv + h.momentum * bv : bv;
      }
      p[k] -= h.lr * gv;
    } else {
      const float pv = p[k] * (1.f - h.lr * h.wd);
      const float mv = h.beta1 * s0[k] + (1.f - h.beta1) * g[k];
      const float vv = h.beta2 * s1[k] + (1.f - h.beta2) * g[k] * g[k];
      s0[k] = mv;
      s1[k] = vv;
      p[k] = pv - (h.lr / h.bc1) * mv / (sqrtf(vv) / h.bc2_sqrt + h.eps);
    }
  }
}

// Optimizer update of one parameter slot.  kPeers = false: the gradient is the (already averaged) local bucket;
// kPeers = true (one-shot): it is summed from every rank's bucket here.  16-byte vectors, kUU of them per thread per trip
// with every load issued before the first use: the scalar one-element-per-trip loop this replaces was latency-bound (one
// DRAM round trip per element per thread - a 27 MB VGG bucket took 685 us and slowed the conv kernels it overlapped).
template <typename T, bool kPeers>
__device__ __forceinline__ void update_slot(const Peers& P, const Hyper& h, const Slot& s) {
  constexpr int V = Vec<T>::N;
  constexpr int kUU = kPeers ? 1 : 4;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
  const bool has0 = (h.opt == 2) || (h.opt == 1 && h.momentum != 0.f);
  const bool has1 = (h.opt == 2);
  const bool rd0 = has0 && !(h.opt == 1 && h.first_step);
  const uintptr_t align = reinterpret_cast<uintptr_t>(s.param) | (has0 ? reinterpret_cast<uintptr_t>(s.state0) : 0) |
                          (has1 ? reinterpret_cast<uintptr_t>(s.state1) : 0) | static_cast<uintptr_t>(s.offset * sizeof(T));
  const long long nv = (align & 15) ? 0 : s.numel / V;
  uint4* pp = reinterpret_cast<uint4*>(s.param);
  uint4* q0 = reinterpret_cast<uint4*>(s.state0);
  uint4* q1 = reinterpret_cast<uint4*>(s.state1);
  for (long long j0 = tid; j0 < nv; j0 += nth * kUU) {
    uint4 gq[kUU][kPeers ? kMaxRanks : 1], pq[kUU], aq[kUU], bq[kUU];
#pragma unroll
    for (int u = 0; u < kUU; ++u) {
      const long long j = j0 + u * nth;
      if (j < nv) {
        if constexpr (kPeers) {
#pragma unroll
          for (int r = 0; r < kMaxRanks; ++r)
            if (r < P.world) gq[u][r] = reinterpret_cast<const uint4*>(static_cast<const T*>(P.buf[r]) + s.offset)[j];
        } else {
          gq[u][0] = reinterpret_cast<const uint4*>(static_cast<const T*>(P.buf[P.rank]) + s.offset)[j];
        }
        pq[u] = pp[j];
        if (rd0) aq[u] = q0[j];
        if (has1) bq[u] = q1[j];
      }
    }
#pragma unroll
    for (int u = 0; u < kUU; ++u) {
      const long long j = j0 + u * nth;
      if (j >= nv) break;
      float g[V], pf[V], s0[V], s1[V];
      if constexpr (kPeers) {
#pragma unroll
        for (int k = 0; k < V; ++k) g[k] = 0.f;
#pragma unroll
        for (int r = 0; r < kMaxRanks; ++r) {
          if (r < P.world) {
            float f[V];
            unpack<T>(gq[u][r], f);
#pragma unroll
            for (int k = 0; k < V; ++k) g[k] += f[k];
          }
        }
      } else {
        unpack<T>(gq[u][0], g);
      }
      unpack<T>(pq[u], pf);
      if (rd0) unpack<T>(aq[u], s0);
      else {
#pragma unroll
        for (int k = 0; k < V; ++k) s0[k] = 0.f;
      }
      if (has1) unpack<T>(bq[u], s1);
      update_regs<V>(h, g, pf, s0, s1);
      pp[j] = pack<T>(pf);
      if (has0) q0[j] = pack<T>(s0);
      if (has1) q1[j] = pack<T>(s1);
    }
  }
  // scalar tail (and unaligned slots)
  for (long long j = nv * V + tid; j < s.numel; j += nth) {
    float g = 0.f;
    if constexpr (kPeers) {
#pragma unroll 1
      for (int r = 0; r < P.world; ++r) g += ldT(static_cast<const T*>(P.buf[r]) + s.offset + j);
    } else {
      g = ldT(static_cast<const T*>(P.buf[P.rank]) + s.offset + j);
    }
    apply_update<T>(h, s, j, g);
  }
}

// two-shot step 2: reduce vectors [v0, v1) of every rank's bucket in fixed rank order and store the sum into every
// rank's bucket.  kW = ranks the loops are unrolled for, kU = vectors per thread per trip: kW * kU = 16 sixteen-byte peer
// loads are in flight per thread whatever the world size (NVLink load latency ~2-3 us).
template <typename T, int kW, int kU>
__device__ __forceinline__ void reduce_shard(const Peers& P, long long v0, long long v1) {
  constexpr int V = Vec<T>::N;
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  for (long long i0 = v0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < v1; i0 += stride * kU) {
    uint4 in[kU][kW];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * stride;
#pragma unroll
      for (int r = 0; r < kW; ++r)
        if (r < P.world && i < v1) in[u][r] = reinterpret_cast<const uint4*>(P.buf[r])[i];
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * stride;
      if (i >= v1) break;
      float acc[V];
#pragma unroll
      for (int k = 0; k < V; ++k) acc[k] = 0.f;
#pragma unroll
      for (int r = 0; r < kW; ++r) {
        if (r < P.world) {
          float f[V];
          unpack<T>(in[u][r], f);
#pragma unroll
          for (int k = 0; k < V; ++k) acc[k] += f[k];
        }
      }
      const uint4 o = pack<T>(acc);
#pragma unroll
      for (int r = 0; r < kW; ++r)
        if (r < P.world) reinterpret_cast<uint4*>(P.buf[r])[i] = o;
    }
  }
}

// NVLS flavour of step 2 (mode 2): the shard is reduced INSIDE the NVSwitch - one multimem.ld_reduce on the multicast
// address returns the sum over all ranks' buckets - and one multimem.st writes the result into every rank's bucket.
// Per-direction traffic is ~S/world in and S/world out per rank instead of (world-1)/world * S each way.
template <typename T>
__device__ __forceinline__ void reduce_shard_nvls(const Peers& P, long long v0, long long v1) {
  constexpr int kU = 8;      // multimem.ld_reduce round trips go through the switch: keep 8 of them in flight per thread
  const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
  uint4* mc = reinterpret_cast<uint4*>(P.mc);
  for (long long i0 = v0 + static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < v1; i0 += stride * kU) {
    uint4 v[kU];
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * stride;
      if (i < v1) {
        if constexpr (sizeof(T) == 4) {
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
                       : "l"(mc + i)
                       : "memory");
        } else {
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                       : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
                       : "l"(mc + i)
                       : "memory");
        }
      }
    }
#pragma unroll
    for (int u = 0; u < kU; ++u) {
      const long long i = i0 + u * stride;
      if (i < v1) {
        if constexpr (sizeof(T) == 4) {
          asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + i), "r"(v[u].x), "r"(v[u].y),
                       "r"(v[u].z), "r"(v[u].w)
                       : "memory");
        } else {
          asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc + i), "r"(v[u].x),
                       "r"(v[u].y), "r"(v[u].z), "r"(v[u].w)
                       : "memory");
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Shard update (two-shot / NVLS with an optimizer): rank r owns elements [e0, e1) of the bucket.  For every parameter
// slot that intersects the shard it sums the gradient from every rank's bucket (or lets the NVSwitch do it), applies
// SGD-momentum / AdamW to ITS slice of the parameter with ITS slice of the optimizer state, and stores the UPDATED
// PARAMETER into every rank's bucket.  After the hand-over barrier every rank copies the slices it does not own from its
// bucket into its parameters.  Compared with "all-gather the averaged gradient, then every rank updates the whole
// bucket" this reads grad + param + state and writes param + state for 1/world of the bucket instead of all of it; the
// rest is a plain copy.  (The full update of a 27 MB VGG bucket ran next to the bandwidth-bound BatchNorm backward and
// cost it 124 instead of 70 us: profiles/raw/trace_w2_round2.log.)  Optimizer state is therefore current only on its
// owner; FusedUpdate.gather_state() re-replicates it before a state_dict() or a bucket rebuild.
// ------------------------------------------------------------------------------------------------
template <typename T, int kW, int kU, bool kNvls>
__device__ __forceinline__ void shard_update(const Peers& P, const Hyper& h, const SlotTable& slots, long long e0, long long e1) {
  constexpr int V = Vec<T>::N;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
  const bool has0 = (h.opt == 2) || (h.opt == 1 && h.momentum != 0.f);
  const bool has1 = (h.opt == 2);
  const bool rd0 = has0 && !(h.opt == 1 && h.first_step);
  for (int si = 0; si < slots.n; ++si) {
    const Slot& s = slots.s[si];
    const long long a = max(s.offset, e0), b = min(s.offset + s.numel, e1);
    if (a >= b) continue;
    const long long j_first = a - s.offset;                       // first element of the slice inside the parameter
    const uintptr_t align = reinterpret_cast<uintptr_t>(s.param) | (has0 ? reinterpret_cast<uintptr_t>(s.state0) : 0) |
                            (has1 ? reinterpret_cast<uintptr_t>(s.state1) : 0) | static_cast<uintptr_t>(j_first * sizeof(T));
    const long long nv = (align & 15) ? 0 : (b - a) / V;
    uint4* pp = reinterpret_cast<uint4*>(static_cast<T*>(s.param) + j_first);
    uint4* q0 = reinterpret_cast<uint4*>(static_cast<T*>(s.state0) + j_first);
    uint4* q1 = reinterpret_cast<uint4*>(static_cast<T*>(s.state1) + j_first);
    const long long vb = a / V;                                   // vector index of the slice inside the bucket (a % V == 0)
    for (long long j0 = tid; j0 < nv; j0 += nth * kU) {
      uint4 gq[kU][kNvls ? 1 : kW], pq[kU], aq[kU], bq[kU];
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long j = j0 + u * nth;
        if (j < nv) {
          if constexpr (kNvls) {
            const uint4* mc = reinterpret_cast<const uint4*>(P.mc) + vb + j;
            if constexpr (sizeof(T) == 4)
              asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                           : "=r"(gq[u][0].x), "=r"(gq[u][0].y), "=r"(gq[u][0].z), "=r"(gq[u][0].w) : "l"(mc) : "memory");
            else
              asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                           : "=r"(gq[u][0].x), "=r"(gq[u][0].y), "=r"(gq[u][0].z), "=r"(gq[u][0].w) : "l"(mc) : "memory");
          } else {
#pragma unroll
            for (int r = 0; r < kW; ++r)
              if (r < P.world) gq[u][r] = reinterpret_cast<const uint4*>(P.buf[r])[vb + j];
          }
          pq[u] = pp[j];
          if (rd0) aq[u] = q0[j];
          if (has1) bq[u] = q1[j];
        }
      }
#pragma unroll
      for (int u = 0; u < kU; ++u) {
        const long long j = j0 + u * nth;
        if (j >= nv) break;
        float g[V], pf[V], s0[V], s1[V];
        if constexpr (kNvls) {
          unpack<T>(gq[u][0], g);
        } else {
#pragma unroll
          for (int k = 0; k < V; ++k) g[k] = 0.f;
#pragma unroll
          for (int r = 0; r < kW; ++r) {
            if (r < P.world) {
              float f[V];
              unpack<T>(gq[u][r], f);
#pragma unroll
              for (int k = 0; k < V; ++k) g[k] += f[k];
            }
          }
        }
        unpack<T>(pq[u], pf);
        if (rd0) unpack<T>(aq[u], s0);
        else {
#pragma unroll
          for (int k = 0; k < V; ++k) s0[k] = 0.f;
        }
        if (has1) unpack<T>(bq[u], s1);
        update_regs<V>(h, g, pf, s0, s1);
        const uint4 o = pack<T>(pf);
        pp[j] = o;
        if (has0) q0[j] = pack<T>(s0);
        if (has1) q1[j] = pack<T>(s1);
        if constexpr (kNvls) {
          uint4* mc = reinterpret_cast<uint4*>(P.mc) + vb + j;
          if constexpr (sizeof(T) == 4)
            asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
          else
            asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(mc), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
        } else {
#pragma unroll
          for (int r = 0; r < kW; ++r)
            if (r < P.world && r != P.rank) reinterpret_cast<uint4*>(P.buf[r])[vb + j] = o;
        }
      }
    }
    // scalar tail of the slice (and slices whose pointers are not 16-byte aligned)
    for (long long j = nv * V + tid; j < b - a; j += nth) {
      float g = 0.f;
#pragma unroll 1
      for (int r = 0; r < P.world; ++r) g += ldT(static_cast<const T*>(P.buf[r]) + a + j);
      apply_update<T>(h, s, j_first + j, g);
      const T pn = static_cast<const T*>(s.param)[j_first + j];
#pragma unroll 1
      for (int r = 0; r < P.world; ++r)
        if (r != P.rank) static_cast<T*>(P.buf[r])[a + j] = pn;
    }
  }
}

// after the hand-over: parameters <- bucket for everything this rank does not own
template <typename T>
__device__ __forceinline__ void copy_foreign_params(const Peers& P, const SlotTable& slots, long long e0, long long e1) {
  constexpr int V = Vec<T>::N;
  constexpr int kU = 4;
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
  const T* bucket = static_cast<const T*>(P.buf[P.rank]);
  for (int si = 0; si < slots.n; ++si) {
    const Slot& s = slots.s[si];
    const long long end = s.offset + s.numel;
#pragma unroll 1
    for (int part = 0; part < 2; ++part) {
      // part 0: the slice before my shard; part 1: the slice after it
      const long long a = part == 0 ? s.offset : max(s.offset, e1);
      const long long b = part == 0 ? min(end, e0) : end;
      if (a >= b) continue;
      const long long j_first = a - s.offset;
      T* dst = static_cast<T*>(s.param) + j_first;
      const bool aligned = ((reinterpret_cast<uintptr_t>(dst) | static_cast<uintptr_t>(a * sizeof(T))) & 15) == 0;
      const long long nv = aligned ? (b - a) / V : 0;
      const uint4* src4 = reinterpret_cast<const uint4*>(bucket + a);
      uint4* dst4 = reinterpret_cast<uint4*>(dst);
      for (long long j0 = tid; j0 < nv; j0 += nth * kU) {
        uint4 v[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) if (j0 + u * nth < nv) v[u] = src4[j0 + u * nth];
#pragma unroll
        for (int u = 0; u < kU; ++u) if (j0 + u * nth < nv) dst4[j0 + u * nth] = v[u];
      }
      for (long long j = nv * V + tid; j < b - a; j += nth) dst[j] = bucket[a + j];
    }
  }
}

// ------------------------------------------------------------------------------------------------
// fused all-reduce (+ optimizer).  n = bucket elements (multiple of the vector width).
// mode 0 = two-shot, mode 1 = one-shot (requires opt != 0: the local bucket is not rewritten)
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(512)
allreduce_fused_kernel(const Peers P, const uint32_t gen, const long long n, const int mode, const Hyper h,
                       const __grid_constant__ SlotTable slots) {
  constexpr int V = Vec<T>::N;
  uint32_t* pad = P.pad[P.rank];
  const long long nvec = n / V;
  // ---- 1. announce "my bucket holds generation gen", wait for every peer's announcement
  if (blockIdx.x == 0 && threadIdx.x < P.world) st_release_sys(P.pad[threadIdx.x] + PAD_ARRIVE + P.rank, gen);
  if (threadIdx.x < P.world) wait_flag(pad + PAD_ARRIVE + threadIdx.x, gen, pad, 0xA1000000u | threadIdx.x);
  __syncthreads();

  if (mode == 0 || mode == 2) {
    const long long per = (nvec + P.world - 1) / P.world;
    const long long v0 = min(per * P.rank, nvec), v1 = min(v0 + per, nvec);
    if (h.opt != 0) {
      // ---- 2. my shard: reduce (peer loads, fixed rank order => identical bits everywhere, or in the NVSwitch), update my
      //         slice of the parameters / optimizer state, hand the updated parameters to every rank's bucket
      const long long e0 = v0 * V, e1 = v1 * V;
      if (mode == 2) shard_update<T, 1, 4, true>(P, h, slots, e0, e1);
      else if (P.world <= 2) shard_update<T, 2, 4, false>(P, h, slots, e0, e1);
      else if (P.world <= 4) shard_update<T, 4, 2, false>(P, h, slots, e0, e1);
      else shard_update<T, kMaxRanks, 2, false>(P, h, slots, e0, e1);
      // ---- 3. when the whole grid is done with the shard, tell every peer; 4. wait for every other shard
      last_cta_does(pad + PAD_COUNTER, [&] {
        for (int r = 0; r < P.world; ++r) st_release_sys(P.pad[r] + PAD_DONE + P.rank, gen);
      });
      if (threadIdx.x < P.world) wait_flag(pad + PAD_DONE + threadIdx.x, gen, pad, 0xA2000000u | threadIdx.x);
      __syncthreads();
      // ---- 5. the parameters this rank does not own come out of its bucket
      if (P.world > 1) copy_foreign_params<T>(P, slots, e0, e1);
      return;
    }
    // ---- plain all-reduce (no optimizer bound): reduce my shard, store the sum into every rank's bucket
    // U vectors per thread per trip, all peer loads issued before the first add: the ~2 us NVLink load latency needs
    // tens of 16-byte requests in flight per thread to reach link bandwidth
    if (mode == 2) reduce_shard_nvls<T>(P, v0, v1);
    else if (P.world <= 2) reduce_shard<T, 2, 8>(P, v0, v1);
    else if (P.world <= 4) reduce_shard<T, 4, 4>(P, v0, v1);
    else reduce_shard<T, kMaxRanks, 2>(P, v0, v1);
    last_cta_does(pad + PAD_COUNTER, [&] {
      for (int r = 0; r < P.world; ++r) st_release_sys(P.pad[r] + PAD_DONE + P.rank, gen);
    });
    if (threadIdx.x < P.world) wait_flag(pad + PAD_DONE + threadIdx.x, gen, pad, 0xA2000000u | threadIdx.x);
    __syncthreads();
  } else {
    // ---- one-shot: every rank reduces the whole bucket from peer memory and updates its replica directly
    for (int si = 0; si < slots.n; ++si) update_slot<T, true>(P, h, slots.s[si]);
    // exit barrier: nobody may refill its bucket before every peer finished reading it
    last_cta_does(pad + PAD_COUNTER, [&] {
      for (int r = 0; r < P.world; ++r) st_release_sys(P.pad[r] + PAD_EXIT + P.rank, gen);
    });
    if (threadIdx.x < P.world) wait_flag(pad + PAD_EXIT + threadIdx.x, gen, pad, 0xA3000000u | threadIdx.x);
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------
// broadcast (pull model).  Source: pack tensors into its staging slot, publish.  Others: wait, copy out of the source's
// staging (peer loads), acknowledge.  Table entries: {tensor ptr, byte offset in staging, bytes} (16-byte granules).
// ------------------------------------------------------------------------------------------------
struct CopyEntry { void* ptr; long long offset; long long bytes; };
struct CopyTable { CopyEntry e[kMaxSlots]; int n; };

__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, long long bytes) {
  const long long tid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  const long long nth = static_cast<long long>(gridDim.x) * blockDim.x;
  if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
    const long long nv = bytes / 16;
    for (long long i = tid; i < nv; i += nth) reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
    for (long long i = nv * 16 + tid; i < bytes; i += nth) dst[i] = src[i];
  } else {
    for (long long i = tid; i < bytes; i += nth) dst[i] = src[i];
  }
}

__global__ void __launch_bounds__(512)
bcast_src_kernel(const Peers P, const uint32_t gen, const uint32_t prev_gen, uint8_t* staging,
                 const __grid_constant__ CopyTable tab) {
  uint32_t* pad = P.pad[P.rank];
  // prev_gen = the generation in which THIS rank last filled this staging slot as a source (0 = never): wait until every
  // peer has consumed that one.  (Not "gen - 2": with changing sources the previous generations were acknowledged on
  // other ranks' pads, and a source waiting for them on its own pad waited forever.)
  if (prev_gen != 0 && threadIdx.x < P.world && static_cast<int>(threadIdx.x) != P.rank)
    wait_flag(pad + PAD_ACK + threadIdx.x, prev_gen, pad, 0xB1000000u | threadIdx.x);
  __syncthreads();
  for (int i = 0; i < tab.n; ++i)
    copy_bytes(staging + tab.e[i].offset, static_cast<const uint8_t*>(tab.e[i].ptr), tab.e[i].bytes);
  last_cta_does(pad + PAD_COUNTER + 1, [&] {
    for (int r = 0; r < P.world; ++r)
      if (r != P.rank) red_max_release_sys(P.pad[r] + PAD_BCAST, gen);
  });
}

__global__ void __launch_bounds__(512)
bcast_dst_kernel(const Peers P, const int src, const uint32_t gen, const uint8_t* src_staging,
                 const __grid_constant__ CopyTable tab) {
  uint32_t* pad = P.pad[P.rank];
  if (threadIdx.x == 0) wait_flag(pad + PAD_BCAST, gen, pad, 0xB2000000u | src);
  __syncthreads();
  for (int i = 0; i < tab.n; ++i)
    copy_bytes(static_cast<uint8_t*>(tab.e[i].ptr), src_staging + tab.e[i].offset, tab.e[i].bytes);
  last_cta_does(pad + PAD_COUNTER + 1, [&] { st_release_sys(P.pad[src] + PAD_ACK + P.rank, gen); });
}

int fill_peers(Peers& P, int rank, int world, void* const* bufs, void* const* pads) {
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world) return DDPX_ERR_ARG;
  memset(&P, 0, sizeof(P));
  P.rank = rank; P.world = world;
  for (int r = 0; r < world; ++r) {
    P.buf[r] = bufs ? bufs[r] : nullptr;
    P.pad[r] = static_cast<uint32_t*>(pads[r]);
  }
  return DDPX_OK;
}

}  // namespace

#define CUDA_OK(x) ((x) == cudaSuccess ? DDPX_OK : DDPX_ERR_LAUNCH)

// ---- memory + IPC ---------------------------------------------------------------------------------
extern "C" int ddpx_p2p_malloc(void** out, long long bytes) {
  void* p = nullptr;
  if (cudaMalloc(&p, static_cast<size_t>(bytes)) != cudaSuccess) return DDPX_ERR_LAUNCH;
  if (cudaMemset(p, 0, static_cast<size_t>(bytes)) != cudaSuccess) return DDPX_ERR_LAUNCH;
  cudaDeviceSynchronize();
  *out = p;
  return DDPX_OK;
}
extern "C" int ddpx_p2p_free(void* p) { return CUDA_OK(cudaFree(p)); }
extern "C" int ddpx_p2p_export(void* p, unsigned char* handle64) {
  cudaIpcMemHandle_t h;
  if (cudaIpcGetMemHandle(&h, p) != cudaSuccess) return DDPX_ERR_DRIVER;
  memcpy(handle64, &h, sizeof(h));
  return DDPX_OK;
}
extern "C" int ddpx_p2p_import(const unsigned char* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  void* p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return DDPX_ERR_DRIVER;
  *out = p;
  return DDPX_OK;
}
extern "C" int ddpx_p2p_unimport(void* p) { return CUDA_OK(cudaIpcCloseMemHandle(p)); }
extern "C" int ddpx_p2p_can_access(int dev, int peer) {
  int ok = 0;
  if (cudaDeviceCanAccessPeer(&ok, dev, peer) != cudaSuccess) return 0;
  return ok;
}
// Collective sequence numbers.  Every rank issues the same collectives in the same order, so a per-process counter gives
// every all-reduce / broadcast the same generation number on every rank, whoever launches it (the Python backend object
// or the native reducer).
static unsigned int g_gen[2] = {0, 0};
extern "C" unsigned int ddpx_p2p_next_gen(void) { g_gen[0] = (g_gen[0] + 1) & 0x7FFFFFFFu; return g_gen[0]; }
extern "C" unsigned int ddpx_p2p_next_bcast_gen(void) { g_gen[1] = (g_gen[1] + 1) & 0x7FFFFFFFu; return g_gen[1]; }
extern "C" void ddpx_p2p_reset_gens(void) { g_gen[0] = g_gen[1] = 0; }

// device-side wait timeout of the current device, in seconds (> 0)
extern "C" int ddpx_p2p_set_timeout(double seconds) {
  if (!(seconds > 0.0)) return DDPX_ERR_ARG;
  const unsigned long long ns = static_cast<unsigned long long>(seconds * 1e9);
  return CUDA_OK(cudaMemcpyToSymbol(g_timeout_ns, &ns, sizeof(ns)));
}
extern "C" unsigned int ddpx_p2p_read_error(void* pad) {
  uint32_t v = 0;
  cudaMemcpy(&v, static_cast<uint32_t*>(pad) + PAD_ERROR, sizeof(v), cudaMemcpyDeviceToHost);
  return v;
}

// ---- collectives -----------------------------------------------------------------------------------
// slots: arrays of length nslots (param / state0 / state1 pointers, element offsets, element counts); may be 0 slots
// when opt == 0.  hyper: [lr, momentum, dampening, wd, beta1, beta2, eps, bc1, bc2_sqrt].
static int allreduce_impl(int rank, int world, void* const* bufs, void* const* pads, unsigned int gen, long long n,
                          int is_bf16, int mode, int opt, const float* hyper, int nesterov, int first_step, int nslots,
                          void* const* params, void* const* state0, void* const* state1, const long long* offsets,
                          const long long* numels, int ctas, void* stream, void* mc) {
  Peers P;
  int rc = fill_peers(P, rank, world, bufs, pads);
  if (rc != DDPX_OK) return rc;
  P.mc = mc;
  if (nslots > kMaxSlots) return DDPX_ERR_LIMIT;
  if (mode == 1 && opt == 0) return DDPX_ERR_ARG;
  if (mode == 2 && mc == nullptr) return DDPX_ERR_ARG;
  const int vec = is_bf16 ? 8 : 4;
  if (n % vec) return DDPX_ERR_ALIGN;
  SlotTable tab;          // filled per call; passed by value (a local: the launcher is reentrant)
  memset(&tab, 0, sizeof(tab));
  tab.n = nslots;
  for (int i = 0; i < nslots; ++i) {
    tab.s[i].param = params[i];
    tab.s[i].state0 = state0 ? state0[i] : nullptr;
    tab.s[i].state1 = state1 ? state1[i] : nullptr;
    tab.s[i].offset = offsets[i];
    tab.s[i].numel = numels[i];
  }
  Hyper h;
  memset(&h, 0, sizeof(h));
  h.opt = opt;
  if (opt != 0) {
    h.lr = hyper[0]; h.momentum = hyper[1]; h.dampening = hyper[2]; h.wd = hyper[3];
    h.beta1 = hyper[4]; h.beta2 = hyper[5]; h.eps = hyper[6]; h.bc1 = hyper[7]; h.bc2_sqrt = hyper[8];
    h.nesterov = nesterov; h.first_step = first_step;
  }
  if (ctas < 1) ctas = 1;
  if (ctas > 148) ctas = 148;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // Block size: 512 threads measured best end to end (2 GPUs, VGG step: 512 x 64 CTAs 4.47 ms, 256 x 64 4.70 ms,
  // 256 x 148 4.57 ms - bench/run_ddp_ab.sh); communication CTAs do run next to the persistent conv CTAs
  // (bench/coresidency_probe.cu), so what matters is how quickly the kernel is gone again.  DDPX_COMM_THREADS overrides.
  static int threads = 0;
  if (threads == 0) {
    const char* e = getenv("DDPX_COMM_THREADS");
    threads = e ? atoi(e) : 512;
    if (threads < 64 || threads > 512 || threads % 32) threads = 512;
  }
  if (is_bf16) allreduce_fused_kernel<__nv_bfloat16><<<ctas, threads, 0, s>>>(P, gen, n, mode, h, tab);
  else allreduce_fused_kernel<float><<<ctas, threads, 0, s>>>(P, gen, n, mode, h, tab);
  return CUDA_OK(cudaGetLastError());
}

extern "C" int ddpx_p2p_allreduce_fused(int rank, int world, void* const* bufs, void* const* pads, unsigned int gen,
                                        long long n, int is_bf16, int mode, int opt, const float* hyper, int nesterov,
                                        int first_step, int nslots, void* const* params, void* const* state0,
                                        void* const* state1, const long long* offsets, const long long* numels, int ctas,
                                        void* stream) {
  if (mode == 2) return DDPX_ERR_ARG;      // NVLS needs the multicast mapping: ddpx_p2p_allreduce_nvls
  return allreduce_impl(rank, world, bufs, pads, gen, n, is_bf16, mode, opt, hyper, nesterov, first_step, nslots, params,
                        state0, state1, offsets, numels, ctas, stream, nullptr);
}

// NVLS all-reduce (+ optimizer): as two-shot, but the shard is reduced in the NVSwitch through `mc_buf`, the multicast
// mapping of the same symmetric buffer (parallel/nvls.py builds it with cuMulticastCreate / cuMulticastBindMem).
extern "C" int ddpx_p2p_allreduce_nvls(int rank, int world, void* const* bufs, void* mc_buf, void* const* pads,
                                       unsigned int gen, long long n, int is_bf16, int opt, const float* hyper, int nesterov,
                                       int first_step, int nslots, void* const* params, void* const* state0,
                                       void* const* state1, const long long* offsets, const long long* numels, int ctas,
                                       void* stream) {
  return allreduce_impl(rank, world, bufs, pads, gen, n, is_bf16, 2, opt, hyper, nesterov, first_step, nslots, params,
                        state0, state1, offsets, numels, ctas, stream, mc_buf);
}

// broadcast one chunk: `ptrs[i]` (bytes[i] long) <-> staging offset offs[i].  is_src selects the role.
extern "C" int ddpx_p2p_broadcast(int rank, int world, void* const* pads, int src, unsigned int gen, unsigned int prev_gen,
                                  void* staging_of_src, int n, void* const* ptrs, const long long* offs,
                                  const long long* bytes, int ctas, void* stream) {
  Peers P;
  int rc = fill_peers(P, rank, world, nullptr, pads);
  if (rc != DDPX_OK) return rc;
  if (n > kMaxSlots) return DDPX_ERR_LIMIT;
  CopyTable tab;
  memset(&tab, 0, sizeof(tab));
  tab.n = n;
  for (int i = 0; i < n; ++i) { tab.e[i].ptr = ptrs[i]; tab.e[i].offset = offs[i]; tab.e[i].bytes = bytes[i]; }
  if (ctas < 1) ctas = 1;
  if (ctas > 148) ctas = 148;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (rank == src) bcast_src_kernel<<<ctas, 512, 0, s>>>(P, gen, prev_gen, static_cast<uint8_t*>(staging_of_src), tab);
  else bcast_dst_kernel<<<ctas, 512, 0, s>>>(P, src, gen, static_cast<const uint8_t*>(staging_of_src), tab);
  return CUDA_OK(cudaGetLastError());
}
