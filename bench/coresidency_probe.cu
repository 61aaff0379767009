// Can a small "communication-shaped" kernel run NEXT TO a persistent kernel that owns ~226 KB of shared memory on every
// SM, or does it wait for that kernel to finish?  (Decides how the fused all-reduce overlaps the conv kernels.)
//   victim  : 148 CTAs x 384 threads, <= 96 registers, DYN bytes of dynamic shared memory, spins ~400 us
//   intruder: 64 CTAs x THREADS threads, ~REGS registers, launched 20 us later on a high-priority stream
// Reports when the intruder finished relative to the victim.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 bench/coresidency_probe.cu -o build/coresidency_probe
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

__global__ void __maxnreg__(96) victim(long long cycles, int* sink) {
  extern __shared__ unsigned char sm[];
  const long long t0 = clock64();
  if (threadIdx.x == 0) sm[0] = 1;
  while (clock64() - t0 < cycles) { __nanosleep(200); }
  if (threadIdx.x == 0 && sm[0] == 77) *sink = 1;
}

template <int kRegs>
__global__ void __launch_bounds__(512) intruder(long long cycles, float* out) {
  float r[kRegs];
#pragma unroll
  for (int i = 0; i < kRegs; ++i) r[i] = threadIdx.x * 0.5f + i;
  const long long t0 = clock64();
  while (clock64() - t0 < cycles) {
#pragma unroll
    for (int i = 0; i < kRegs; ++i) r[i] = r[i] * 1.0001f + r[(i + 1) % kRegs];
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < kRegs; ++i) s += r[i];
  if (s == 12345.678f) out[0] = s;
}

template <int kRegs>
static void run(const char* name, size_t victim_dyn, int threads, int carveout, cudaStream_t sv, cudaStream_t si, int* sink,
                float* out) {
  cudaFuncSetAttribute(victim, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)victim_dyn);
  if (carveout >= 0) cudaFuncSetAttribute(intruder<kRegs>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
  cudaEvent_t v0, v1, i0, i1;
  cudaEventCreate(&v0); cudaEventCreate(&v1); cudaEventCreate(&i0); cudaEventCreate(&i1);
  for (int rep = 0; rep < 2; ++rep) {
    cudaDeviceSynchronize();
    cudaEventRecord(v0, sv);
    victim<<<148, 384, victim_dyn, sv>>>(800000, sink);       // ~400 us
    cudaEventRecord(v1, sv);
    cudaEventRecord(i0, si);
    intruder<kRegs><<<64, threads, 0, si>>>(40000, out);      // ~20 us of work
    cudaEventRecord(i1, si);
    cudaDeviceSynchronize();
  }
  float tv, ti_end, ti;
  cudaEventElapsedTime(&tv, v0, v1);
  cudaEventElapsedTime(&ti_end, v0, i1);
  cudaEventElapsedTime(&ti, i0, i1);
  cudaFuncAttributes fa;
  cudaFuncGetAttributes(&fa, intruder<kRegs>);
  printf("%-46s victim %.0f us | intruder (%d regs) done %.0f us after victim start (own span %.0f us) -> %s\n", name,
         tv * 1e3, fa.numRegs, ti_end * 1e3, ti * 1e3, ti_end < tv * 0.8f ? "CO-RESIDENT" : "waited for the victim");
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) printf("  CUDA error: %s\n", cudaGetErrorString(e));
}

int main() {
  int lo, hi;
  cudaDeviceGetStreamPriorityRange(&lo, &hi);
  cudaStream_t sv, si;
  cudaStreamCreateWithFlags(&sv, cudaStreamNonBlocking);
  cudaStreamCreateWithPriority(&si, cudaStreamNonBlocking, hi);
  int* sink; float* out;
  cudaMalloc(&sink, 4); cudaMalloc(&out, 4);
  run<8>("dyn 226KB, intruder 256thr small, default", 226 * 1024, 256, -1, sv, si, sink, out);
  run<8>("dyn 226KB, intruder 256thr small, carveout100", 226 * 1024, 256, 100, sv, si, sink, out);
  run<8>("dyn 200KB, intruder 256thr small, carveout100", 200 * 1024, 256, 100, sv, si, sink, out);
  run<8>("dyn 100KB, intruder 256thr small, carveout100", 100 * 1024, 256, 100, sv, si, sink, out);
  run<80>("dyn 200KB, intruder 256thr big, carveout100", 200 * 1024, 256, 100, sv, si, sink, out);
  run<80>("dyn 226KB, intruder 256thr big, carveout100", 226 * 1024, 256, 100, sv, si, sink, out);
  run<80>("dyn 200KB, intruder 512thr big, carveout100", 200 * 1024, 512, 100, sv, si, sink, out);
  run<8>("dyn 16KB, intruder 256thr small, default", 16 * 1024, 256, -1, sv, si, sink, out);
  return 0;
}
