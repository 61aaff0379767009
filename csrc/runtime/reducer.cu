// Native gradient reducer: everything that happens per gradient and per bucket during backward, without Python.
//
// Counterpart of the hot half of c10d::Reducer (c10d/reducer.hpp: autograd_hook :74, mark_variable_ready :276,
// mark_bucket_ready / all_reduce_bucket :282,524, finalize_backward :286) that torch DDP gives the reference at
// /root/reference/multigpu.py:89.  The autograd post-hooks themselves are C++ too (csrc/torch_ext/autograd_hooks.cpp);
// they call ddpx_nreducer_mark_ready with the gradient's device pointer.  When a parameter completes a bucket - and all
// earlier buckets have been launched (strict bucket order, identical on every rank) - this file
//   1. gathers the bucket's gradients into their slots, pre-divided by the world size (one multi-tensor kernel on the
//      compute stream),
//   2. records an event, makes the communication stream wait for it,
//   3. launches the fused all-reduce (+ shard / replica optimizer update) kernel of csrc/runtime/p2p.cu on the
//      communication stream, bracketed by timing events (per-bucket timers: c10d::Timer, reducer_timer.hpp:13-74),
//   4. records the bucket's completion event; ddpx_nreducer_wait makes the compute stream wait for all of them.
// Python (parallel/reducer.py) keeps construction, the one-time rebuild and the once-per-iteration bookkeeping.
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <algorithm>
#include <vector>

#include "../../include/ddpx_kernels.h"

namespace {

constexpr int kMaxRanks = 8;

struct NBucket {
  std::vector<int> pids;
  std::vector<long long> offsets, numels;
  std::vector<void*> params, state0, state1;     // optimizer slot tables (may stay empty: plain all-reduce)
  void* bufs[kMaxRanks] = {nullptr};
  void* mc = nullptr;
  long long numel = 0;                           // bucket elements (padded)
  int is_bf16 = 0;
  int pending = 0;
  // per iteration
  int mode = 0, opt = 0, nesterov = 0, first_step = 0;
  float hyper[9] = {0};
  cudaEvent_t ready = nullptr, t0 = nullptr, done = nullptr;
  bool launched = false;
  unsigned int gen = 0;
};

struct NReducer {
  int n_params = 0, rank = 0, world = 1, ctas = 64, timing = 0;
  std::vector<int> bucket_of;
  std::vector<const void*> grad;                 // device pointer of every parameter's gradient, this iteration
  std::vector<uint8_t> marked;
  std::vector<int> ready_order;
  std::vector<NBucket> buckets;
  void* pads[kMaxRanks] = {nullptr};
  cudaStream_t comm = nullptr;
  int next_bucket = 0, num_marked = 0, launched = 0;
  float scale = 1.f;
  int last_error = 0;
};

int launch_bucket(NReducer* r, NBucket& b, cudaStream_t compute) {
  // 1. gather: gradients -> bucket slots, times 1/world
  const int n = static_cast<int>(b.pids.size());
  std::vector<void*> dst, src;
  std::vector<long long> numel;
  dst.reserve(n); src.reserve(n); numel.reserve(n);
  char* base = static_cast<char*>(b.bufs[r->rank]);
  const size_t es = b.is_bf16 ? 2 : 4;
  for (int i = 0; i < n; ++i) {
    void* d = base + static_cast<size_t>(b.offsets[i]) * es;
    const void* g = r->grad[b.pids[i]];
    if (g == nullptr) return DDPX_ERR_ARG;
    if (g == d && r->scale == 1.f) continue;     // the gradient already lives in its slot
    dst.push_back(d); src.push_back(const_cast<void*>(g)); numel.push_back(b.numels[i]);
  }
  int rc = DDPX_OK;
  for (size_t i = 0; i < dst.size(); i += 96) {  // ddpx_scale_copy_multi takes up to 96 tensors per launch
    const int m = static_cast<int>(dst.size() - i < 96 ? dst.size() - i : 96);
    rc = ddpx_scale_copy_multi(m, dst.data() + i, src.data() + i, numel.data() + i, r->scale, b.is_bf16, compute);
    if (rc != DDPX_OK) return rc;
  }
  // 2. the communication stream picks the bucket up when the gather is done
  if (cudaEventRecord(b.ready, compute) != cudaSuccess) return DDPX_ERR_LAUNCH;
  if (cudaStreamWaitEvent(r->comm, b.ready, 0) != cudaSuccess) return DDPX_ERR_LAUNCH;
  if (r->timing && cudaEventRecord(b.t0, r->comm) != cudaSuccess) return DDPX_ERR_LAUNCH;
  // 3. one kernel: all-reduce over peer memory + optimizer update
  b.gen = ddpx_p2p_next_gen();
  const int ns = static_cast<int>(b.params.size());
  const long long bytes = b.numel * static_cast<long long>(es);
  int ctas = static_cast<int>(bytes >> 15);      // small buckets are pure latency: few CTAs hand over sooner
  if (ctas < 1) ctas = 1;
  if (ctas > r->ctas) ctas = r->ctas;
  const int opt = ns > 0 ? b.opt : 0;
  int mode = b.mode;
  if (mode == 2 && b.mc == nullptr) mode = 0;
  if (mode == 1 && opt == 0) mode = 0;
  if (mode == 2)
    rc = ddpx_p2p_allreduce_nvls(r->rank, r->world, b.bufs, b.mc, r->pads, b.gen, b.numel, b.is_bf16, opt, b.hyper, b.nesterov,
                                 b.first_step, ns, b.params.data(), ns ? b.state0.data() : nullptr,
                                 ns ? b.state1.data() : nullptr, b.offsets.data(), b.numels.data(), ctas, r->comm);
  else
    rc = ddpx_p2p_allreduce_fused(r->rank, r->world, b.bufs, r->pads, b.gen, b.numel, b.is_bf16, mode, opt, b.hyper,
                                  b.nesterov, b.first_step, ns, b.params.data(), ns ? b.state0.data() : nullptr,
                                  ns ? b.state1.data() : nullptr, b.offsets.data(), b.numels.data(), ctas, r->comm);
  if (rc != DDPX_OK) return rc;
  // 4. completion
  if (cudaEventRecord(b.done, r->comm) != cudaSuccess) return DDPX_ERR_LAUNCH;
  b.launched = true;
  b.mode = mode;
  r->launched += 1;
  return DDPX_OK;
}

}  // namespace

extern "C" void* ddpx_nreducer_create(int n_params, int n_buckets, const int* bucket_of_param, int rank, int world,
                                      void* const* pads, void* comm_stream, int ctas) {
  if (world < 1 || world > kMaxRanks || rank < 0 || rank >= world || n_buckets < 1) return nullptr;
  auto* r = new NReducer();
  r->n_params = n_params; r->rank = rank; r->world = world; r->ctas = ctas > 0 ? ctas : 64;
  r->bucket_of.assign(bucket_of_param, bucket_of_param + n_params);
  r->grad.assign(n_params, nullptr);
  r->marked.assign(n_params, 0);
  r->buckets.resize(n_buckets);
  for (int i = 0; i < world; ++i) r->pads[i] = pads[i];
  r->comm = static_cast<cudaStream_t>(comm_stream);
  r->scale = 1.f / static_cast<float>(world);
  for (auto& b : r->buckets) {
    cudaEventCreateWithFlags(&b.ready, cudaEventDisableTiming);
    cudaEventCreate(&b.t0);
    cudaEventCreate(&b.done);
  }
  return r;
}

extern "C" void ddpx_nreducer_destroy(void* h) {
  auto* r = static_cast<NReducer*>(h);
  if (!r) return;
  for (auto& b : r->buckets) {
    if (b.ready) cudaEventDestroy(b.ready);
    if (b.t0) cudaEventDestroy(b.t0);
    if (b.done) cudaEventDestroy(b.done);
  }
  delete r;
}

// geometry of one bucket: parameters in slot order, their element offsets / counts, every rank's mapping of the flat buffer
extern "C" int ddpx_nreducer_set_bucket(void* h, int b, int n, const int* pids, const long long* offsets, const long long* numels,
                                        void* const* bufs, void* mc, long long numel, int is_bf16) {
  auto* r = static_cast<NReducer*>(h);
  if (b < 0 || b >= static_cast<int>(r->buckets.size()) || n < 1) return DDPX_ERR_ARG;
  NBucket& B = r->buckets[b];
  B.pids.assign(pids, pids + n);
  B.offsets.assign(offsets, offsets + n);
  B.numels.assign(numels, numels + n);
  for (int i = 0; i < r->world; ++i) B.bufs[i] = bufs[i];
  B.mc = mc; B.numel = numel; B.is_bf16 = is_bf16;
  B.params.clear(); B.state0.clear(); B.state1.clear();
  return DDPX_OK;
}

// optimizer slot table of a bucket (n = 0 clears it: the bucket is only averaged)
extern "C" int ddpx_nreducer_set_slots(void* h, int b, int n, void* const* params, void* const* state0, void* const* state1) {
  auto* r = static_cast<NReducer*>(h);
  if (b < 0 || b >= static_cast<int>(r->buckets.size())) return DDPX_ERR_ARG;
  NBucket& B = r->buckets[b];
  if (n != 0 && n != static_cast<int>(B.pids.size())) return DDPX_ERR_ARG;
  B.params.clear();
  if (n) B.params.assign(params, params + n);
  B.state0.resize(n); B.state1.resize(n);
  for (int i = 0; i < n; ++i) { B.state0[i] = state0 ? state0[i] : nullptr; B.state1[i] = state1 ? state1[i] : nullptr; }
  return DDPX_OK;
}

// per iteration, before backward: algorithm and hyper-parameters of a bucket (LambdaLR changes lr every step)
extern "C" int ddpx_nreducer_set_step(void* h, int b, int mode, int opt, const float* hyper9, int nesterov, int first_step) {
  auto* r = static_cast<NReducer*>(h);
  if (b < 0 || b >= static_cast<int>(r->buckets.size())) return DDPX_ERR_ARG;
  NBucket& B = r->buckets[b];
  B.mode = mode; B.opt = opt; B.nesterov = nesterov; B.first_step = first_step;
  if (hyper9) memcpy(B.hyper, hyper9, sizeof(B.hyper)); else memset(B.hyper, 0, sizeof(B.hyper));
  return DDPX_OK;
}

// prepare_for_backward
extern "C" void ddpx_nreducer_arm(void* h, int timing) {
  auto* r = static_cast<NReducer*>(h);
  for (auto& b : r->buckets) { b.pending = static_cast<int>(b.pids.size()); b.launched = false; }
  std::fill(r->marked.begin(), r->marked.end(), 0);
  std::fill(r->grad.begin(), r->grad.end(), nullptr);
  r->ready_order.clear();
  r->next_bucket = 0; r->num_marked = 0; r->launched = 0; r->timing = timing; r->last_error = 0;
}

// One gradient is ready.  Returns a bit mask: 1 = some bucket was launched by this call, 2 = every bucket has now been
// launched (the caller queues the finalisation), 4 = the first bucket of this iteration was launched by this call;
// negative: -1 marked twice, -2 bad index, -(16 + DDPX_ERR_*) a launch failed.
extern "C" int ddpx_nreducer_mark_ready(void* h, int pid, const void* grad_ptr, void* compute_stream) {
  auto* r = static_cast<NReducer*>(h);
  if (pid < 0 || pid >= r->n_params) return -2;
  if (r->marked[pid]) return -1;
  r->marked[pid] = 1;
  r->num_marked += 1;
  r->grad[pid] = grad_ptr;
  r->ready_order.push_back(pid);
  r->buckets[r->bucket_of[pid]].pending -= 1;
  int flags = 0;
  const int nb = static_cast<int>(r->buckets.size());
  while (r->next_bucket < nb && r->buckets[r->next_bucket].pending == 0) {
    const int rc = launch_bucket(r, r->buckets[r->next_bucket], static_cast<cudaStream_t>(compute_stream));
    if (rc != DDPX_OK) { r->last_error = rc; return -(16 + rc); }
    if (r->next_bucket == 0) flags |= 4;
    flags |= 1;
    r->next_bucket += 1;
  }
  if (r->next_bucket == nb) flags |= 2;
  return flags;
}

// finalize_backward: the compute stream waits for every bucket's kernel
extern "C" int ddpx_nreducer_wait(void* h, void* compute_stream) {
  auto* r = static_cast<NReducer*>(h);
  for (auto& b : r->buckets)
    if (b.launched && cudaStreamWaitEvent(static_cast<cudaStream_t>(compute_stream), b.done, 0) != cudaSuccess) return DDPX_ERR_LAUNCH;
  return DDPX_OK;
}

// Device time of every bucket's kernel in the last iteration armed with timing = 1 (milliseconds; -1 = not launched or
// not timed).  Synchronises on the buckets' completion events.  Also reports the mode each bucket actually ran with.
extern "C" int ddpx_nreducer_bucket_times(void* h, float* out_ms, int* out_mode, int max_out) {
  auto* r = static_cast<NReducer*>(h);
  const int nb = static_cast<int>(r->buckets.size());
  for (int i = 0; i < nb && i < max_out; ++i) {
    NBucket& b = r->buckets[i];
    out_ms[i] = -1.f;
    if (out_mode) out_mode[i] = b.mode;
    if (!b.launched || !r->timing) continue;
    if (cudaEventSynchronize(b.done) != cudaSuccess) return -1;
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, b.t0, b.done) == cudaSuccess) out_ms[i] = ms;
  }
  return nb;
}

extern "C" int ddpx_nreducer_ready_order(void* h, int* out, int max_out) {
  auto* r = static_cast<NReducer*>(h);
  const int n = static_cast<int>(r->ready_order.size());
  for (int i = 0; i < n && i < max_out; ++i) out[i] = r->ready_order[i];
  return n;
}

extern "C" int ddpx_nreducer_unmarked(void* h, int* out, int max_out) {
  auto* r = static_cast<NReducer*>(h);
  int n = 0;
  for (int i = 0; i < r->n_params; ++i)
    if (!r->marked[i]) { if (n < max_out) out[n] = i; ++n; }
  return n;
}

extern "C" int ddpx_nreducer_launched(void* h) { return static_cast<NReducer*>(h)->launched; }
