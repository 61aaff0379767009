// Host-side core of the gradient reducer: bucket assignment and the ready/launch state machine.
//
// Plays the role of c10d's compute_bucket_assignment_by_size (c10d/reducer.hpp:592-598) and of the bookkeeping
// half of c10d::Reducer (mark_variable_ready -> mark_bucket_ready -> in-order launch, reducer.hpp:276-282, next_bucket_
// :231) that torch DDP instantiates for the reference at multigpu.py:89.  Plain C ABI; the tensor side (bucket
// memory, kernels, streams) lives in parallel/reducer.py and the communication backend.
#include <algorithm>
#include <cstdint>
#include <map>
#include <vector>

#include "../../include/ddpx_kernels.h"

namespace {

struct Acc {
  std::vector<int> positions;
  long long size = 0;
  size_t limit_idx = 0;
};

struct ReducerCore {
  int n_params = 0;
  int n_buckets = 0;
  std::vector<int> bucket_of;      // param -> bucket
  std::vector<int> bucket_count;   // params per bucket
  std::vector<int> pending;        // per bucket, this iteration
  std::vector<uint8_t> marked;     // per param, this iteration
  std::vector<int> ready_order;    // params in the order they were marked
  int next_bucket = 0;             // buckets launch strictly in index order
  int num_marked = 0;
};

}  // namespace

// Greedy size-capped bucketing by key (dtype/device id), in the order given.  limits[] is consumed per key: a key's
// first bucket closes at limits[0], its next at limits[1], ... the last limit repeats.  A bucket closes when its size
// reaches its limit (so one large tensor may exceed it).  Leftovers are emitted in first-seen key order.  When
// sort_by_min_position != 0 the buckets are ordered by their smallest position (torch does this when no explicit
// tensor order is supplied).  Outputs: bucket id per position, the byte limit per bucket, the bucket count.
extern "C" int ddpx_bucket_assign(int n, const long long* nbytes, const int* keys, int nlimits, const long long* limits,
                                  int sort_by_min_position, int* out_bucket_of_position, long long* out_bucket_limit,
                                  int* out_num_buckets) {
  if (n < 0 || nlimits < 1) return DDPX_ERR_ARG;
  std::map<int, Acc> open;
  std::map<int, size_t> limit_iter;
  std::vector<int> key_order;
  std::vector<std::pair<std::vector<int>, long long>> result;
  for (int i = 0; i < n; ++i) {
    const int key = keys ? keys[i] : 0;
    if (!open.count(key)) {
      if (std::find(key_order.begin(), key_order.end(), key) == key_order.end()) key_order.push_back(key);
    }
    Acc& a = open[key];
    a.positions.push_back(i);
    a.size += nbytes[i];
    if (!limit_iter.count(key)) limit_iter[key] = 0;
    const long long lim = limits[limit_iter[key]];
    if (a.size >= lim) {
      result.emplace_back(a.positions, lim);
      open.erase(key);
      if (limit_iter[key] + 1 < static_cast<size_t>(nlimits)) limit_iter[key] += 1;
    }
  }
  for (int key : key_order) {
    auto it = open.find(key);
    if (it != open.end() && !it->second.positions.empty())
      result.emplace_back(it->second.positions, limits[limit_iter[key]]);
  }
  if (sort_by_min_position) {
    std::stable_sort(result.begin(), result.end(), [](const auto& a, const auto& b) {
      return *std::min_element(a.first.begin(), a.first.end()) < *std::min_element(b.first.begin(), b.first.end());
    });
  }
  for (size_t b = 0; b < result.size(); ++b) {
    for (int pos : result[b].first) out_bucket_of_position[pos] = static_cast<int>(b);
    out_bucket_limit[b] = result[b].second;
  }
  *out_num_buckets = static_cast<int>(result.size());
  return DDPX_OK;
}

extern "C" void* ddpx_reducer_create(int n_params, int n_buckets, const int* bucket_of_param) {
  auto* r = new ReducerCore();
  r->n_params = n_params;
  r->n_buckets = n_buckets;
  r->bucket_of.assign(bucket_of_param, bucket_of_param + n_params);
  r->bucket_count.assign(n_buckets, 0);
  for (int i = 0; i < n_params; ++i) {
    if (r->bucket_of[i] < 0 || r->bucket_of[i] >= n_buckets) { delete r; return nullptr; }
    r->bucket_count[r->bucket_of[i]] += 1;
  }
  r->pending = r->bucket_count;
  r->marked.assign(n_params, 0);
  return r;
}

extern "C" void ddpx_reducer_destroy(void* h) { delete static_cast<ReducerCore*>(h); }

// prepare_for_backward: re-arm every bucket
extern "C" void ddpx_reducer_reset(void* h) {
  auto* r = static_cast<ReducerCore*>(h);
  r->pending = r->bucket_count;
  std::fill(r->marked.begin(), r->marked.end(), 0);
  r->ready_order.clear();
  r->next_bucket = 0;
  r->num_marked = 0;
}

// Marks one parameter's gradient ready.  Returns the number of buckets that became launchable (written, in launch
// order, to out_buckets); -1 if the parameter was already marked this iteration (e.g. used twice without
// static-graph support, reducer.hpp:559-568), -2 on a bad index.
extern "C" int ddpx_reducer_mark_ready(void* h, int param, int* out_buckets, int max_out) {
  auto* r = static_cast<ReducerCore*>(h);
  if (param < 0 || param >= r->n_params) return -2;
  if (r->marked[param]) return -1;
  r->marked[param] = 1;
  r->num_marked += 1;
  r->ready_order.push_back(param);
  r->pending[r->bucket_of[param]] -= 1;
  int n = 0;
  while (r->next_bucket < r->n_buckets && r->pending[r->next_bucket] == 0) {
    if (n < max_out) out_buckets[n] = r->next_bucket;
    ++n;
    r->next_bucket += 1;
  }
  return n;
}

extern "C" int ddpx_reducer_all_launched(void* h) {
  auto* r = static_cast<ReducerCore*>(h);
  return r->next_bucket == r->n_buckets ? 1 : 0;
}

extern "C" int ddpx_reducer_num_marked(void* h) { return static_cast<ReducerCore*>(h)->num_marked; }

extern "C" int ddpx_reducer_ready_order(void* h, int* out, int max_out) {
  auto* r = static_cast<ReducerCore*>(h);
  const int n = static_cast<int>(r->ready_order.size());
  for (int i = 0; i < n && i < max_out; ++i) out[i] = r->ready_order[i];
  return n;
}

// parameters that never produced a gradient this iteration (diagnostics for the "unused parameter" error)
extern "C" int ddpx_reducer_unmarked(void* h, int* out, int max_out) {
  auto* r = static_cast<ReducerCore*>(h);
  int n = 0;
  for (int i = 0; i < r->n_params; ++i)
    if (!r->marked[i]) {
      if (n < max_out) out[n] = i;
      ++n;
    }
  return n;
}
