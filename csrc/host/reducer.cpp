// Reducer core: the bookkeeping of one data-parallel backward pass.
//
// What the reference gets from torch's C++ Reducer (constructed by DistributedDataParallel at
// launch_dist.py:61 / example_launch.py:29): every parameter's gradient is marked ready once per
// iteration by an autograd hook; a bucket may be reduced when all of its gradients are ready AND every
// earlier bucket has been launched (all ranks must issue the collectives in the same order); at the end
// of backward the buckets that never filled up (parameters without a gradient this iteration) are
// launched anyway; the order in which gradients arrived during the first iteration is what the one-off
// bucket rebuild is computed from.  This file is that state machine; tensors, streams and kernels stay
// on the Python side (pytorch_distributed_learning_b200/parallel/ddp.py).
#include <cstdint>
#include <vector>

namespace {

struct Reducer {
  std::vector<int32_t> bucket_of;  // parameter -> bucket
  std::vector<int32_t> size;       // bucket -> number of parameters
  std::vector<int32_t> pending;    // bucket -> gradients still missing this iteration
  std::vector<uint8_t> ready;      // parameter -> already marked this iteration
  std::vector<int32_t> arrival;    // parameters in the order they were marked (when recording)
  int next = 0;                    // first bucket not launched yet
  bool record = false;
  bool armed = false;              // prepare() was called and finalize() has not run yet
};

void set_layout(Reducer* r, int num_params, const int32_t* bucket_of, int num_buckets) {
  r->bucket_of.assign(bucket_of, bucket_of + num_params);
  r->size.assign(num_buckets, 0);
  for (int i = 0; i < num_params; ++i) ++r->size[bucket_of[i]];
  r->pending = r->size;
  r->ready.assign(num_params, 0);
  r->arrival.clear();
  r->next = 0;
  r->armed = false;
}

bool valid_layout(int num_params, const int32_t* bucket_of, int num_buckets) {
  if (num_params <= 0 || num_buckets <= 0 || !bucket_of) return false;
  for (int i = 0; i < num_params; ++i)
    if (bucket_of[i] < 0 || bucket_of[i] >= num_buckets) return false;
  return true;
}

}  // namespace

extern "C" {

void* pdl_reducer_create(int num_params, const int32_t* bucket_of, int num_buckets) {
  if (!valid_layout(num_params, bucket_of, num_buckets)) return nullptr;
  Reducer* r = new Reducer();
  set_layout(r, num_params, bucket_of, num_buckets);
  return r;
}

void pdl_reducer_destroy(void* h) { delete static_cast<Reducer*>(h); }

// New bucket layout (after the rebuild that follows the first iteration).  Returns 0, or -1 for an invalid layout.
int pdl_reducer_set_layout(void* h, int num_params, const int32_t* bucket_of, int num_buckets) {
  if (!h || !valid_layout(num_params, bucket_of, num_buckets)) return -1;
  set_layout(static_cast<Reducer*>(h), num_params, bucket_of, num_buckets);
  return 0;
}

// Start of an iteration (DDP forward with gradient synchronisation enabled).
void pdl_reducer_prepare(void* h, int record_arrival) {
  Reducer* r = static_cast<Reducer*>(h);
  r->pending = r->size;
  r->ready.assign(r->ready.size(), 0);
  r->next = 0;
  r->record = record_arrival != 0;
  if (r->record) r->arrival.clear();
  r->armed = true;
}

// Gradient of `param` is final.  Writes the buckets that may be launched NOW, in launch order, to out[0..cap) and
// returns their count.  -1: unknown parameter; -2: second gradient for this parameter in one iteration; -3: out too small.
int pdl_reducer_mark_ready(void* h, int param, int32_t* out, int cap) {
  Reducer* r = static_cast<Reducer*>(h);
  if (param < 0 || param >= static_cast<int>(r->bucket_of.size())) return -1;
  if (!r->armed) return 0;  // a backward without a preceding forward through the wrapper: not synchronised (as in torch)
  if (r->ready[param]) return -2;
  r->ready[param] = 1;
  if (r->record) r->arrival.push_back(param);
  --r->pending[r->bucket_of[param]];
  int n = 0;
  const int nb = static_cast<int>(r->size.size());
  while (r->next < nb && r->pending[r->next] == 0) {
    if (n >= cap) return -3;
    out[n++] = r->next++;
  }
  return n;
}

// End of backward: the buckets not launched yet, in order (their missing gradients count as zeros).  -3: out too small.
int pdl_reducer_finalize(void* h, int32_t* out, int cap) {
  Reducer* r = static_cast<Reducer*>(h);
  if (!r->armed) return 0;
  const int nb = static_cast<int>(r->size.size());
  int n = 0;
  while (r->next < nb) {
    if (n >= cap) return -3;
    r->pending[r->next] = 0;
    out[n++] = r->next++;
  }
  r->armed = false;
  return n;
}

// 1 if `param` has delivered its gradient this iteration.
int pdl_reducer_is_ready(void* h, int param) {
  Reducer* r = static_cast<Reducer*>(h);
  return (param >= 0 && param < static_cast<int>(r->ready.size()) && r->ready[param]) ? 1 : 0;
}

int pdl_reducer_next_bucket(void* h) { return static_cast<Reducer*>(h)->next; }

// Arrival order recorded since the last prepare(record_arrival = 1).  Returns the count (copies at most cap entries).
int pdl_reducer_arrival(void* h, int32_t* out, int cap) {
  Reducer* r = static_cast<Reducer*>(h);
  const int n = static_cast<int>(r->arrival.size());
  for (int i = 0; i < n && i < cap; ++i) out[i] = r->arrival[i];
  return n;
}

}  // extern "C"
