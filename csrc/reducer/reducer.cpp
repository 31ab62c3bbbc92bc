// Reducer core: bucket planning and the per-iteration readiness state machine.
//
// The library counterpart is c10d::Reducer, which the reference drives implicitly through
// DistributedDataParallel (ref:pytorch/resnet/main.py:44-46,131; ref:pytorch/unet/train.py:68-70,191).
// Differences by design:
//   * gradients ARE views into the flat bucket arena, so there is no copy-in / copy-out;
//   * buckets are planned once, in reverse registration order (~ gradient-ready order), with the same
//     size rule as torch (first bucket 1 MiB, then 25 MiB; a tensor is added before the cap is checked);
//   * buckets are launched in a fixed LAUNCH ORDER (a bucket goes out once it and every bucket before it in that
//     order is complete), so every rank issues the collectives in the same sequence (required by the peer-to-peer
//     flag protocol).  The order starts as 0, 1, 2, ... and is re-derived once from the gradient-ready order
//     observed in the first backward pass (dmpi_reducer_rebuild, the counterpart of c10d's bucket rebuild): buckets
//     are contiguous ranges of the gradient arena and stay where they are, but a bucket that holds a late gradient
//     (a parked downsample dgrad, a UNet skip branch) no longer holds back the buckets behind it;
//   * the launch itself is a C callback (CUDA glue in reducer_cuda.cu) — Python only forwards the
//     autograd hook.  On the CPU/gloo path no callback is installed and the ready buckets are returned.
#include <stdint.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <vector>

extern "C" {

typedef void (*dmpi_launch_cb)(void* ctx, int bucket);

// sizes: bytes per tensor in bucket-fill order.  Writes bucket_of[i]; returns the number of buckets.
int dmpi_plan_buckets(const long long* sizes, int n, long long first_cap, long long cap, int* bucket_of) {
  int bucket = 0;
  long long fill = 0;
  bool open = false;
  for (int i = 0; i < n; ++i) {
    bucket_of[i] = bucket;
    fill += sizes[i];
    open = true;
    const long long limit = bucket == 0 ? first_cap : cap;
    if (fill >= limit) {
      ++bucket;
      fill = 0;
      open = false;
    }
  }
  return bucket + (open ? 1 : 0);
}

struct Reducer {
  std::mutex mu;
  int n_params = 0, n_buckets = 0;
  std::vector<int> bucket_of;       // param -> bucket
  std::vector<int> bucket_size;     // params per bucket
  std::vector<int> pending;         // per bucket, this iteration
  std::vector<uint8_t> ready;       // per param, this iteration
  std::vector<int> ready_order;     // params in the order their gradients arrived (for plan rebuild)
  std::vector<int> launch_order;    // buckets in the order they are launched
  int next_bucket = 0;              // position in launch_order of the first bucket not yet launched
  long long double_marks = 0;
  dmpi_launch_cb cb = nullptr;
  void* cb_ctx = nullptr;
};

void* dmpi_reducer_create(int n_params, const int* bucket_of, int n_buckets) {
  Reducer* r = new Reducer();
  r->n_params = n_params;
  r->n_buckets = n_buckets;
  r->bucket_of.assign(bucket_of, bucket_of + n_params);
  r->bucket_size.assign(n_buckets, 0);
  for (int i = 0; i < n_params; ++i) r->bucket_size[bucket_of[i]]++;
  r->pending = r->bucket_size;
  r->ready.assign(n_params, 0);
  r->ready_order.reserve(n_params);
  r->launch_order.resize(n_buckets);
  for (int b = 0; b < n_buckets; ++b) r->launch_order[b] = b;
  return r;
}

void dmpi_reducer_destroy(void* h) { delete static_cast<Reducer*>(h); }

void dmpi_reducer_set_launch_cb(void* h, void* cb, void* ctx) {
  Reducer* r = static_cast<Reducer*>(h);
  r->cb = reinterpret_cast<dmpi_launch_cb>(cb);
  r->cb_ctx = ctx;
}

// Start of an iteration that will synchronise gradients.
void dmpi_reducer_reset(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> lk(r->mu);
  r->pending = r->bucket_size;
  std::fill(r->ready.begin(), r->ready.end(), 0);
  r->ready_order.clear();
  r->next_bucket = 0;
}

// Called from the autograd hook of parameter `param`.  Launches (callback) or reports (out_buckets) every
// bucket that became launchable, in bucket order.  Returns how many; -1 if the parameter was already marked
// (the same parameter used twice without gradient accumulation semantics -> caller raises).
int dmpi_reducer_mark_ready(void* h, int param, int* out_buckets, int max_out) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> lk(r->mu);
  if (param < 0 || param >= r->n_params) return -2;
  if (r->ready[param]) {
    r->double_marks++;
    return -1;
  }
  r->ready[param] = 1;
  r->ready_order.push_back(param);
  r->pending[r->bucket_of[param]]--;
  int n = 0;
  while (r->next_bucket < r->n_buckets && r->pending[r->launch_order[r->next_bucket]] == 0) {
    const int b = r->launch_order[r->next_bucket++];
    if (r->cb) r->cb(r->cb_ctx, b);
    if (n < max_out) out_buckets[n] = b;
    ++n;
  }
  return n;
}

int dmpi_reducer_num_buckets(void* h) { return static_cast<Reducer*>(h)->n_buckets; }

// Buckets not launched yet this iteration (0 after a complete backward; > 0 means unused parameters).
int dmpi_reducer_pending(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> lk(r->mu);
  return r->n_buckets - r->next_bucket;
}

int dmpi_reducer_ready_order(void* h, int* out, int max_out) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> lk(r->mu);
  const int n = static_cast<int>(r->ready_order.size());
  for (int i = 0; i < n && i < max_out; ++i) out[i] = r->ready_order[i];
  return n;
}

// Re-derives the launch order from the gradient-ready order of the iteration that just finished (call it only after a
// complete backward pass): buckets sorted by the arrival position of their LAST gradient, ties by index.  Returns the
// number of buckets whose position changed, or -1 if the iteration was incomplete (order left as it was).
int dmpi_reducer_rebuild(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> lk(r->mu);
  if (static_cast<int>(r->ready_order.size()) != r->n_params) return -1;
  std::vector<int> last(r->n_buckets, -1);
  for (int pos = 0; pos < r->n_params; ++pos) last[r->bucket_of[r->ready_order[pos]]] = pos;
  std::vector<int> order(r->n_buckets);
  for (int b = 0; b < r->n_buckets; ++b) order[b] = b;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return last[a] < last[b]; });
  int moved = 0;
  for (int i = 0; i < r->n_buckets; ++i) moved += order[i] != r->launch_order[i];
  r->launch_order = order;
  return moved;
}

// Installs a launch order (a permutation of 0..n_buckets-1; e.g. back to the identity when ranks disagree).
int dmpi_reducer_set_launch_order(void* h, const int* order, int n) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> lk(r->mu);
  if (n != r->n_buckets) return -1;
  std::vector<uint8_t> seen(n, 0);
  for (int i = 0; i < n; ++i) {
    if (order[i] < 0 || order[i] >= n || seen[order[i]]) return -2;
    seen[order[i]] = 1;
  }
  r->launch_order.assign(order, order + n);
  return 0;
}

int dmpi_reducer_launch_order(void* h, int* out, int max_out) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> lk(r->mu);
  for (int i = 0; i < r->n_buckets && i < max_out; ++i) out[i] = r->launch_order[i];
  return r->n_buckets;
}

// FNV-1a over the bucket layout, exchanged through the store so ranks can assert identical plans
// (analogue of Reducer::sync_bucket_indices + _verify_param_shape_across_processes).
unsigned long long dmpi_plan_hash(const long long* sizes, const int* bucket_of, int n) {
  unsigned long long hsh = 1469598103934665603ull;
  auto mix = [&](unsigned long long v) {
    for (int i = 0; i < 8; ++i) {
      hsh ^= (v >> (8 * i)) & 0xffull;
      hsh *= 1099511628211ull;
    }
  };
  mix(static_cast<unsigned long long>(n));
  for (int i = 0; i < n; ++i) {
    mix(static_cast<unsigned long long>(sizes[i]));
    mix(static_cast<unsigned long long>(bucket_of[i]));
  }
  return hsh;
}

}  // extern "C"
