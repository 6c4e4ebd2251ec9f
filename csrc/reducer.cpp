// Host-side gradient bucket planner and ready-counting reducer core.
//
// Reference behaviour being reproduced (torch DDP, reached from `DDP(model)` at
// src/train_dist.py:63; semantics pinned from torch/nn/parallel/distributed.py:1252-1276):
// parameters are bucketed in REVERSE registration order (gradients become ready roughly
// back-to-front), the first bucket is capped at 1 MiB and the others at 25 MiB, and a
// bucket closes once its size reaches the cap. For `Net` everything lands in one bucket.
//
// The reducer core counts per-bucket "gradient ready" marks during one backward pass and
// reports when a bucket is complete, in which order buckets completed, and whether the
// pass finished. Stream/event wiring and kernel launches live in Python
// (parallel/reducer.py); this part is pure host logic and is unit-tested without a GPU.
#include <cstdint>
#include <cstring>
#include <mutex>
#include <vector>

extern "C" void b200_set_error(const char* fmt, ...);

namespace {

struct Reducer {
  std::vector<int> bucket_of;       // param index -> bucket
  std::vector<int> bucket_size;     // params per bucket
  std::vector<int> pending;         // params still missing per bucket in this pass
  std::vector<uint8_t> marked;      // per-param: already marked in this pass
  std::vector<int> completion_order;
  int buckets_done = 0;
  std::mutex mu;                    // autograd hooks may fire from another thread
};

}  // namespace

extern "C" {

// Assign n parameters (byte sizes in registration order) to buckets.
// bucket_of[i] receives the bucket index of parameter i; bucket 0 is the one that fills
// first during backward (it holds the LAST registered parameters). Returns the number of
// buckets, or -1 on error.
int b200_plan_buckets(const long long* bytes, int n, long long first_cap, long long cap,
                      int* bucket_of) {
  if (n <= 0 || !bytes || !bucket_of || first_cap <= 0 || cap <= 0) {
    b200_set_error("b200_plan_buckets: bad arguments");
    return -1;
  }
  int bucket = 0;
  long long acc = 0;
  long long limit = first_cap;
  for (int i = n - 1; i >= 0; --i) {
    bucket_of[i] = bucket;
    acc += bytes[i];
    if (acc >= limit && i > 0) {  // bucket closes only after reaching the cap
      ++bucket;
      acc = 0;
      limit = cap;
    }
  }
  return bucket + 1;
}

// One bucket per parameter group boundary given explicitly (e.g. per layer):
// boundaries[k] = index of the first parameter (registration order) of group k, ascending.
// Groups are numbered back-to-front like b200_plan_buckets.
int b200_plan_buckets_by_group(int n, const int* boundaries, int ngroups, int* bucket_of) {
  if (n <= 0 || ngroups <= 0 || !boundaries || !bucket_of || boundaries[0] != 0) {
    b200_set_error("b200_plan_buckets_by_group: bad arguments");
    return -1;
  }
  for (int k = 0; k < ngroups; ++k) {
    const int lo = boundaries[k];
    const int hi = (k + 1 < ngroups) ? boundaries[k + 1] : n;
    if (lo >= hi || hi > n) {
      b200_set_error("b200_plan_buckets_by_group: boundaries must be strictly ascending");
      return -1;
    }
    for (int i = lo; i < hi; ++i) bucket_of[i] = ngroups - 1 - k;
  }
  return ngroups;
}

void* b200_reducer_create(const int* bucket_of, int n, int nbuckets) {
  if (n <= 0 || nbuckets <= 0 || !bucket_of) {
    b200_set_error("b200_reducer_create: bad arguments");
    return nullptr;
  }
  Reducer* r = new Reducer();
  r->bucket_of.assign(bucket_of, bucket_of + n);
  r->bucket_size.assign(nbuckets, 0);
  for (int i = 0; i < n; ++i) {
    if (bucket_of[i] < 0 || bucket_of[i] >= nbuckets) {
      b200_set_error("b200_reducer_create: bucket index out of range");
      delete r;
      return nullptr;
    }
    r->bucket_size[bucket_of[i]]++;
  }
  r->pending = r->bucket_size;
  r->marked.assign(n, 0);
  return r;
}

void b200_reducer_destroy(void* h) { delete static_cast<Reducer*>(h); }

// Start a backward pass: every bucket expects all of its parameters again.
void b200_reducer_reset(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  r->pending = r->bucket_size;
  std::fill(r->marked.begin(), r->marked.end(), 0);
  r->completion_order.clear();
  r->buckets_done = 0;
}

// Mark parameter i's gradient ready. Returns the bucket index if that completed the
// bucket, -1 if the bucket is still waiting, -2 if the parameter was already marked in
// this pass (gradient accumulated twice: the caller must treat it as an error, like
// torch's "marked ready twice"), -3 on a bad index.
int b200_reducer_mark_ready(void* h, int i) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  if (i < 0 || i >= (int)r->bucket_of.size()) return -3;
  if (r->marked[i]) return -2;
  r->marked[i] = 1;
  const int b = r->bucket_of[i];
  if (--r->pending[b] == 0) {
    r->completion_order.push_back(b);
    r->buckets_done++;
    return b;
  }
  return -1;
}

int b200_reducer_all_done(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  return r->buckets_done == (int)r->bucket_size.size() ? 1 : 0;
}

// Buckets that were not completed in this pass (unused parameters); returns the count and
// writes up to cap indices.
int b200_reducer_pending_buckets(void* h, int* out, int cap) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  int k = 0;
  for (int b = 0; b < (int)r->pending.size(); ++b)
    if (r->pending[b] != 0) {
      if (k < cap) out[k] = b;
      ++k;
    }
  return k;
}

int b200_reducer_completion_order(void* h, int* out, int cap) {
  Reducer* r = static_cast<Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  const int n = (int)r->completion_order.size();
  for (int i = 0; i < n && i < cap; ++i) out[i] = r->completion_order[i];
  return n;
}

}  // extern "C"
