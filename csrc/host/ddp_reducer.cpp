// Host-side core of the bucketed data-parallel gradient reducer (C ABI, no torch dependency).
//
// Stands in for the bookkeeping half of torch's C++ c10d::Reducer (bucket assignment by
// size in reverse registration order with [first, rest] caps, per-parameter ready marking, in-order
// bucket launch, bucket rebuild from the observed ready order after the first iteration).  The
// communication itself is launched by Python on the returned bucket ids: on B200 that is the fused
// NVLS all-reduce + scale + cast kernel (csrc/comm.cu), on CPU it is gloo.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

extern "C" {

// Greedy size-capped assignment.  Tensors are visited in the order given; tensors with different
// `key` (dtype/device class) never share a bucket.  The first bucket of every key closes at
// `first_cap` bytes, later ones at `cap`.  Buckets are finally ordered by their smallest visit
// position, so bucket 0 holds the tensors visited first (= gradients that become ready first).
// Returns the number of buckets; bucket_of[i] = bucket index of the i-th visited tensor.
int b2h_ddp_assign_buckets(const int64_t* nbytes, const int32_t* key, int32_t n, int64_t first_cap, int64_t cap,
                           int32_t* bucket_of) {
  struct Open { std::vector<int32_t> idx; int64_t size = 0; int nclosed = 0; };
  std::map<int32_t, Open> open;
  std::vector<std::vector<int32_t>> closed;
  for (int32_t i = 0; i < n; ++i) {
    Open& o = open[key ? key[i] : 0];
    o.idx.push_back(i);
    o.size += nbytes[i];
    const int64_t limit = o.nclosed == 0 ? first_cap : cap;
    if (o.size >= limit) {
      closed.push_back(std::move(o.idx));
      o.idx.clear();
      o.size = 0;
      o.nclosed++;
    }
  }
  for (auto& kv : open)
    if (!kv.second.idx.empty()) closed.push_back(std::move(kv.second.idx));
  std::sort(closed.begin(), closed.end(), [](const std::vector<int32_t>& a, const std::vector<int32_t>& b) {
    return *std::min_element(a.begin(), a.end()) < *std::min_element(b.begin(), b.end());
  });
  for (size_t b = 0; b < closed.size(); ++b)
    for (int32_t i : closed[b]) bucket_of[i] = static_cast<int32_t>(b);
  return static_cast<int>(closed.size());
}

// ---------------------------------------------------------------------------------------------
// Ready tracking for one backward pass.
// ---------------------------------------------------------------------------------------------
struct B2Reducer {
  std::vector<int32_t> bucket_of;      // param -> bucket
  std::vector<int32_t> bucket_size;    // params per bucket
  std::vector<int32_t> pending;        // params not yet ready, per bucket
  std::vector<uint8_t> ready;          // per param
  std::vector<int32_t> ready_order;    // params in the order they became ready (for rebuild)
  int32_t next_bucket = 0;             // buckets must launch in index order on every rank
  std::mutex mu;
};

void* b2h_reducer_create(const int32_t* bucket_of, int32_t nparams, int32_t nbuckets) {
  auto* r = new B2Reducer();
  r->bucket_of.assign(bucket_of, bucket_of + nparams);
  r->bucket_size.assign(nbuckets, 0);
  for (int32_t i = 0; i < nparams; ++i) r->bucket_size[bucket_of[i]]++;
  r->pending = r->bucket_size;
  r->ready.assign(nparams, 0);
  return r;
}
void b2h_reducer_destroy(void* h) { delete static_cast<B2Reducer*>(h); }

void b2h_reducer_reset(void* h) {
  auto* r = static_cast<B2Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  r->pending = r->bucket_size;
  std::fill(r->ready.begin(), r->ready.end(), 0);
  r->ready_order.clear();
  r->next_bucket = 0;
}

// Marks a parameter's gradient ready.  Writes the ids of buckets that may be launched NOW (complete,
// and every earlier bucket already launched) into `launch` (capacity = nbuckets); returns how many.
// Returns -1 if the parameter was already marked in this pass (a double-ready is a usage error).
int b2h_reducer_mark_ready(void* h, int32_t param, int32_t* launch) {
  auto* r = static_cast<B2Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  if (r->ready[param]) return -1;
  r->ready[param] = 1;
  r->ready_order.push_back(param);
  r->pending[r->bucket_of[param]]--;
  int n = 0;
  while (r->next_bucket < static_cast<int32_t>(r->pending.size()) && r->pending[r->next_bucket] == 0) {
    launch[n++] = r->next_bucket++;
  }
  return n;
}

// End of backward: parameters that never became ready (unused in this iteration) are force-marked;
// returns the remaining launchable buckets.  `unused` receives those parameter ids, *nunused their count.
int b2h_reducer_finalize(void* h, int32_t* launch, int32_t* unused, int32_t* nunused) {
  auto* r = static_cast<B2Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  int nu = 0;
  for (int32_t p = 0; p < static_cast<int32_t>(r->ready.size()); ++p) {
    if (!r->ready[p]) {
      r->ready[p] = 1;
      r->pending[r->bucket_of[p]]--;
      if (unused) unused[nu] = p;
      nu++;
    }
  }
  if (nunused) *nunused = nu;
  int n = 0;
  while (r->next_bucket < static_cast<int32_t>(r->pending.size()) && r->pending[r->next_bucket] == 0) {
    launch[n++] = r->next_bucket++;
  }
  return n;
}

// Copies the observed ready order (param ids) into out; returns its length.
int b2h_reducer_ready_order(void* h, int32_t* out) {
  auto* r = static_cast<B2Reducer*>(h);
  std::lock_guard<std::mutex> g(r->mu);
  std::memcpy(out, r->ready_order.data(), r->ready_order.size() * sizeof(int32_t));
  return static_cast<int>(r->ready_order.size());
}

}  // extern "C"
