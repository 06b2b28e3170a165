// b200flow — readiness bookkeeping of the bucketed gradient reducer (host C++).
//
// The part of torch's C++ `Reducer` (and of Horovod's background coordinator) that runs on every
// gradient-ready event: count down the bucket the parameter belongs to and say when a bucket has
// become complete, so that its all-reduce (kernel K1) can be launched while backward continues
// (SURVEY.md §2E "DDP Reducer", §2F "DP", §2H C1/C2).  The bucket plan comes from bucket_plan.cpp;
// launching stays with the caller, which owns the streams.
//
// A parameter may be used several times per backward (tied weights, `ddp_uses`): it counts as ready
// after its last use.  A bucket is handed out exactly once per step, either by the mark that
// completes it or by the sweep at the end of backward that collects buckets some of whose
// parameters received no gradient.
//
// Buckets are handed out strictly in bucket-index order: bucket b only after buckets 0..b-1.  The
// cross-rank matching of the reductions (epoch-keyed flags of K1, NCCL call order) depends on every
// rank launching the same sequence, and the order in which autograd hooks fire may differ between
// ranks (data-dependent branches, unused heads).  A complete bucket therefore waits behind an
// incomplete earlier one, exactly as in torch's Reducer.
#include <cstdint>
#include <new>
#include <vector>

namespace {

struct Reducer {
  std::vector<int32_t> slot_bucket;   // slot -> bucket
  std::vector<int32_t> bucket_slots;  // parameters per bucket
  std::vector<int32_t> uses_init;     // per slot: gradient-ready events expected per step
  std::vector<int32_t> uses;          // ... still outstanding
  std::vector<int32_t> pending;       // per bucket: slots not yet ready
  std::vector<uint8_t> launched;      // per bucket: already handed out this step
  int next = 0;                       // first bucket not handed out yet (in-order cursor)
};

}  // namespace

extern "C" {

// plan: nb rows of {first_slot, last_slot_exclusive, elem_begin, elem_end} (b2f_plan_buckets);
// uses: nslots entries or NULL (= 1 each).  Returns NULL when the plan does not cover every slot once.
void* b2f_reducer_create(const int64_t* plan, int nb, int nslots, const int32_t* uses) {
  if (!plan || nb <= 0 || nslots <= 0) return nullptr;
  Reducer* r = new (std::nothrow) Reducer();
  if (!r) return nullptr;
  r->slot_bucket.assign(nslots, -1);
  r->bucket_slots.assign(nb, 0);
  for (int b = 0; b < nb; ++b) {
    const int64_t s0 = plan[4 * b], s1 = plan[4 * b + 1];
    if (s0 < 0 || s1 > nslots || s0 >= s1) { delete r; return nullptr; }
    for (int64_t s = s0; s < s1; ++s) {
      if (r->slot_bucket[s] != -1) { delete r; return nullptr; }
      r->slot_bucket[s] = b;
    }
    r->bucket_slots[b] = static_cast<int32_t>(s1 - s0);
  }
  for (int s = 0; s < nslots; ++s)
    if (r->slot_bucket[s] < 0) { delete r; return nullptr; }
  r->uses_init.assign(nslots, 1);
  if (uses)
    for (int s = 0; s < nslots; ++s) r->uses_init[s] = uses[s] > 0 ? uses[s] : 1;
  r->uses = r->uses_init;
  r->pending = r->bucket_slots;
  r->launched.assign(nb, 0);
  return r;
}

void b2f_reducer_destroy(void* h) { delete static_cast<Reducer*>(h); }

// Re-arm for the next step.
void b2f_reducer_reset(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  r->uses = r->uses_init;
  r->pending = r->bucket_slots;
  r->launched.assign(r->launched.size(), 0);
  r->next = 0;
}

// The next bucket in index order if it is complete (now marked as handed out), or -1.  Call until -1
// after every mark that returned a bucket: one event can release a run of consecutive buckets.
int b2f_reducer_next_ready(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  const int nb = static_cast<int>(r->launched.size());
  if (r->next >= nb || r->pending[r->next] != 0) return -1;
  const int b = r->next++;
  r->launched[b] = 1;
  return b;
}

// One gradient-ready event of `slot`.  Returns the next bucket to launch (in index order, now marked
// as handed out) if this event made one available, or -1.  Events beyond the expected count are ignored.
int b2f_reducer_mark(void* h, int slot) {
  Reducer* r = static_cast<Reducer*>(h);
  if (slot < 0 || slot >= static_cast<int>(r->uses.size()) || r->uses[slot] <= 0) return -1;
  if (--r->uses[slot] > 0) return -1;
  const int b = r->slot_bucket[slot];
  if (--r->pending[b] != 0 || b != r->next) return -1;
  return b2f_reducer_next_ready(h);
}

// End-of-backward sweep: the next bucket that has not been handed out yet (now marked), complete or
// not, in index order; -1 when all are out.
int b2f_reducer_next_unlaunched(void* h) {
  Reducer* r = static_cast<Reducer*>(h);
  const int nb = static_cast<int>(r->launched.size());
  if (r->next >= nb) return -1;
  const int b = r->next++;
  r->launched[b] = 1;
  return b;
}

int b2f_reducer_num_buckets(void* h) { return static_cast<int>(static_cast<Reducer*>(h)->launched.size()); }

}  // extern "C"
