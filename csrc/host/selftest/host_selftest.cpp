// Self-test of the C++ host runtime (dataset index builders, bucket planner, shared-memory
// mailbox), meant to be built together with csrc/host/*.cpp under
//   g++ -std=c++17 -g -O1 -fsanitize=address,undefined -fno-omit-frame-pointer -pthread
// and run by tests/test_host_sanitizers.py.  The reference has no sanitizer target at all
// (SURVEY.md §5.2); this is the host-side half of ours (the device side is the cycle-budgeted
// spin-waits, the loopback tests and the delay-injection build).
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <thread>
#include <vector>

#include <unistd.h>

extern "C" {
long long b2f_build_sample_idx(const int32_t* sizes, const int32_t* doc_idx, long long doc_idx_len,
                               int32_t seq_length, long long num_samples, int64_t* sample_idx);
void b2f_build_blending_indices(uint8_t* dataset_index, int64_t* dataset_sample_index,
                                const double* weights, int32_t num_datasets, long long size);
void b2f_build_shuffle_idx(int64_t* out, long long n, uint64_t seed);
void b2f_build_doc_idx(int32_t* out, const int32_t* documents, long long num_docs, int32_t num_epochs,
                       uint64_t seed);
int b2f_plan_buckets(const int64_t* offsets, const int64_t* numels, int nslots, int64_t total_elems,
                     int64_t elem_bytes, int64_t cap_bytes, int64_t first_cap_bytes, int64_t* out,
                     int max_buckets);
void* b2f_reducer_create(const int64_t* plan, int nb, int nslots, const int32_t* uses);
void b2f_reducer_destroy(void* h);
void b2f_reducer_reset(void* h);
int b2f_reducer_mark(void* h, int slot);
int b2f_reducer_next_unlaunched(void* h);
int b2f_reducer_next_ready(void* h);
int b2f_reducer_num_buckets(void* h);
void* b2f_chan_create(const char* name, uint64_t capacity);
void* b2f_chan_open(const char* name);
int b2f_chan_send(void* ch, const void* buf, uint32_t len, long timeout_ms);
long long b2f_chan_recv(void* ch, void* buf, uint32_t buf_len, long timeout_ms, uint32_t* need_out);
void b2f_chan_close(void* ch);
void b2f_chan_free(void* ch);
}

static int failures = 0;
#define CHECK(cond)                                                        \
  do {                                                                     \
    if (!(cond)) {                                                         \
      std::fprintf(stderr, "FAIL %s:%d  %s\n", __FILE__, __LINE__, #cond); \
      ++failures;                                                          \
    }                                                                      \
  } while (0)

static void test_indices() {
  const int ndocs = 37, epochs = 3, seq = 16;
  std::vector<int32_t> sizes(ndocs), docs(ndocs), doc_idx(ndocs * epochs);
  long long tokens = 0;
  for (int i = 0; i < ndocs; ++i) { sizes[i] = 3 + (i * 7) % 29; docs[i] = i; tokens += sizes[i]; }
  b2f_build_doc_idx(doc_idx.data(), docs.data(), ndocs, epochs, 42);
  for (int e = 0; e < epochs; ++e) {
    std::vector<int> seen(ndocs, 0);
    for (int i = 0; i < ndocs; ++i) ++seen[doc_idx[e * ndocs + i]];
    for (int i = 0; i < ndocs; ++i) CHECK(seen[i] == 1);
  }
  const long long total = (epochs * tokens - 1) / seq;
  std::vector<int64_t> sample_idx((total + 1) * 2, -1);
  const long long got = b2f_build_sample_idx(sizes.data(), doc_idx.data(), (long long)doc_idx.size(), seq, total,
                                             sample_idx.data());
  CHECK(got == total + 1);
  CHECK(sample_idx[0] == 0 && sample_idx[1] == 0);
  for (long long s = 1; s <= total; ++s) {
    CHECK(sample_idx[2 * s] >= sample_idx[2 * (s - 1)]);
    CHECK(sample_idx[2 * s] < (int64_t)doc_idx.size());
    CHECK(sample_idx[2 * s + 1] >= 0 && sample_idx[2 * s + 1] < sizes[doc_idx[sample_idx[2 * s]]]);
  }
  // asking for more samples than the documents hold is reported, not overrun
  std::vector<int64_t> small(((total + 50) + 1) * 2, -1);
  CHECK(b2f_build_sample_idx(sizes.data(), doc_idx.data(), (long long)doc_idx.size(), seq, total + 50, small.data()) == -1);
  std::vector<int64_t> shuf(total);
  b2f_build_shuffle_idx(shuf.data(), total, 7);
  std::vector<int> seen(total, 0);
  for (auto v : shuf) { CHECK(v >= 0 && v < total); if (v >= 0 && v < total) ++seen[v]; }
  for (auto c : seen) CHECK(c == 1);
  // blending: realised shares follow the weights
  const double w[3] = {0.5, 0.3, 0.2};
  const long long n = 1000;
  std::vector<uint8_t> which(n);
  std::vector<int64_t> idx(n);
  b2f_build_blending_indices(which.data(), idx.data(), w, 3, n);
  long long cnt[3] = {0, 0, 0};
  for (long long i = 0; i < n; ++i) { CHECK(which[i] < 3); CHECK(idx[i] == cnt[which[i]]); ++cnt[which[i]]; }
  for (int d = 0; d < 3; ++d) CHECK(std::llabs(cnt[d] - (long long)(w[d] * n)) <= 2);
}

static void test_buckets() {
  const int n = 50;
  std::vector<int64_t> offs(n), nums(n);
  int64_t total = 0;
  for (int i = 0; i < n; ++i) { offs[i] = total; nums[i] = 1000 + 37 * i; total += nums[i]; }
  std::vector<int64_t> out((n + 1) * 4, -1);
  const int nb = b2f_plan_buckets(offs.data(), nums.data(), n, total, 4, 40000, 8000, out.data(), n + 1);
  CHECK(nb > 1);
  int next_hi = n;
  int64_t next_hi_elem = total;
  for (int b = 0; b < nb; ++b) {
    const int64_t s0 = out[4 * b], s1 = out[4 * b + 1], e0 = out[4 * b + 2], e1 = out[4 * b + 3];
    CHECK(s1 == next_hi && e1 == next_hi_elem);       // contiguous, last parameters first
    CHECK(s0 < s1 && e0 < e1 && e0 == offs[s0]);
    next_hi = (int)s0;
    next_hi_elem = e0;
  }
  CHECK(next_hi == 0 && next_hi_elem == 0);
  CHECK(b2f_plan_buckets(offs.data(), nums.data(), n, total, 4, 4000, 4000, out.data(), 2) == -1);  // too few rows
}

static void test_reducer() {
  // 3 buckets over 7 slots, backward order: {5,6} {2,3,4} {0,1}; slot 3 is used twice (tied weight)
  const int64_t plan[12] = {5, 7, 50, 70, 2, 5, 20, 50, 0, 2, 0, 20};
  const int32_t uses[7] = {1, 1, 1, 2, 1, 1, 1};
  void* r = b2f_reducer_create(plan, 3, 7, uses);
  CHECK(r != nullptr);
  if (!r) return;
  CHECK(b2f_reducer_num_buckets(r) == 3);
  for (int step = 0; step < 2; ++step) {
    CHECK(b2f_reducer_mark(r, 6) == -1);
    CHECK(b2f_reducer_mark(r, 5) == 0);          // bucket 0 complete
    CHECK(b2f_reducer_mark(r, 5) == -1);         // surplus event ignored
    CHECK(b2f_reducer_mark(r, 4) == -1);
    CHECK(b2f_reducer_mark(r, 3) == -1);         // first of two uses
    CHECK(b2f_reducer_mark(r, 2) == -1);
    CHECK(b2f_reducer_mark(r, 3) == 1);          // last use completes bucket 1
    CHECK(b2f_reducer_mark(r, 1) == -1);         // slot 0 never gets a gradient this step
    CHECK(b2f_reducer_next_unlaunched(r) == 2);  // the sweep hands out what is left, once
    CHECK(b2f_reducer_next_unlaunched(r) == -1);
    CHECK(b2f_reducer_mark(r, 0) == -1);         // a late event cannot hand bucket 2 out again
    b2f_reducer_reset(r);
  }
  // hooks firing out of bucket order (a rank whose head ran late): bucket 1 completes first but is only
  // handed out behind bucket 0, so the launch sequence is 0, 1, 2 on every rank
  CHECK(b2f_reducer_mark(r, 4) == -1);
  CHECK(b2f_reducer_mark(r, 3) == -1);
  CHECK(b2f_reducer_mark(r, 3) == -1);
  CHECK(b2f_reducer_mark(r, 2) == -1);           // bucket 1 complete, bucket 0 not: nothing leaves
  CHECK(b2f_reducer_next_ready(r) == -1);
  CHECK(b2f_reducer_mark(r, 6) == -1);
  CHECK(b2f_reducer_mark(r, 5) == 0);            // bucket 0 leaves ...
  CHECK(b2f_reducer_next_ready(r) == 1);         // ... and releases bucket 1 behind it
  CHECK(b2f_reducer_next_ready(r) == -1);
  CHECK(b2f_reducer_mark(r, 1) == -1);
  CHECK(b2f_reducer_mark(r, 0) == 2);
  CHECK(b2f_reducer_next_unlaunched(r) == -1);
  b2f_reducer_reset(r);
  CHECK(b2f_reducer_mark(r, 99) == -1 && b2f_reducer_mark(r, -1) == -1);
  b2f_reducer_destroy(r);
  const int64_t overlap[8] = {0, 3, 0, 30, 2, 4, 20, 40};   // slot 2 in two buckets
  CHECK(b2f_reducer_create(overlap, 2, 4, nullptr) == nullptr);
  const int64_t gap[4] = {1, 4, 10, 40};                    // slot 0 in no bucket
  CHECK(b2f_reducer_create(gap, 1, 4, nullptr) == nullptr);
}

static void test_mailbox() {
  const std::string name = "/b2f_selftest_" + std::to_string((long)getpid());
  void* rx = b2f_chan_create(name.c_str(), 1 << 16);
  CHECK(rx != nullptr);
  if (!rx) return;
  const int nmsg = 200;
  std::thread producer([&] {
    void* tx = b2f_chan_open(name.c_str());
    if (!tx) { ++failures; return; }
    std::vector<uint8_t> buf;
    for (int i = 0; i < nmsg; ++i) {
      buf.assign(1 + (i * 131) % 5000, (uint8_t)(i & 0xff));   // wraps the 64 KB ring many times
      if (b2f_chan_send(tx, buf.data(), (uint32_t)buf.size(), 10000) != 0) { ++failures; break; }
    }
    b2f_chan_close(tx);   // marks the stream finished for the reader ...
    b2f_chan_free(tx);    // ... and a non-owner's free only unmaps its view
  });
  std::vector<uint8_t> buf(64);   // deliberately small: exercises the "need a bigger buffer" path
  int received = 0;
  for (int spins = 0; spins < 100000 && received < nmsg; ++spins) {
    uint32_t need = 0;
    long long got = b2f_chan_recv(rx, buf.data(), (uint32_t)buf.size(), 50, &need);
    if (got == -4) { buf.resize(need); continue; }
    if (got == -2) continue;        // timeout
    if (got < 0) break;             // closed
    const size_t want = 1 + (received * 131) % 5000;
    CHECK((size_t)got == want);
    bool same = true;
    for (long long j = 0; j < got; ++j) same &= buf[j] == (uint8_t)(received & 0xff);
    CHECK(same);
    ++received;
  }
  producer.join();
  CHECK(received == nmsg);
  b2f_chan_free(rx);
}

int main() {
  test_indices();
  test_buckets();
  test_reducer();
  test_mailbox();
  if (failures) {
    std::fprintf(stderr, "host_selftest: %d failure(s)\n", failures);
    return 1;
  }
  std::printf("host_selftest: ok\n");
  return 0;
}
