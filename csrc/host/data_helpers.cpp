// b200flow — dataset index builders for the mmap'd GPT dataset (host C++).
//
// Equivalent in purpose to Megatron's pybind11 `helpers.cpp`
// (build_sample_idx / build_blending_indices; SURVEY.md §2E, §2H C31 — compiled at
// run time by `_compile_dependencies` in the reference's Megatron-DeepSpeed pin,
// containers/megatron-deepspeed/Dockerfile:13).  Written against a plain C ABI so
// it is callable through ctypes with numpy buffers.
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <random>
#include <vector>

extern "C" {

// Walk the (shuffled) document order and cut it into samples of seq_length+1 tokens.
// sample_idx has (num_samples + 1) rows of {index into doc_idx, token offset in that doc}.
// Returns the number of rows written, or -1 if the documents run out before num_samples.
long long b2f_build_sample_idx(const int32_t* sizes, const int32_t* doc_idx, long long doc_idx_len,
                               int32_t seq_length, long long num_samples, int64_t* sample_idx) {
  long long sample = 0, di = 0;
  int64_t offset = 0;
  sample_idx[0] = 0;
  sample_idx[1] = 0;
  ++sample;
  while (sample <= num_samples) {
    int64_t remaining = static_cast<int64_t>(seq_length) + 1;
    while (remaining != 0) {
      if (di >= doc_idx_len) return -1;
      const int64_t doc_len = static_cast<int64_t>(sizes[doc_idx[di]]) - offset;
      remaining -= doc_len;
      if (remaining <= 0) {
        // sample ends inside this document; the last token is shared with the next sample
        offset += remaining + doc_len - 1;
        remaining = 0;
      } else {
        ++di;
        offset = 0;
      }
    }
    sample_idx[2 * sample] = di;
    sample_idx[2 * sample + 1] = offset;
    ++sample;
  }
  return sample;
}

// Blend several datasets by weight: greedily pick, for every global sample, the
// dataset whose realised share lags its target share the most.
void b2f_build_blending_indices(uint8_t* dataset_index, int64_t* dataset_sample_index,
                                const double* weights, int32_t num_datasets, long long size) {
  std::vector<int64_t> taken(num_datasets, 0);
  for (long long i = 0; i < size; ++i) {
    const double denom = std::max<double>(static_cast<double>(i), 1.0);
    int best = 0;
    double best_err = weights[0] * denom - static_cast<double>(taken[0]);
    for (int d = 1; d < num_datasets; ++d) {
      const double err = weights[d] * denom - static_cast<double>(taken[d]);
      if (err > best_err) { best_err = err; best = d; }
    }
    dataset_index[i] = static_cast<uint8_t>(best);
    dataset_sample_index[i] = taken[best];
    ++taken[best];
  }
}

// Deterministic Fisher-Yates permutation of [0, n) (shuffle_idx).
void b2f_build_shuffle_idx(int64_t* out, long long n, uint64_t seed) {
  for (long long i = 0; i < n; ++i) out[i] = i;
  std::mt19937_64 rng(seed);
  for (long long i = n - 1; i > 0; --i) {
    const uint64_t j = rng() % static_cast<uint64_t>(i + 1);
    std::swap(out[i], out[j]);
  }
}

// doc_idx: `num_epochs` concatenated permutations of the document ids.
void b2f_build_doc_idx(int32_t* out, const int32_t* documents, long long num_docs,
                       int32_t num_epochs, uint64_t seed) {
  std::mt19937_64 rng(seed);
  for (int e = 0; e < num_epochs; ++e) {
    int32_t* dst = out + static_cast<long long>(e) * num_docs;
    std::memcpy(dst, documents, sizeof(int32_t) * num_docs);
    for (long long i = num_docs - 1; i > 0; --i) {
      const uint64_t j = rng() % static_cast<uint64_t>(i + 1);
      std::swap(dst[i], dst[j]);
    }
  }
}

}  // extern "C"
