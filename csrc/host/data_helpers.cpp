// b200flow — dataset index builders for the mmap'd GPT dataset (host C++).
//
// Equivalent in purpose to Megatron's pybind11 `helpers.cpp`
// (build_sample_idx / build_blending_indices / build_mapping; SURVEY.md §2E, §2H C31 — compiled at
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

// Sentence-pair sample map for BERT-style pre-training over a corpus whose items are sentences
// (Megatron `build_mapping`).  A document d owns the sentences [docs[d], docs[d + 1]).  Each epoch
// walks the documents, packing consecutive sentences into one sample until the sample's target
// length is reached — the target is max_seq_length, or with probability short_seq_prob a uniform
// length in [2, max_seq_length] — while at least two sentences remain for the next sample; the
// sentences left at the end of a document close its last sample.  Documents with fewer than
// min_num_sent sentences, or with a sentence longer than `long_sentence_len` tokens, are skipped.
// Rows are {first sentence, last sentence + 1, target length}; they are shuffled at the end.
// Pass maps == NULL to count: the return value is the number of rows (at most max_num_samples),
// and a second call with a buffer of that many rows fills them identically.
long long b2f_build_bert_mapping(const int64_t* docs, long long num_docs, const int32_t* sizes,
                                 int32_t num_epochs, long long max_num_samples, int32_t max_seq_length,
                                 double short_seq_prob, uint64_t seed, int32_t min_num_sent,
                                 int32_t long_sentence_len, int64_t* maps) {
  const int32_t short_every = short_seq_prob > 0.0 ? static_cast<int32_t>(1.0 / short_seq_prob + 0.5) : 0;
  std::mt19937 rng(static_cast<uint32_t>(seed));
  auto next_target = [&]() -> int32_t {
    const uint32_t r = rng();
    if (short_every > 0 && r % static_cast<uint32_t>(short_every) == 0)
      return 2 + static_cast<int32_t>(rng() % static_cast<uint32_t>(max_seq_length - 1));
    return max_seq_length;
  };
  long long rows = 0;
  for (int32_t epoch = 0; epoch < num_epochs && rows < max_num_samples; ++epoch) {
    for (long long d = 0; d < num_docs && rows < max_num_samples; ++d) {
      const int64_t first = docs[d], last = docs[d + 1];
      if (last - first < min_num_sent) continue;
      bool too_long = false;
      for (int64_t s = first; s < last; ++s)
        if (sizes[s] > long_sentence_len) { too_long = true; break; }
      if (too_long) continue;
      int64_t start = first;
      int32_t tokens = 0, nsent = 0, target = next_target();
      for (int64_t s = first; s < last && rows < max_num_samples; ++s) {
        tokens += sizes[s];
        ++nsent;
        const int64_t remaining = last - s - 1;
        if ((tokens >= target && remaining > 1 && nsent >= min_num_sent) || remaining == 0) {
          if (maps) {
            maps[3 * rows + 0] = start;
            maps[3 * rows + 1] = s + 1;
            maps[3 * rows + 2] = target;
          }
          ++rows;
          start = s + 1;
          tokens = 0;
          nsent = 0;
          target = next_target();
        }
      }
    }
  }
  if (maps) {
    std::mt19937_64 shuf(seed + 1);
    for (long long i = rows - 1; i > 0; --i) {
      const uint64_t j = shuf() % static_cast<uint64_t>(i + 1);
      for (int c = 0; c < 3; ++c) std::swap(maps[3 * i + c], maps[3 * j + c]);
    }
  }
  return rows;
}

}  // extern "C"
