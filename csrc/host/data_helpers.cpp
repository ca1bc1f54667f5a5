// Dataset index builders (C ABI; numpy arrays are passed as raw pointers through ctypes).
//
// Native equivalent of Megatron's pybind11 `helpers.cpp` that the reference calls from
// gpt_dataset.py:433-437:
//   * build_sample_idx       - walk the shuffled document order and cut it into samples of
//                              seq_length + 1 tokens that may span documents
//   * build_blending_indices - greedy interleave of several datasets to match target weights
//   * build_mapping          - sentence-pair samples for BERT-style datasets: consecutive sentences of one document
//                              up to a (sometimes shortened) target length -> (first sentence, end sentence, target length)
//   * build_blocks_mapping   - retrieval blocks (ICT / REALM): consecutive sentences up to max_seq_length minus the title
//                              -> (first sentence, end sentence, document, block id)
// gpt_dataset.py:533-579 (`_build_sample_idx`) is the Python twin used as the oracle in tests; the mapping builders have
// Python twins in data/dataset_utils.py.
#include <algorithm>
#include <cstdint>
#include <cmath>
#include <random>
#include <vector>

namespace {
constexpr int32_t kLongSentenceLen = 512;      // documents containing a longer sentence are skipped (upstream constant)

inline int32_t target_sample_len(int32_t short_seq_ratio, int32_t max_length, std::mt19937& rng) {
  if (short_seq_ratio == 0) return max_length;
  const uint32_t r = rng();
  if (r % static_cast<uint32_t>(short_seq_ratio) == 0) return 2 + static_cast<int32_t>(r % static_cast<uint32_t>(max_length - 1));
  return max_length;
}

template <int W, typename T>
void shuffle_rows(T* rows, int64_t n, uint64_t seed) {      // Fisher-Yates with a 64-bit Mersenne twister
  std::mt19937_64 rng(seed);
  for (int64_t i = n - 1; i > 0; --i) {
    const int64_t j = static_cast<int64_t>(rng() % static_cast<uint64_t>(i + 1));
    for (int w = 0; w < W; ++w) std::swap(rows[W * i + w], rows[W * j + w]);
  }
}

inline bool has_long_sentence(const int32_t* sizes, int64_t first, int64_t last) {
  for (int64_t s = first; s < last; ++s) if (sizes[s] > kLongSentenceLen) return true;
  return false;
}
}  // namespace

extern "C" {

// sizes: tokens per document; doc_idx: shuffled document order over all epochs (length n_doc_idx).
// out: int32 [num_samples + 1, 2] = (index into doc_idx, token offset inside that document).
// returns number of rows written (num_samples + 1) or -1 if the token stream is too short.
int64_t b2h_build_sample_idx(const int32_t* sizes, const int32_t* doc_idx, int64_t n_doc_idx, int32_t seq_length,
                             int32_t num_epochs, int64_t tokens_per_epoch, int32_t* out, int64_t out_rows) {
  const int64_t num_samples = (static_cast<int64_t>(num_epochs) * tokens_per_epoch - 1) / seq_length;
  if (out_rows < num_samples + 1) return -1;
  int64_t sample = 0, di = 0;
  int32_t off = 0;
  out[0] = 0; out[1] = 0;
  ++sample;
  while (sample <= num_samples) {
    int32_t remaining = seq_length + 1;
    while (remaining != 0) {
      if (di >= n_doc_idx) return -1;
      const int32_t doc_len = sizes[doc_idx[di]] - off;
      remaining -= doc_len;
      if (remaining <= 0) {
        // sample ends inside this document; the next one re-uses its last token as its first input
        off += remaining + doc_len - 1;
        remaining = 0;
      } else {
        ++di;
        off = 0;
      }
    }
    out[2 * sample] = static_cast<int32_t>(di);
    out[2 * sample + 1] = off;
    ++sample;
  }
  return num_samples + 1;
}

// dataset_index[i] in [0, n), dataset_sample_index[i] = running per-dataset counter; at each step pick
// the dataset whose achieved share lags its weight the most.
void b2h_build_blending_indices(uint8_t* dataset_index, int64_t* dataset_sample_index, const double* weights,
                                int32_t num_datasets, int64_t size) {
  std::vector<int64_t> cur(num_datasets, 0);
  for (int64_t i = 0; i < size; ++i) {
    const double denom = static_cast<double>(i > 0 ? i : 1);
    int best = 0;
    double best_err = weights[0] * denom - static_cast<double>(cur[0]);
    for (int d = 1; d < num_datasets; ++d) {
      const double err = weights[d] * denom - static_cast<double>(cur[d]);
      if (err > best_err) { best_err = err; best = d; }
    }
    dataset_index[i] = static_cast<uint8_t>(best);
    dataset_sample_index[i] = cur[best];
    cur[best]++;
  }
}

// docs: [n_docs + 1] sentence-index boundaries of the documents; sizes: tokens per sentence.  out: int64 [rows, 3] =
// (first sentence, one-past-last sentence, target sequence length) or NULL to only count.  Returns the number of samples
// (capped by max_num_samples across epochs like upstream: the epoch loop stops once the cap is reached), -1 if out is
// too small.  Rows are shuffled with seed + 1 unless shuffle == 0.
int64_t b2h_build_mapping(const int64_t* docs, int64_t n_docs, const int32_t* sizes, int32_t num_epochs, int64_t max_num_samples,
                          int32_t max_seq_length, double short_seq_prob, int32_t seed, int32_t min_num_sent, int64_t* out,
                          int64_t out_rows, int32_t shuffle) {
  const int32_t short_seq_ratio = short_seq_prob > 0 ? static_cast<int32_t>(std::lround(1.0 / short_seq_prob)) : 0;
  std::mt19937 rng(static_cast<uint32_t>(seed));
  int64_t idx = 0;
  for (int32_t epoch = 0; epoch < num_epochs; ++epoch) {
    if (idx >= max_num_samples) break;
    for (int64_t d = 0; d < n_docs; ++d) {
      const int64_t first = docs[d], last = docs[d + 1];
      int64_t prev_start = first, remain = last - first;
      if (remain < std::max<int32_t>(min_num_sent, 2) || has_long_sentence(sizes, first, last)) continue;
      int32_t seq_len = 0, num_sent = 0;
      int32_t target = target_sample_len(short_seq_ratio, max_seq_length, rng);
      for (int64_t s = first; s < last; ++s) {
        seq_len += sizes[s]; ++num_sent; --remain;
        if ((seq_len >= target && remain > 1 && num_sent >= min_num_sent) || remain == 0) {
          if (out) {
            if (idx >= out_rows) return -1;
            out[3 * idx] = prev_start; out[3 * idx + 1] = s + 1; out[3 * idx + 2] = target;
          }
          ++idx;
          prev_start = s + 1;
          target = target_sample_len(short_seq_ratio, max_seq_length, rng);
          seq_len = 0; num_sent = 0;
        }
      }
    }
  }
  if (out && shuffle) shuffle_rows<3>(out, idx, static_cast<uint64_t>(seed) + 1);
  return idx;
}

// titles_sizes: tokens of each document's title (the block budget is max_seq_length - title).  out: int64 [rows, 4] =
// (first sentence, one-past-last sentence, document, block id within the epoch) or NULL to count.
int64_t b2h_build_blocks_mapping(const int64_t* docs, int64_t n_docs, const int32_t* sizes, const int32_t* titles_sizes,
                                 int32_t num_epochs, int64_t max_num_samples, int32_t max_seq_length, int32_t seed,
                                 int32_t use_one_sent_blocks, int64_t* out, int64_t out_rows, int32_t shuffle) {
  const int32_t min_num_sent = use_one_sent_blocks ? 1 : 2;
  int64_t idx = 0;
  for (int32_t epoch = 0; epoch < num_epochs; ++epoch) {
    int64_t block_id = 0;
    if (idx >= max_num_samples) break;
    for (int64_t d = 0; d < n_docs; ++d) {
      const int64_t first = docs[d], last = docs[d + 1];
      const int32_t target = max_seq_length - titles_sizes[d];
      int64_t prev_start = first, remain = last - first;
      if (remain < min_num_sent || has_long_sentence(sizes, first, last)) continue;
      int32_t seq_len = 0, num_sent = 0;
      for (int64_t s = first; s < last; ++s) {
        seq_len += sizes[s]; ++num_sent; --remain;
        if ((seq_len >= target && remain >= min_num_sent && num_sent >= min_num_sent) || remain == 0) {
          if (out) {
            if (idx >= out_rows) return -1;
            out[4 * idx] = prev_start; out[4 * idx + 1] = s + 1; out[4 * idx + 2] = d; out[4 * idx + 3] = block_id;
          }
          ++idx; ++block_id;
          prev_start = s + 1;
          seq_len = 0; num_sent = 0;
        }
      }
    }
  }
  if (out && shuffle) shuffle_rows<4>(out, idx, static_cast<uint64_t>(seed) + 1);
  return idx;
}

}  // extern "C"
