// Dataset index builders (C ABI; numpy arrays are passed as raw pointers through ctypes).
//
// Native equivalent of Megatron's pybind11 `helpers.cpp` that the reference calls from
// gpt_dataset.py:433-437:
//   * build_sample_idx       - walk the shuffled document order and cut it into samples of
//                              seq_length + 1 tokens that may span documents
//   * build_blending_indices - greedy interleave of several datasets to match target weights
// gpt_dataset.py:533-579 (`_build_sample_idx`) is the Python twin used as the oracle in tests.
#include <cstdint>
#include <cmath>
#include <vector>

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

}  // extern "C"
