// CPU index builders for the dataset pipeline, exported with a C ABI and bound via
// ctypes + numpy buffers (no pybind11, builds with plain g++ against any python).
//
// Same job as the reference's pybind module megatron/data/helpers.cpp
// (build_sample_idx :83-169, build_blending_indices :20-80, build_mapping :187-451,
// build_blocks_mapping :453-694); the algorithms are restated from their behaviour.
#include <algorithm>
#include <cstdint>
#include <cmath>
#include <cstring>
#include <random>
#include <vector>

#define DATA_API extern "C" __attribute__((visibility("default")))

// GPT sample boundaries. A sample is seq_length+1 tokens and consecutive samples overlap
// by one token. out: int32 [num_samples+1, 2] = (index into doc_idx, token offset in
// that document). Returns the number of rows written.
DATA_API int64_t b200_build_sample_idx(const int32_t* sizes, const int32_t* doc_idx,
                                       int64_t doc_idx_len, int32_t seq_length,
                                       int64_t num_samples, int32_t* out) {
  int64_t di = 0;
  int32_t off = 0;
  out[0] = 0;
  out[1] = 0;
  int64_t row = 1;
  for (; row <= num_samples; ++row) {
    int64_t need = (int64_t)seq_length + 1;
    while (need > 0) {
      if (di >= doc_idx_len) return -1;  // ran out of tokens: caller mis-sized num_samples
      const int64_t avail = (int64_t)sizes[doc_idx[di]] - off;
      if (avail >= need) {
        // sample ends inside this document; keep the last token for the next sample
        off += (int32_t)(need - 1);
        need = 0;
      } else {
        need -= avail;
        ++di;
        off = 0;
      }
    }
    out[2 * row] = (int32_t)di;
    out[2 * row + 1] = off;
  }
  return row;
}

// Weighted mixture: sample i goes to the dataset whose realised share lags its target
// weight the most.
DATA_API void b200_build_blending_indices(uint8_t* dataset_index, int64_t* dataset_sample_index,
                                          const double* weights, int32_t num_datasets,
                                          int64_t size) {
  std::vector<int64_t> taken(num_datasets, 0);
  for (int64_t i = 0; i < size; ++i) {
    const double denom = std::max<double>((double)i, 1.0);
    int best = 0;
    double best_err = weights[0] * denom - (double)taken[0];
    for (int d = 1; d < num_datasets; ++d) {
      const double err = weights[d] * denom - (double)taken[d];
      if (err > best_err) {
        best_err = err;
        best = d;
      }
    }
    dataset_index[i] = (uint8_t)best;
    dataset_sample_index[i] = taken[best];
    ++taken[best];
  }
}

namespace {

inline int32_t target_len(int32_t short_seq_ratio, int32_t max_len, std::mt19937& gen) {
  // With probability 1/short_seq_ratio pick a random shorter target (BERT trick).
  if (short_seq_ratio == 0) return max_len;
  const uint32_t r = gen();
  if (r % short_seq_ratio == 0) return 2 + r % (max_len - 1);
  return max_len;
}

template <typename Emit>
int64_t walk_sentence_windows(const int64_t* docs, int64_t n_docs, const int32_t* sizes,
                              int32_t num_epochs, int64_t max_num_samples, int32_t max_seq_length,
                              double short_seq_prob, int32_t seed, int32_t min_num_sent,
                              Emit emit) {
  constexpr int32_t kLongSentence = 512;
  const int32_t ratio = short_seq_prob > 0 ? (int32_t)std::lround(1.0 / short_seq_prob) : 0;
  int64_t count = 0;
  std::mt19937 gen(seed);
  for (int32_t epoch = 0; epoch < num_epochs && count < max_num_samples; ++epoch) {
    for (int64_t doc = 0; doc < n_docs && count < max_num_samples; ++doc) {
      const int64_t first = docs[doc], last = docs[doc + 1];
      int64_t remain = last - first;
      if (remain < min_num_sent) continue;
      bool has_long = false;
      for (int64_t s = first; s < last; ++s)
        if (sizes[s] > kLongSentence) has_long = true;
      if (has_long) continue;
      int64_t start = first;
      int32_t seq_len = 0, num_sent = 0;
      int32_t target = target_len(ratio, max_seq_length, gen);
      for (int64_t s = first; s < last; ++s) {
        seq_len += sizes[s];
        ++num_sent;
        --remain;
        const bool window_full = seq_len >= target && remain >= min_num_sent &&
                                 num_sent >= min_num_sent;
        if (window_full || remain == 0) {
          if (num_sent >= min_num_sent) {
            emit(count, start, s + 1, target);
            ++count;
          }
          start = s + 1;
          seq_len = 0;
          num_sent = 0;
          target = target_len(ratio, max_seq_length, gen);
        }
      }
    }
  }
  return count;
}

}  // namespace

// BERT-style sentence windows: rows of (first sentence, one-past-last sentence, target
// length), shuffled with a 64-bit Mersenne twister seeded by seed+1. Call with
// out == nullptr to get the row count, then again with a buffer of 3*count entries.
DATA_API int64_t b200_build_mapping(const int64_t* docs, int64_t n_docs, const int32_t* sizes,
                                    int32_t num_epochs, int64_t max_num_samples,
                                    int32_t max_seq_length, double short_seq_prob, int32_t seed,
                                    int32_t min_num_sent, int64_t* out) {
  int64_t n = walk_sentence_windows(
      docs, n_docs, sizes, num_epochs, max_num_samples, max_seq_length, short_seq_prob, seed,
      min_num_sent, [&](int64_t i, int64_t a, int64_t b, int32_t t) {
        if (out) {
          out[3 * i] = a;
          out[3 * i + 1] = b;
          out[3 * i + 2] = t;
        }
      });
  if (out) {
    std::mt19937_64 rng(seed + 1);
    for (int64_t i = n - 1; i > 0; --i) {
      const int64_t j = (int64_t)(rng() % (uint64_t)(i + 1));
      for (int k = 0; k < 3; ++k) std::swap(out[3 * i + k], out[3 * j + k]);
    }
  }
  return n;
}

// ICT / REALM blocks: rows of (first sentence, one-past-last, document, block id).
DATA_API int64_t b200_build_blocks_mapping(const int64_t* docs, int64_t n_docs,
                                           const int32_t* sizes, const int32_t* title_sizes,
                                           int32_t num_epochs, int64_t max_num_samples,
                                           int32_t max_seq_length, int32_t seed,
                                           int32_t use_one_sent_blocks, int64_t* out) {
  constexpr int32_t kLongSentence = 512;
  const int32_t min_sent = use_one_sent_blocks ? 1 : 2;
  int64_t count = 0;
  for (int32_t epoch = 0; epoch < num_epochs && count < max_num_samples; ++epoch) {
    int64_t block_id = 0;
    for (int64_t doc = 0; doc < n_docs && count < max_num_samples; ++doc) {
      const int64_t first = docs[doc], last = docs[doc + 1];
      const int32_t budget = max_seq_length - title_sizes[doc];
      int64_t remain = last - first;
      if (remain < min_sent) continue;
      bool has_long = false;
      for (int64_t s = first; s < last; ++s)
        if (sizes[s] > kLongSentence) has_long = true;
      if (has_long) continue;
      int64_t start = first;
      int32_t seq_len = 0, num_sent = 0;
      for (int64_t s = first; s < last; ++s) {
        seq_len += sizes[s];
        ++num_sent;
        --remain;
        const bool full = seq_len >= budget && remain >= min_sent && num_sent >= min_sent;
        if (full || remain == 0) {
          if (num_sent >= min_sent) {
            if (out) {
              out[4 * count] = start;
              out[4 * count + 1] = s + 1;
              out[4 * count + 2] = doc;
              out[4 * count + 3] = block_id;
            }
            ++count;
            ++block_id;
          }
          start = s + 1;
          seq_len = 0;
          num_sent = 0;
        }
      }
    }
  }
  if (out) {
    std::mt19937_64 rng(seed + 1);
    for (int64_t i = count - 1; i > 0; --i) {
      const int64_t j = (int64_t)(rng() % (uint64_t)(i + 1));
      for (int k = 0; k < 4; ++k) std::swap(out[4 * i + k], out[4 * j + k]);
    }
  }
  return count;
}
