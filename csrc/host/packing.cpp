// CPU-side packers for the data pipeline (C ABI, loaded with ctypes).
//
//  * llmt_best_fit_pack      best-fit bin packing: item i goes into the open bin with the least remaining
//                            space that still fits it (ties: lowest bin index), else opens a new bin.
//                            Same result as the reference's O(n * bins) scan
//                            (src/llm_training/data/pre_training/pre_training_datamodule.py:144-179) in
//                            O(n log bins) via an ordered set of (remaining, bin).
//  * llmt_group_by_length    greedy grouping of items sorted by length, one slot reserved per document
//                            (src/llm_training/data/instruction_tuning/instruction_tuning_datamodule.py:102-121).
#include <algorithm>
#include <cstdint>
#include <numeric>
#include <set>
#include <utility>
#include <vector>

extern "C" {

// lengths[n] (caller sorts if "decreasing" is wanted) -> bin_of[n]; returns number of bins.
int64_t llmt_best_fit_pack(const int64_t* lengths, int64_t n, int64_t capacity, int64_t* bin_of) {
  std::set<std::pair<int64_t, int64_t>> open;   // (remaining, bin index)
  std::vector<int64_t> remaining;
  for (int64_t i = 0; i < n; ++i) {
    const int64_t len = lengths[i];
    auto it = open.lower_bound({len, -1});
    if (it != open.end()) {
      const int64_t b = it->second;
      open.erase(it);
      remaining[b] -= len;
      open.insert({remaining[b], b});
      bin_of[i] = b;
    } else {
      const int64_t b = (int64_t)remaining.size();
      remaining.push_back(capacity - len);
      open.insert({capacity - len, b});
      bin_of[i] = b;
    }
  }
  return (int64_t)remaining.size();
}

// order[n] receives the indices sorted by length (stable); group_of[n] is indexed by ORIGINAL index.
int64_t llmt_group_by_length(const int64_t* lengths, int64_t n, int64_t max_length, int64_t* order,
                             int64_t* group_of) {
  std::iota(order, order + n, (int64_t)0);
  std::stable_sort(order, order + n, [&](int64_t a, int64_t b) { return lengths[a] < lengths[b]; });
  int64_t group = 0, sum = 0, count = 0;
  for (int64_t k = 0; k < n; ++k) {
    const int64_t i = order[k];
    if (count > 0 && sum + lengths[i] + count > max_length) {
      ++group;
      sum = 0;
      count = 0;
    }
    group_of[i] = group;
    sum += lengths[i];
    ++count;
  }
  return n == 0 ? 0 : group + 1;
}

// Concatenate variable-length int32 rows into a [rows, width] padded matrix (collator hot loop).
// offsets[rows + 1] index into `flat`; left_pad != 0 pads on the left.
void llmt_pad_rows_i32(const int32_t* flat, const int64_t* offsets, int64_t rows, int64_t width, int32_t pad_value,
                       int left_pad, int64_t* out) {
  for (int64_t r = 0; r < rows; ++r) {
    const int64_t len = offsets[r + 1] - offsets[r];
    int64_t* dst = out + r * width;
    const int64_t start = left_pad ? width - len : 0;
    for (int64_t c = 0; c < width; ++c) dst[c] = pad_value;
    for (int64_t c = 0; c < len; ++c) dst[start + c] = flat[offsets[r] + c];
  }
}

}  // extern "C"
