// peng_oracle_select.cpp — TEST INFRASTRUCTURE ONLY.  See peng_oracle.h.
//
// Restatement of BasePattern::select_base_patterns (/root/reference/src/base_pattern.cpp:443-515).
// Kept in C++ because the selected list depends on the order libstdc++'s (unstable) std::sort leaves
// exactly-tied z-scores in — every reverse-complement pair ties — so the restatement has to run the
// same algorithm over the same index array with the same comparator (src/base_pattern.h:166-172).
#include <algorithm>
#include <cstdint>
#include <vector>

#include "peng_oracle.h"

namespace {
template <typename REAL>
struct by_z_desc {
  const REAL* z;
  bool operator()(const size_t i, const size_t j) const { return z[i] > z[j]; }
};

template <typename REAL>
uint64_t select_impl(const REAL* z, const uint64_t* counts, int W, REAL z_threshold, uint64_t count_threshold,
                     int single_stranded, int filter_neighbors, uint64_t* out, uint64_t cap) {
  const size_t n = (size_t)1 << (2 * W);
  std::vector<bool> seen(n, false);
  std::vector<size_t> order(n);
  for (size_t i = 0; i < n; i++) order[i] = i;
  std::sort(order.begin(), order.end(), by_z_desc<REAL>{z});
  uint64_t n_sel = 0;
  for (size_t r = 0; r < n; r++) {
    const size_t id = order[r];
    if (z[id] < z_threshold) break;                 // (:463)
    if (counts[id] < count_threshold) continue;     // (:466)
    bool take = !seen[id];
    if (!single_stranded) take = take && !seen[po_revcomp(id, W)];  // (:490-491)
    if (!take) continue;
    if (n_sel < cap) out[n_sel] = id;
    n_sel++;
    seen[id] = true;
    if (filter_neighbors) {  // all ids at Hamming distance <= 1 (:495-507)
      for (int p = 0; p < W; p++) {
        const size_t masked = id & ~((size_t)3 << (2 * p));
        for (size_t c = 0; c < 4; c++) seen[masked | (c << (2 * p))] = true;
      }
    }
  }
  return n_sel;
}
}  // namespace

extern "C" uint64_t po_select_f32(const float* z, const uint64_t* counts, int W, float zt, uint64_t ct, int ss,
                                  int fn, uint64_t* out, uint64_t cap) {
  return select_impl<float>(z, counts, W, zt, ct, ss, fn, out, cap);
}
extern "C" uint64_t po_select_f64(const double* z, const uint64_t* counts, int W, double zt, uint64_t ct, int ss,
                                  int fn, uint64_t* out, uint64_t cap) {
  return select_impl<double>(z, counts, W, zt, ct, ss, fn, out, cap);
}
