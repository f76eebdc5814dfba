// host_ingest.h — host-side pieces of libpeng_cuda.so that are inherently serial or tiny:
//   * FASTA record splitting with the reference's line rules (the letters themselves are translated and
//     packed on the device),
//   * BackgroundModel::calculateV on the <= 1364 device-counted (k+1)-mer totals,
//   * select_base_patterns (must replay libstdc++'s std::sort tie order, see below),
//   * IUPAC id -> per-position letter-set mask.
// Reference citations are relative to the reference tree.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace peng_host {

// SequenceSet::readFASTA (src/shared/SequenceSet.cpp:285-447) as a single pass over the file image that hands the
// sequence bytes to a sink as it finds them (the sink of libpeng_cuda copies them into pinned staging buffers whose
// DMA to the device overlaps the scan).  Returns 0 ok, 2 malformed (message in *msg).
//   - lines end at '\n' only ('\r' stays in the line and later becomes a non-ACGT letter);
//   - `while (getline(file, line).good())`: a last line without '\n' is dropped (:304);
//   - blank lines are skipped (:306); '>' starts a record (:308); sequence lines are concatenated (:366);
//   - a record without sequence is ignored with a warning (:346, :417);
//   - a space inside a sequence line, or a sequence line before any header, is fatal (:361-373).
// Sink: void letters(const char* p, size_t n)  — next n sequence bytes of the current record
//       void end_record(const char* header, size_t header_len) — the current record is complete (it had >= 1 byte)
//       void end_of_file(bool last_record_kept) — whether the file's final record was a complete one
// warn != NULL: warnings are appended there instead of going to stderr (parallel scan: printed in file order afterwards)
template <typename Sink>
inline int scan_fasta(const char* text, size_t size, const char* path, Sink& sink, std::string* msg, std::string* warn = nullptr) {
  bool in_record = false;
  uint64_t record_bytes = 0;
  const char* header = nullptr;
  size_t header_len = 0;
  const char* p = text;
  const char* end = text + size;
  auto close_record = [&]() {
    if (record_bytes) sink.end_record(header, header_len);
    else if (warn) *warn += std::string("Warning: Ignore FASTA entry without sequence: ") + path + "\n";
    else fprintf(stderr, "Warning: Ignore FASTA entry without sequence: %s\n", path);
  };
  while (p < end) {
    const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
    if (!nl) break;  // unterminated last line: never delivered by getline().good()
    const size_t len = (size_t)(nl - p);
    if (len) {
      if (p[0] == '>') {
        if (in_record) close_record();
        in_record = true;
        record_bytes = 0;
        header = p + 1;
        header_len = len - 1;
      } else if (in_record) {
        if (memchr(p, ' ', len)) {
          *msg = std::string("Error: FASTA sequence contains space character: ") + path;
          return 2;
        }
        sink.letters(p, len);
        record_bytes += len;
      } else {
        *msg = std::string("Error: Wrong FASTA format: ") + path;
        return 2;
      }
    }
    p = nl + 1;
  }
  if (in_record) close_record();
  sink.end_of_file(in_record && record_bytes > 0);
  return 0;
}

// First position >= pos where a line starts with '>' (a record can only begin there), or size.
inline size_t next_header_line(const char* text, size_t size, size_t pos) {
  if (pos == 0) return 0;  // whatever the file starts with belongs to the first chunk
  const char* p = text + pos - 1;  // the line start at pos counts if the byte in front of it is a newline
  const char* end = text + size;
  while (p < end) {
    const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
    if (!nl || nl + 1 >= end) return size;
    if (nl[1] == '>') return (size_t)(nl + 1 - text);
    p = nl + 1;
  }
  return size;
}

struct RecordSink {
  const char* text = nullptr;
  size_t keep_lo = 0, keep_hi = (size_t)-1;  // shard: records whose first sequence line starts in [keep_lo, keep_hi)
  std::vector<const uint8_t*> ptrs;
  std::vector<uint32_t> lens;
  std::vector<std::unique_ptr<char[]>> arena;
  std::vector<std::pair<const char*, size_t>> segments, last_segments;
  const char* last_header = nullptr;
  size_t last_header_len = 0;
  bool cur_keep = true, last_mine = false, last_kept = false, too_long = false;
  void letters(const char* p, size_t n) {
    if (segments.empty()) cur_keep = (size_t)(p - text) >= keep_lo && (size_t)(p - text) < keep_hi;
    segments.emplace_back(p, n);
  }
  void end_record(const char* header, size_t header_len) {
    last_mine = cur_keep;
    if (cur_keep) {
      size_t total = 0;
      for (const auto& sg : segments) total += sg.second;
      if (total > 0xFFFFFFFFull) too_long = true;
      if (segments.size() == 1) {
        ptrs.push_back((const uint8_t*)segments[0].first);
      } else {
        arena.emplace_back(new char[total + 64]);
        char* w = arena.back().get();
        ptrs.push_back((const uint8_t*)w);
        for (const auto& sg : segments) {
          memcpy(w, sg.first, sg.second);
          w += sg.second;
        }
      }
      lens.push_back((uint32_t)total);
      last_header = header;
      last_header_len = header_len;
      last_segments.swap(segments);
    }
    segments.clear();
  }
  void end_of_file(bool last_record_kept) { last_kept = last_record_kept; }
};
struct ChunkedScan {
  const char* text;
  const char* path;
  std::vector<size_t> cut;  // chunk i = [cut[i], cut[i + 1])
  std::vector<RecordSink> sinks;
  std::vector<std::string> msgs, warns;
  std::vector<int> rcs;
  // n chunks cut at header lines near equal file positions; shard (part, n_parts) as in peng_cuda_ingest_fasta_shard
  void prepare(const char* text_, size_t size, const char* path_, int n, int part, int n_parts) {
    text = text_;
    path = path_;
    cut.assign(1, 0);
    for (int i = 1; i < n; i++) cut.push_back(std::max(cut.back(), next_header_line(text, size, size * (size_t)i / (size_t)n)));
    cut.push_back(size);
    sinks.clear();
    sinks.resize((size_t)n);
    msgs.assign((size_t)n, std::string());
    warns.assign((size_t)n, std::string());
    rcs.assign((size_t)n, 0);
    for (auto& k : sinks) {
      k.text = text;
      if (n_parts > 1) {
        k.keep_lo = (size_t)((unsigned __int128)size * (unsigned)part / (unsigned)n_parts);
        k.keep_hi = part == n_parts - 1 ? (size_t)-1 : (size_t)((unsigned __int128)size * (unsigned)(part + 1) / (unsigned)n_parts);
      }
    }
  }
  void scan_chunk(int i) {
    const size_t a = cut[(size_t)i], b = cut[(size_t)i + 1];
    rcs[(size_t)i] = b > a ? scan_fasta(text + a, b - a, path, sinks[(size_t)i], &msgs[(size_t)i], &warns[(size_t)i]) : 0;
  }
};


// The whole file as one read-only mapping (page cache pages, no copy).  An empty file maps to size 0.
struct MappedFile {
  const char* data = nullptr;
  size_t size = 0;
  int open_file(const char* path, bool populate = true);  // 0 ok, 1 cannot open; populate: map every page now (one serial pass)
  ~MappedFile();
};
inline int MappedFile::open_file(const char* path, bool populate) {
  const int fd = open(path, O_RDONLY);
  if (fd < 0) return 1;
  struct stat st;
  if (fstat(fd, &st) != 0 || !S_ISREG(st.st_mode)) {
    close(fd);
    return 1;
  }
  size = (size_t)st.st_size;
  if (size) {
    void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE | (populate ? MAP_POPULATE : 0), fd, 0);
    if (m == MAP_FAILED) {
      close(fd);
      size = 0;
      return 1;
    }
    if (populate) madvise(m, size, MADV_SEQUENTIAL);
    data = static_cast<const char*>(m);
  }
  close(fd);
  return 0;
}
inline MappedFile::~MappedFile() {
  if (data) munmap(const_cast<char*>(data), size);
}

// BackgroundModel::calculateV (src/shared/BackgroundModel.cpp:490-530) on counts laid out v[0]|v[1]|...|v[K].
template <typename R>
inline void calculate_v(const uint64_t* n_all, int K, const float* alpha, bool interpolate, R* v_all) {
  auto off = [](int k) { return (size_t)(((1u << (2 * (k + 1))) - 4u) / 3u); };
  const uint64_t* n0 = n_all;
  uint64_t base_counts = 0;
  for (int y = 0; y < 4; y++) base_counts += n0[y];
  R* v0 = v_all;
  for (int y = 0; y < 4; y++) v0[y] = ((R)n0[y] + (R)alpha[0] * 0.25f) / ((R)base_counts + (R)alpha[0]);
  for (int k = 1; k <= K; k++) {
    const uint64_t* nk = n_all + off(k);
    const uint64_t* nk1 = n_all + off(k - 1);
    R* vk = v_all + off(k);
    const R* vk1 = v_all + off(k - 1);
    const size_t Yk1 = (size_t)1 << (2 * (k + 1)), Yk = (size_t)1 << (2 * k);
    for (size_t y = 0; y < Yk1; y++) {
      const size_t y2 = y % Yk, yk = y / 4;
      if (interpolate) vk[y] = ((R)nk[y] + (R)alpha[k] * vk1[y2]) / ((R)nk1[yk] + (R)alpha[k]);
      else vk[y] = ((R)nk[y] + (R)alpha[k] * 0.25f) / ((R)nk1[yk] + (R)alpha[k]);
    }
    R factor = 0;
    for (size_t y = 0; y < Yk1; y++) {
      factor += vk[y];
      if ((y + 1) % 4 == 0) {
        for (int i = 0; i < 4; i++) vk[y - i] /= factor;
        factor = 0;
      }
    }
  }
}

inline uint64_t revcomp_id64(uint64_t id, int W) {
  uint64_t r = 0;
  for (int p = 0; p < W; p++) r |= (3 - ((id >> (2 * p)) & 3)) << (2 * (W - 1 - p));
  return r;
}

// BasePattern::select_base_patterns (src/base_pattern.cpp:443-515).  std::sort is unstable and every
// reverse-complement pair has bit-identical z-scores, so which member of a pair comes first — and is
// therefore reported — is decided by libstdc++'s introsort over the full index array; the only way to
// reproduce the reference's list is to run that same sort with the same comparator (src/base_pattern.h:166-172).
struct ByZDescending {
  const float* z;
  bool operator()(const size_t i, const size_t j) const { return z[i] > z[j]; }
};

// The walk below only looks at the sorted array until the first z below the threshold, i.e. at the K ids with
// z >= threshold (a few hundred of 4^W).  std::sort (libstdc++ introsort: median-of-3 partitions down to ranges of 16,
// then one insertion pass) never moves an element across a partition boundary once it is made, and the right part of
// a partition holds only values <= its pivot.  So a right part whose pivot is already below the threshold can be left
// unsorted without changing where, and in which tie order, the K leading ids end up: the same partition steps are
// made on every range that can still hold one of them, by calling libstdc++'s own building blocks.  This turns
// ~n log n comparator calls into ~2 n (104 ms -> 10 ms at W = 10, 2 s -> 0.2 s at W = 12) with the identical result
// (tests/test_kernel_logic_sim.py::test_pruned_sort_matches_std_sort).  Any NaN, other standard libraries and small
// tables take the plain std::sort.
#if defined(__GLIBCXX__)
template <typename Comp>
inline void pruned_introsort_loop(size_t* first, size_t* last, long depth_limit, Comp comp, const float* z, float threshold,
                                  size_t** cold_begin) {
  while (last - first > 16) {
    if (depth_limit == 0) {
      std::__partial_sort(first, last, last, comp);
      return;
    }
    --depth_limit;
    size_t* cut = std::__unguarded_partition_pivot(first, last, comp);  // pivot is left at *first
    if (z[*first] < threshold) {
      if (cut < *cold_begin) *cold_begin = cut;  // [cut, last): every z <= z[pivot] < threshold — never looked at
    } else {
      pruned_introsort_loop(cut, last, depth_limit, comp, z, threshold, cold_begin);
    }
    last = cut;
  }
}
#endif

// order[0 .. returned) is what std::sort(order, order + n, ByZDescending) would leave there, and every id behind it
// has z < threshold
inline size_t sort_leading_ids(std::vector<size_t>& order, const float* z, float threshold) {
  const size_t n = order.size();
  ByZDescending by_z{z};
#if defined(__GLIBCXX__)
  bool plain = n <= 4096 || threshold != threshold;
  for (size_t i = 0; i < n && !plain; i++) plain = z[i] != z[i];
  if (!plain) {
    auto comp = __gnu_cxx::__ops::__iter_comp_iter(by_z);
    size_t* first = order.data();
    size_t* cold_begin = first + n;
    pruned_introsort_loop(first, first + n, (long)std::__lg((long)n) * 2, comp, z, threshold, &cold_begin);
    // std::__final_insertion_sort restricted to the leading part
    const size_t lead = (size_t)(cold_begin - first);
    if (lead > 16) {
      std::__insertion_sort(first, first + 16, comp);
      std::__unguarded_insertion_sort(first + 16, first + lead, comp);
    } else {
      std::__insertion_sort(first, first + lead, comp);
    }
    return lead;
  }
#endif
  std::sort(order.begin(), order.end(), by_z);
  return n;
}

inline std::vector<uint64_t> select_base_patterns(const float* z, const uint64_t* counts, int W, float z_threshold,
                                                  uint64_t count_threshold, bool single_stranded, bool filter_neighbors) {
  const size_t n = (size_t)1 << (2 * W);
  std::vector<size_t> order(n);
  for (size_t i = 0; i < n; i++) order[i] = i;
  const size_t n_sorted = sort_leading_ids(order, z, z_threshold);
  std::vector<bool> seen(n, false);
  std::vector<uint64_t> selected;
  for (size_t r = 0; r < n_sorted; r++) {
    const size_t id = order[r];
    if (z[id] < z_threshold) break;
    if (counts[id] < count_threshold) continue;
    if (seen[id]) continue;
    if (!single_stranded && seen[revcomp_id64(id, W)]) continue;
    selected.push_back(id);
    seen[id] = true;
    if (filter_neighbors) {
      for (int p = 0; p < W; p++) {
        const size_t masked = id & ~((size_t)3 << (2 * p));
        for (size_t c = 0; c < 4; c++) seen[masked | (c << (2 * p))] = true;
      }
    }
  }
  return selected;
}

// ---- seed selection with the top partition levels on the device -----------------------------------------
// The 4^W-element introsort above spends nearly all of its time in the first few partition passes, which touch the whole
// index array only to throw the right part away (its pivot lies below the threshold).  One pass of libstdc++'s
//   std::__unguarded_partition(first + 1, last, pivot = *first)       [after std::__move_median_to_first]
// is data-parallel: with  A = positions (ascending) whose element stops the left scan  (!(z > z_pivot)),
//                         B = positions (descending) whose element stops the right scan (!(z_pivot > z)),
// the k-th swap exchanges A[k] and B[k] for every k < m, m = #{k : A[k] < B[k]} (the predicate is monotone), and the cut is
// min(A[m], B[m - 1]): while k < m the left pointer has only seen original elements in front of B[k] and the right
// pointer only original elements behind A[k], so the sequential scans stop exactly at the list entries.  The device runs
// these passes on its resident z table (flags -> scan -> compact -> pair swap, peng_cuda.cu: k_sel_*), follows the left
// part while it is long, and hands the host only the short leading range — as (z, id) pairs, so that the host needs no
// 4^W table at all — to finish with libstdc++'s own steps as above.  `Engine` is the device (peng_cuda.cu) or, for the
// GPU-less tests, a host restatement of the same list formulation (tests/sim/sim_kernels.cpp).
struct ZId {
  float z;
  uint32_t id;
};
struct ZIdDescending {
  bool operator()(const ZId& a, const ZId& b) const { return a.z > b.z; }
};

// the list formulation of one partition pass on order[0 .. n) (host restatement of k_sel_pivot / _flags / _compact /
// _find / _swap; n > 16 as in the introsort loop): returns the cut, *pivot_z = z of the pivot (left at order[0])
inline size_t partition_level_lists(uint32_t* order, const float* z, size_t n, float* pivot_z) {
  auto gt = [&](size_t i, size_t j) { return z[order[i]] > z[order[j]]; };
  const size_t a = 1, b = n / 2, c = n - 1;
  size_t sel;
  if (gt(a, b)) sel = gt(b, c) ? b : (gt(a, c) ? c : a);
  else sel = gt(a, c) ? a : (gt(b, c) ? c : b);
  std::swap(order[0], order[sel]);
  const float zp = z[order[0]];
  *pivot_z = zp;
  std::vector<uint32_t> A, B;
  for (size_t i = 1; i < n; i++)
    if (!(z[order[i]] > zp)) A.push_back((uint32_t)i);
  for (size_t i = n - 1; i >= 1; i--)
    if (!(zp > z[order[i]])) B.push_back((uint32_t)i);
  size_t lo = 0, hi = std::min(A.size(), B.size());  // m = number of k with A[k] < B[k]
  while (lo < hi) {
    const size_t mid = (lo + hi) / 2;
    if (A[mid] < B[mid]) lo = mid + 1; else hi = mid;
  }
  const size_t m = lo;
  for (size_t k = 0; k < m; k++) std::swap(order[A[k]], order[B[k]]);
  size_t cut = (size_t)-1;
  if (m < A.size()) cut = A[m];
  if (m >= 1 && B[m - 1] < cut) cut = B[m - 1];
  return cut;
}

#if defined(__GLIBCXX__)
template <typename Comp>
inline void pruned_introsort_loop_pairs(ZId* first, ZId* last, long depth_limit, Comp comp, float threshold, ZId** cold_begin) {
  while (last - first > 16) {
    if (depth_limit == 0) {
      std::__partial_sort(first, last, last, comp);
      return;
    }
    --depth_limit;
    ZId* cut = std::__unguarded_partition_pivot(first, last, comp);  // pivot is left at *first
    if (first->z < threshold) {
      if (cut < *cold_begin) *cold_begin = cut;
    } else {
      pruned_introsort_loop_pairs(cut, last, depth_limit, comp, threshold, cold_begin);
    }
    last = cut;
  }
}

// Engine: size_t level(size_t n_cur, float* pivot_z)   one partition pass on the leading n_cur positions -> cut
//         void   fetch(size_t n, ZId* out)            the leading n positions as (z, id) pairs
// -> lead[0 .. returned) is what std::sort over all n ids leaves in front, every id behind it has z < threshold.
// n > 4096, no NaN in the table (the caller checks).
template <typename Engine>
inline size_t sort_leading_by_levels(Engine& eng, size_t n, float threshold, size_t stop_below, std::vector<ZId>* lead) {
  struct Range { size_t b, e; long depth; };
  std::vector<Range> todo;  // right parts that may hold an id above the threshold: sorted on the host with the depth they had
  long depth = (long)std::__lg((long)n) * 2;
  size_t n_cur = n, cold = n;
  while (n_cur > stop_below && n_cur > 16 && depth > 0) {
    --depth;
    float pz;
    const size_t cut = eng.level(n_cur, &pz);
    if (pz < threshold) { if (cut < cold) cold = cut; }
    else todo.push_back(Range{cut, n_cur, depth});
    n_cur = cut;
  }
  lead->resize(cold);
  eng.fetch(cold, lead->data());
  auto comp = __gnu_cxx::__ops::__iter_comp_iter(ZIdDescending{});
  ZId* first = lead->data();
  ZId* cold_begin = first + cold;
  for (const Range& r : todo) pruned_introsort_loop_pairs(first + r.b, first + r.e, r.depth, comp, threshold, &cold_begin);
  pruned_introsort_loop_pairs(first, first + n_cur, depth, comp, threshold, &cold_begin);
  const size_t n_lead = (size_t)(cold_begin - first);
  if (n_lead > 16) {
    std::__insertion_sort(first, first + 16, comp);
    std::__unguarded_insertion_sort(first + 16, first + n_lead, comp);
  } else {
    std::__insertion_sort(first, first + n_lead, comp);
  }
  return n_lead;
}
#endif

// the walk of select_base_patterns over the sorted leading ids whose z >= threshold (src/base_pattern.cpp:470-515);
// counts[r] = count of sorted[r].id
inline std::vector<uint64_t> select_walk(const ZId* sorted, const uint64_t* counts, size_t n_candidates, int W,
                                         uint64_t count_threshold, bool single_stranded, bool filter_neighbors) {
  const size_t n = (size_t)1 << (2 * W);
  std::vector<bool> seen(n, false);
  std::vector<uint64_t> selected;
  for (size_t r = 0; r < n_candidates; r++) {
    const size_t id = sorted[r].id;
    if (counts[r] < count_threshold) continue;
    if (seen[id]) continue;
    if (!single_stranded && seen[revcomp_id64(id, W)]) continue;
    selected.push_back(id);
    seen[id] = true;
    if (filter_neighbors) {
      for (int p = 0; p < W; p++) {
        const size_t masked = id & ~((size_t)3 << (2 * p));
        for (size_t c = 0; c < 4; c++) seen[masked | (c << (2 * p))] = true;
      }
    }
  }
  return selected;
}

// IUPAC id (sum c_p 11^p, A,C,G,T,S,W,R,Y,M,K,N = 0..10; src/iupac_pattern.h:26-30) -> nibble p = set of
// bases the letter at position p stands for (representative_iupac_nucleotides, src/iupac_alphabet.cpp:138-180)
inline bool iupac_letter_mask(uint64_t iupac_id, int W, uint64_t* mask) {
  static const uint64_t sets[11] = {0x1, 0x2, 0x4, 0x8, 0x6 /*S=CG*/, 0x9 /*W=AT*/, 0x5 /*R=AG*/, 0xA /*Y=CT*/,
                                    0x3 /*M=AC*/, 0xC /*K=GT*/, 0xF /*N*/};
  uint64_t m = 0;
  for (int p = 0; p < W; p++) {
    m |= sets[iupac_id % 11] << (4 * p);
    iupac_id /= 11;
  }
  *mask = m;
  return iupac_id == 0;
}

}  // namespace peng_host
