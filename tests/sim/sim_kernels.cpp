// sim_kernels.cpp — TEST INFRASTRUCTURE ONLY (never shipped, never loaded by the product).
//
// Compiles the per-thread kernel bodies of peng-motif_b200/csrc/core_seq.cuh and core_tab.cuh as plain
// host C++ and drives them "thread by thread" in a loop, reproducing the launch geometry of the
// __global__ wrappers in peng_cuda.cu.  This lets the kernel LOGIC (stream layout, k-mer start marking,
// exact non-overlap counting incl. warm-up / slow resolve, background counts, table formulas, IUPAC
// enumeration orders, EM chain order) be checked against the oracle on machines without a GPU.
// It is a debugging aid for `pytest -m "not gpu"`; parity proper is the `-m gpu` suite through the C ABI.
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

#define PENG_HD inline
#include "../../peng-motif_b200/csrc/core_seq.cuh"
#include "../../peng-motif_b200/csrc/core_tab.cuh"

using namespace peng;

struct EmitHost {
  uint64_t* hist;
  void operator()(uint32_t key, uint32_t) const { hist[key]++; }
};

template <int W>
static void count_W(const uint32_t* codes, const uint32_t* kstart, uint64_t stream_end, int both, uint32_t wpc, uint64_t* hist) {
  const uint64_t n_words16 = (stream_end - PENG_FRONT_PAD + 15) / 16 + 1;
  const uint64_t n_chunks = (n_words16 + wpc - 1) / wpc;
  for (uint64_t c = 0; c < n_chunks; c++) {
    const uint64_t e0 = PENG_FRONT_PAD + c * 16ull * wpc;
    if (both) count_chunk<W, true>(codes, kstart, e0, wpc, EmitHost{hist});
    else count_chunk<W, false>(codes, kstart, e0, wpc, EmitHost{hist});
  }
}

struct EmitKeys {
  std::vector<uint32_t>* keys;
  void operator()(uint32_t key, uint32_t) const { keys->push_back(key); }
};
template <int W>
static void count_keys_W(const uint32_t* codes, const uint32_t* kstart, uint64_t stream_end, int both, uint32_t wpc,
                         std::vector<uint32_t>* keys) {
  const uint64_t n_words16 = (stream_end - PENG_FRONT_PAD + 15) / 16 + 1;
  const uint64_t n_chunks = (n_words16 + wpc - 1) / wpc;
  for (uint64_t c = 0; c < n_chunks; c++) {
    const uint64_t e0 = PENG_FRONT_PAD + c * 16ull * wpc;
    if (both) count_chunk<W, true>(codes, kstart, e0, wpc, EmitKeys{keys});
    else count_chunk<W, false>(codes, kstart, e0, wpc, EmitKeys{keys});
  }
}

extern "C" {

// large W without a 4^W table: the counted keys in emission order (out may be NULL to size)
uint64_t sim_count_keys(const uint32_t* codes, const uint32_t* kstart, uint64_t stream_end, int W, int both, uint32_t wpc,
                        uint32_t* out, uint64_t cap) {
  std::vector<uint32_t> keys;
  switch (W) {
    case 12: count_keys_W<12>(codes, kstart, stream_end, both, wpc, &keys); break;
    case 14: count_keys_W<14>(codes, kstart, stream_end, both, wpc, &keys); break;
    case 16: count_keys_W<16>(codes, kstart, stream_end, both, wpc, &keys); break;
    default: return 0;
  }
  for (uint64_t i = 0; i < keys.size() && i < cap; i++) out[i] = keys[i];
  return keys.size();
}

uint64_t sim_n_words32(uint64_t n_seq, uint64_t total) {
  return (PENG_FRONT_PAD + total + 2 * n_seq + PENG_TAIL_PAD + 31) / 32 + 8;
}

// arrays must hold n_words32*2+8 (codes) / n_words32+8 (valid, sstart) words, zero-initialised by the caller
void sim_pack(const uint8_t* src, const uint64_t* offsets, uint64_t n_seq, int ascii, uint32_t* codes, uint32_t* valid,
              uint32_t* sstart) {
  const uint64_t nw = sim_n_words32(n_seq, n_seq ? offsets[n_seq] : 0);
  for (uint64_t t = 0; t < nw; t++) {
    if (ascii) pack_word<true>(t, src, offsets, n_seq, codes, valid, sstart);
    else pack_word<false>(t, src, offsets, n_seq, codes, valid, sstart);
  }
}

// nibble path of the hybrid ingest: the host's nibble_pack of the bases [src_base, total) (host_pack.cpp, the pool's threads)
// followed by pack_word<2> on every word — must leave what pack_word<0> leaves for the words that hold no base below src_base
void sim_pack_nibbles(const uint8_t* src, const uint64_t* offsets, uint64_t n_seq, uint64_t src_base, int n_threads, uint32_t* codes,
                      uint32_t* valid, uint32_t* sstart, uint64_t* first_word);

void sim_synth(uint64_t n_seq, uint32_t seq_len, uint64_t seed, uint64_t first_seq, const uint8_t* motif,
               uint32_t motif_len, double plant_fraction, double n_fraction, uint32_t* codes, uint32_t* valid,
               uint32_t* sstart) {
  SynthParams P;
  P.n_seq = n_seq; P.first_seq = first_seq; P.seed = seed; P.seq_len = seq_len; P.motif_len = motif ? motif_len : 0;
  P.plant_threshold = (uint32_t)(plant_fraction * 16777216.0);
  P.n_threshold = (uint32_t)(n_fraction * 16777216.0);
  P.motif_bits = 0; P.degenerate_pos = 0xFFFFFFFFu;
  for (uint32_t i = 0; i < P.motif_len; i++) {
    uint32_t c = motif[i];
    if (c > 3) { P.degenerate_pos = i; c = 1; }
    P.motif_bits |= (uint64_t)c << (2 * i);
  }
  const uint64_t nw = sim_n_words32(n_seq, n_seq * (uint64_t)seq_len);
  for (uint64_t t = 0; t < nw; t++) synth_word(t, P, codes, valid, sstart);
}

void sim_unpack(const uint32_t* codes, const uint32_t* valid, const uint64_t* offsets, uint64_t n_seq, uint8_t* out) {
  for (uint64_t s = 0; s < n_seq; s++)
    for (uint64_t i = 0; i < offsets[s + 1] - offsets[s]; i++)
      out[offsets[s] + i] = unpack_code(codes, valid, PENG_FRONT_PAD + offsets[s] + 2 * s + i);
}

uint64_t sim_mark(const uint32_t* valid, uint64_t n_words32, int W, uint32_t* kstart) {
  uint64_t ltot = 0;
  for (uint64_t t = 0; t < n_words32; t++) ltot += (uint64_t)mark_kstart_word(t, n_words32, valid, W, kstart);
  return ltot;
}

// hist: 4^W uint64, zeroed by the caller; unmirrored (canonical bins only for BOTH)
void sim_count(const uint32_t* codes, const uint32_t* kstart, uint64_t stream_end, int W, int both, uint32_t wpc, uint64_t* hist) {
  switch (W) {
    case 2: count_W<2>(codes, kstart, stream_end, both, wpc, hist); break;
    case 4: count_W<4>(codes, kstart, stream_end, both, wpc, hist); break;
    case 6: count_W<6>(codes, kstart, stream_end, both, wpc, hist); break;
    case 8: count_W<8>(codes, kstart, stream_end, both, wpc, hist); break;
    case 10: count_W<10>(codes, kstart, stream_end, both, wpc, hist); break;
    case 12: count_W<12>(codes, kstart, stream_end, both, wpc, hist); break;
    case 14: count_W<14>(codes, kstart, stream_end, both, wpc, hist); break;
    case 16: count_W<16>(codes, kstart, stream_end, both, wpc, hist); break;
  }
}

void sim_mirror(uint64_t* counts, int W) {
  const uint64_t n = 1ull << (2 * W);
  for (uint64_t i = 0; i < n; i++) {
    const uint32_t rc = revcomp_id((uint32_t)i, W);
    if ((uint32_t)i > rc) counts[i] = counts[rc];
  }
}

uint32_t sim_revcomp(uint32_t id, int W) { return revcomp_id(id, W); }

void sim_bg_counts(const uint32_t* codes, const uint32_t* valid, const uint32_t* sstart, uint64_t stream_end, int K,
                   uint64_t* out) {
  for (uint64_t q = PENG_FRONT_PAD; q < stream_end; q++) {
    const bool sep = q + 2 >= stream_end || bit_at(sstart, q + 1) || bit_at(sstart, q + 2);
    if (sep) continue;
    uint32_t ids[8];
    const uint32_t m = background_ids(codes, valid, sstart, q, K, ids);
    for (int k = 0; k <= K; k++)
      if ((m >> k) & 1u) out[v_offset(k) + ids[k]]++;
  }
}

void sim_bgprob_f32(const float* v, int W, int order, int both, float* out) {
  for (uint64_t i = 0; i < (1ull << (2 * W)); i++) out[i] = bg_prob_entry<float>(v, W, order, (uint32_t)i, both);
}
void sim_bgprob_f64(const double* v, int W, int order, int both, double* out) {
  for (uint64_t i = 0; i < (1ull << (2 * W)); i++) out[i] = bg_prob_entry<double>(v, W, order, (uint32_t)i, both);
}
void sim_stats_f32(const uint64_t* counts, const float* bg, uint64_t ltot, int W, float* e, float* z, float* lp) {
  for (uint64_t i = 0; i < (1ull << (2 * W)); i++) stats_of<float>(counts[i], bg[i], ltot, e + i, z + i, lp + i);
}
void sim_stats_f64(const uint64_t* counts, const double* bg, uint64_t ltot, int W, double* e, double* z, double* lp) {
  for (uint64_t i = 0; i < (1ull << (2 * W)); i++) stats_of<double>(counts[i], bg[i], ltot, e + i, z + i, lp + i);
}

// IUPAC: expansion list in the reference order (PLUS: dfs; BOTH: ascending distinct canonical via the
// representative rule + a sort done by the caller)
uint64_t sim_iupac_list(uint64_t mask, int W, int mode /*0 dfs, 1 any, 2 representatives only*/, uint32_t* out) {
  const IupacShape s = iupac_shape(mask, W);
  const uint64_t N = 1ull << s.total_bits;
  uint64_t n = 0;
  for (uint64_t t = 0; t < N; t++) {
    if (mode == 0) out[n++] = iupac_expansion_dfs(mask, W, s, t);
    else {
      const uint32_t e = iupac_expansion_any(mask, W, s, t);
      if (mode == 1 || iupac_is_representative(mask, W, e)) out[n++] = e;
    }
  }
  return n;
}

// EM float32: terms + chains exactly as k_em_terms_f32 / k_em_chains_f32 order them
void sim_em_acc_f32(const float* pwm, int W, const uint64_t* counts, const float* bg, float s, float* acc) {
  const uint64_t n = 1ull << (2 * W);
  std::vector<float> terms(n);
  for (uint64_t id = 0; id < n; id++) terms[id] = em_term_f32(pwm, W, (uint32_t)id, counts[id], bg[id], s);
  for (int p = 0; p < W; p++)
    for (uint32_t a = 0; a < 4; a++) {
      float sum = 0.f;
      for (uint32_t i = 0; i < (uint32_t)(n >> 2); i++) sum += terms[em_chain_id(p, a, i)];
      acc[p * 4 + a] = sum;
    }
}
}

// ---- partition path (k_count_part / k_count_slow / k_fold_canon of peng_cuda.cu), thread by thread ----------
static inline uint32_t kmer_at(const uint32_t* codes, uint64_t s, int W) {
  const uint64_t w = s >> 4;
  const uint32_t sh = 2u * (uint32_t)(s & 15);
  const uint64_t both = (uint64_t)codes[w] | ((uint64_t)codes[w + 1] << 32);
  return (uint32_t)(both >> sh) & (W == 16 ? 0xFFFFFFFFu : ((1u << (2 * W)) - 1u));
}

template <int W, bool BOTH>
static void count_part_T(const uint32_t* codes, const uint32_t* valid, uint64_t n_words32, uint64_t* hist, uint64_t* stats,
                         std::vector<uint32_t>* keys = nullptr) {
  std::vector<uint32_t> kstart(n_words32 + 8, 0);
  for (uint64_t u = 0; u < n_words32; u++) {
    kstart[u] = kstart_word(u, n_words32, valid, W);
    stats[0] += (uint64_t)popc32(kstart[u]);
  }
  for (uint64_t u = 0; u < n_words32; u++) {
    const uint32_t ks = kstart[u];
    if (!ks) continue;
    uint32_t cw[5];
    for (int j = 0; j < 5; j++) cw[j] = (2 * u + j >= 1) ? codes[2 * u - 1 + j] : 0u;
    const uint32_t cand = conflict_candidates<W, BOTH>(cw);
    const uint32_t fast = ks & ~cand, slow = ks & cand;
    // brute force: does an equal key start within the W-1 positions in front?
    for (uint32_t b = 0; b < 32; b++) {
      const uint64_t sp = u * 32 + b;
      const uint32_t k0 = kmer_at(codes, sp, W);
      const uint32_t c0 = BOTH ? std::min(k0, revcomp_id(k0, W)) : k0;
      bool truth = false;
      for (int d = 1; d < W && (uint64_t)d <= sp; d++) {
        const uint32_t k1 = kmer_at(codes, sp - d, W);
        const uint32_t c1 = BOTH ? std::min(k1, revcomp_id(k1, W)) : k1;
        truth |= (c0 == c1);
      }
      const bool flagged = (cand >> b) & 1u;
      if (truth && !flagged) stats[1]++;  // missed conflict: must stay 0
      if (!truth && flagged && ((ks >> b) & 1u)) stats[2]++;  // looked-at W-mers sent to the slow path needlessly
    }
    stats[3] += (uint64_t)popc32(fast);
    for (uint32_t b = 0; b < 32; b++)
      if ((fast >> b) & 1u) {
        const uint32_t k = kmer_at(codes, u * 32 + b, W);
        if (keys) keys->push_back(BOTH ? std::min(k, revcomp_id(k, W)) : k);  // sparse mode: canonical keys
        else hist[k]++;
      }
    if (slow) {
      stats[4]++;
      auto emit = [&](uint32_t key, uint32_t i) {
        const int32_t b = (int32_t)i - (W - 1);
        if (b >= 0 && b < 32 && ((slow >> b) & 1u)) {
          if (keys) keys->push_back(key);
          else hist[key]++;
        }
      };
      count_chunk<W, BOTH>(codes, kstart.data(), u * 32, 3, emit);
    }
  }
}

extern "C" {
// hist: 4^W uint64 zeroed by the caller; BOTH: forward-id bins for fast positions, canonical for slow ones (fold with
// sim_fold_canon).  stats[0] ltot, [1] missed conflicts (must be 0), [2] needless slow positions, [3] fast positions,
// [4] slow units
void sim_count_part(const uint32_t* codes, const uint32_t* valid, uint64_t n_words32, int W, int both, uint64_t* hist,
                    uint64_t* stats) {
  for (int i = 0; i < 5; i++) stats[i] = 0;
  switch (W) {
    case 4: both ? count_part_T<4, true>(codes, valid, n_words32, hist, stats) : count_part_T<4, false>(codes, valid, n_words32, hist, stats); break;
    case 6: both ? count_part_T<6, true>(codes, valid, n_words32, hist, stats) : count_part_T<6, false>(codes, valid, n_words32, hist, stats); break;
    case 8: both ? count_part_T<8, true>(codes, valid, n_words32, hist, stats) : count_part_T<8, false>(codes, valid, n_words32, hist, stats); break;
    case 10: both ? count_part_T<10, true>(codes, valid, n_words32, hist, stats) : count_part_T<10, false>(codes, valid, n_words32, hist, stats); break;
    case 12: both ? count_part_T<12, true>(codes, valid, n_words32, hist, stats) : count_part_T<12, false>(codes, valid, n_words32, hist, stats); break;
  }
}
// W in (12, 14, 16) without a 4^W table: the canonical keys of the counted occurrences (fast + slow), unordered
uint64_t sim_count_part_keys(const uint32_t* codes, const uint32_t* valid, uint64_t n_words32, int W, int both, uint32_t* out,
                             uint64_t cap, uint64_t* stats) {
  for (int i = 0; i < 5; i++) stats[i] = 0;
  std::vector<uint32_t> keys;
  switch (W) {
    case 12: both ? count_part_T<12, true>(codes, valid, n_words32, nullptr, stats, &keys) : count_part_T<12, false>(codes, valid, n_words32, nullptr, stats, &keys); break;
    case 14: both ? count_part_T<14, true>(codes, valid, n_words32, nullptr, stats, &keys) : count_part_T<14, false>(codes, valid, n_words32, nullptr, stats, &keys); break;
    case 16: both ? count_part_T<16, true>(codes, valid, n_words32, nullptr, stats, &keys) : count_part_T<16, false>(codes, valid, n_words32, nullptr, stats, &keys); break;
  }
  if (out) for (uint64_t i = 0; i < keys.size() && i < cap; i++) out[i] = keys[i];
  return keys.size();
}
}  // extern "C"
// bit-plane form of the conflict scan (conflict_candidates_planes) against the 2-bit form, unit by unit: number of units
// whose 32 flags differ (must be 0); *n_flagged = flags set in total (so that the test knows the scan saw conflicts)
template <int W, bool BOTH>
static uint64_t cand_planes_T(const uint32_t* codes, uint64_t n_words32, uint64_t* n_flagged) {
  std::vector<uint32_t> hi(n_words32 + 2, 0), lo(n_words32 + 2, 0);
  for (uint64_t u = 0; u < n_words32; u++) planes_of_unit(codes[2 * u], codes[2 * u + 1], &hi[u], &lo[u]);
  uint64_t bad = 0;
  for (uint64_t u = 0; u + 1 < n_words32; u++) {
    uint32_t cw[5];
    for (int j = 0; j < 5; j++) cw[j] = (2 * u + j >= 1) ? codes[2 * u - 1 + j] : 0u;
    const uint32_t pl[6] = {u ? hi[u - 1] : 0u, hi[u], hi[u + 1], u ? lo[u - 1] : 0u, lo[u], lo[u + 1]};
    const uint32_t a = conflict_candidates<W, BOTH>(cw), b = conflict_candidates_planes<W, BOTH>(pl);
    bad += a != b;
    *n_flagged += (uint64_t)popc32(b);
  }
  return bad;
}
extern "C" uint64_t sim_cand_planes_mismatch(const uint32_t* codes, uint64_t n_words32, int W, int both, uint64_t* n_flagged) {
  *n_flagged = 0;
  switch (W) {
    case 4: return both ? cand_planes_T<4, true>(codes, n_words32, n_flagged) : cand_planes_T<4, false>(codes, n_words32, n_flagged);
    case 6: return both ? cand_planes_T<6, true>(codes, n_words32, n_flagged) : cand_planes_T<6, false>(codes, n_words32, n_flagged);
    case 8: return both ? cand_planes_T<8, true>(codes, n_words32, n_flagged) : cand_planes_T<8, false>(codes, n_words32, n_flagged);
    case 10: return both ? cand_planes_T<10, true>(codes, n_words32, n_flagged) : cand_planes_T<10, false>(codes, n_words32, n_flagged);
    case 12: return both ? cand_planes_T<12, true>(codes, n_words32, n_flagged) : cand_planes_T<12, false>(codes, n_words32, n_flagged);
    case 14: return both ? cand_planes_T<14, true>(codes, n_words32, n_flagged) : cand_planes_T<14, false>(codes, n_words32, n_flagged);
    case 16: return both ? cand_planes_T<16, true>(codes, n_words32, n_flagged) : cand_planes_T<16, false>(codes, n_words32, n_flagged);
  }
  return ~0ull;
}
extern "C" {
// BOTH: pair sums into the canonical bin (then sim_mirror copies them to the other id)
void sim_fold_canon(uint64_t* hist, int W) {
  const uint64_t n = 1ull << (2 * W);
  for (uint64_t i = 0; i < n; i++) {
    const uint32_t rc = revcomp_id((uint32_t)i, W);
    if ((uint32_t)i < rc) { hist[i] += hist[rc]; hist[rc] = 0; }
  }
}
}

// ---- FASTA record splitting (host_ingest.h: scan_fasta + MappedFile), the host half of peng_cuda_ingest_fasta ----
#include "../../peng-motif_b200/csrc/host_ingest.h"
namespace {
struct VectorSink {
  std::vector<char> letters_;
  std::vector<uint64_t> offsets{0};
  bool last_kept = false;
  void letters(const char* p, size_t n) { letters_.insert(letters_.end(), p, p + n); }
  void end_record(const char*, size_t) { offsets.push_back(letters_.size()); }
  void end_of_file(bool k) { last_kept = k; }
};
}  // namespace
extern "C" {
// returns 0 ok, 1 cannot open, 2 malformed; letters (raw bytes) / offsets may be NULL to size
int sim_scan_fasta(const char* path, uint64_t* n_seq, uint64_t* n_letters, char* letters, uint64_t* offsets) {
  peng_host::MappedFile f;
  if (f.open_file(path) != 0) return 1;
  VectorSink sink;
  std::string msg;
  const int rc = peng_host::scan_fasta(f.data, f.size, path, sink, &msg);
  if (rc) return rc;
  *n_seq = sink.offsets.size() - 1;
  *n_letters = sink.letters_.size();
  if (letters) memcpy(letters, sink.letters_.data(), sink.letters_.size());
  if (offsets) memcpy(offsets, sink.offsets.data(), sink.offsets.size() * 8);
  return 0;
}
}

// the chunked scan of the parallel FASTA ingest (host_ingest.h: ChunkedScan, chunks scanned one after the other here):
// same return codes / outputs as sim_scan_fasta; *n_warn = warnings collected; shard (part, n_parts)
extern "C" int sim_scan_fasta_chunked(const char* path, int n_chunks, int part, int n_parts, uint64_t* n_seq, uint64_t* n_letters,
                                      char* letters, uint64_t* offsets, uint64_t* n_warn, int* last_kept_and_mine) {
  peng_host::MappedFile f;
  if (f.open_file(path, false) != 0) return 1;
  peng_host::ChunkedScan cs;
  cs.prepare(f.data, f.size, path, n_chunks, part, n_parts);
  for (int i = 0; i < n_chunks; i++) cs.scan_chunk(i);
  *n_warn = 0;
  int last_chunk = -1;
  for (int i = 0; i < n_chunks; i++) {
    for (char c : cs.warns[(size_t)i]) *n_warn += c == '\n';
    if (cs.rcs[(size_t)i]) return cs.rcs[(size_t)i];
    if (cs.cut[(size_t)i + 1] > cs.cut[(size_t)i]) last_chunk = i;
  }
  uint64_t ns = 0, nl = 0;
  for (auto& k : cs.sinks)
    for (size_t r = 0; r < k.ptrs.size(); r++) {
      if (letters) memcpy(letters + nl, k.ptrs[r], k.lens[r]);
      nl += k.lens[r];
      ns++;
      if (offsets) offsets[ns] = nl;
    }
  if (offsets) offsets[0] = 0;
  *n_seq = ns;
  *n_letters = nl;
  *last_kept_and_mine = last_chunk >= 0 && cs.sinks[(size_t)last_chunk].last_kept && cs.sinks[(size_t)last_chunk].last_mine;
  return 0;
}

// ---- seed selection (host_ingest.h: select_base_patterns with the pruned introsort) and the leading part of the sort ----
extern "C" {
uint64_t sim_select(const float* z, const uint64_t* counts, int W, float z_threshold, uint64_t count_threshold, int single_stranded,
                    int filter_neighbors, uint64_t* out, uint64_t cap) {
  std::vector<uint64_t> sel = peng_host::select_base_patterns(z, counts, W, z_threshold, count_threshold, single_stranded != 0,
                                                              filter_neighbors != 0);
  for (size_t i = 0; i < sel.size() && i < cap; i++) out[i] = sel[i];
  return sel.size();
}
// leading = ids the pruned sort guarantees, full = plain std::sort of all n ids; returns the guaranteed length
uint64_t sim_sort_leading(const float* z, uint64_t n, float threshold, uint64_t* leading, uint64_t* full) {
  std::vector<size_t> a(n), b(n);
  for (size_t i = 0; i < n; i++) a[i] = b[i] = i;
  const size_t lead = peng_host::sort_leading_ids(a, z, threshold);
  std::sort(b.begin(), b.end(), peng_host::ByZDescending{z});
  for (size_t i = 0; i < n; i++) { leading[i] = a[i]; full[i] = b[i]; }
  return lead;
}
}

// ---- seed selection through partition levels in the device's list formulation (host_ingest.h: sort_leading_by_levels with a
// host engine that runs partition_level_lists) ----
struct HostLevelEngine {
  std::vector<uint32_t> order;
  const float* z;
  size_t levels = 0;
  size_t level(size_t n_cur, float* pivot_z) {
    levels++;
    return peng_host::partition_level_lists(order.data(), z, n_cur, pivot_z);
  }
  void fetch(size_t n, peng_host::ZId* out) {
    for (size_t i = 0; i < n; i++) out[i] = peng_host::ZId{z[order[i]], order[i]};
  }
};
extern "C" {
// one pass: list formulation vs std::__unguarded_partition_pivot on the same array; returns 1 when cut and arrays agree
int sim_partition_level_agrees(const float* z, uint64_t n) {
  std::vector<uint32_t> a(n);
  std::vector<size_t> b(n);
  for (size_t i = 0; i < n; i++) a[i] = (uint32_t)i, b[i] = i;
  // a few nested levels on the shrinking left part, as the device runs them
  size_t na = n;
  while (na > 16) {
    float pz;
    const size_t cut_a = peng_host::partition_level_lists(a.data(), z, na, &pz);
    auto comp = __gnu_cxx::__ops::__iter_comp_iter(peng_host::ByZDescending{z});
    size_t* cut_b = std::__unguarded_partition_pivot(b.data(), b.data() + na, comp);
    if ((size_t)(cut_b - b.data()) != cut_a || pz != z[b[0]]) return 0;
    for (size_t i = 0; i < na; i++) if (a[i] != b[i]) return 0;
    na = cut_a;
  }
  return 1;
}
uint64_t sim_select_levels(const float* z, const uint64_t* counts, int W, float z_threshold, uint64_t count_threshold, int single_stranded,
                           int filter_neighbors, uint64_t stop_below, uint64_t* out, uint64_t cap, uint64_t* levels) {
  const size_t n = (size_t)1 << (2 * W);
  HostLevelEngine eng;
  eng.order.resize(n);
  for (size_t i = 0; i < n; i++) eng.order[i] = (uint32_t)i;
  eng.z = z;
  std::vector<peng_host::ZId> lead;
  const size_t n_lead = peng_host::sort_leading_by_levels(eng, n, z_threshold, stop_below, &lead);
  size_t k = 0;
  while (k < n_lead && !(lead[k].z < z_threshold)) k++;
  std::vector<uint64_t> c(k);
  for (size_t r = 0; r < k; r++) c[r] = counts[lead[r].id];
  std::vector<uint64_t> sel = peng_host::select_walk(lead.data(), c.data(), k, W, count_threshold, single_stranded != 0, filter_neighbors != 0);
  for (size_t i = 0; i < sel.size() && i < cap; i++) out[i] = sel[i];
  if (levels) *levels = eng.levels;
  return sel.size();
}
}

// ---- host half of the packed ingest (host_pack.cpp): all phases into one host image of the device arrays ----------------
#include "../../peng-motif_b200/csrc/host_pack.cpp"
extern "C" {
void sim_pack_nibbles(const uint8_t* src, const uint64_t* offsets, uint64_t n_seq, uint64_t src_base, int n_threads, uint32_t* codes,
                      uint32_t* valid, uint32_t* sstart, uint64_t* first_word) {
  const uint64_t total = n_seq ? offsets[n_seq] : 0;
  std::vector<uint8_t> nib((total - src_base + 1) / 2 + 64, 0xEE);
  peng_hostpack::Pool pool(n_threads);
  peng_hostpack::nibble_pack(pool, src + src_base, total - src_base, nib.data());
  // first word whose positions all lie at or behind the first base >= src_base
  uint64_t s = 0;
  while (s < n_seq && offsets[s] < src_base) s++;
  const uint64_t q = PENG_FRONT_PAD + (s < n_seq ? offsets[s] : total) + 2 * s;
  *first_word = (q + 31) / 32;
  const uint64_t nw = sim_n_words32(n_seq, total);
  for (uint64_t t = *first_word; t < nw; t++) pack_word<2>(t, nib.data(), offsets, n_seq, codes, valid, sstart, nullptr, src_base);
}

// codes32: 2 * n_words32 + 8 words, valid32: n_words32 + 8 words (as the device allocates them); returns the number of
// phases, or -1 when a sequence does not fit a phase
int sim_host_pack(const uint8_t* src, const uint64_t* offsets, uint64_t n_seq, int ascii, int n_threads, uint64_t max_blocks,
                  uint64_t n_words32, uint32_t* codes32, uint32_t* valid32, uint64_t* n_exc) {
  using namespace peng_hostpack;
  Pool pool(n_threads);
  if (n_exc) *n_exc = 0;
  const uint64_t q_total = ((n_words32 * 32 + 63) / 64) * 64;
  const uint64_t n_blocks = q_total / 64;
  std::vector<uint64_t> sc(2 * max_blocks + 2), sv(max_blocks + 1);
  bool has_carry = false;
  Partial carry{};
  int phases = 0;
  uint64_t s = 0;
  bool first = true;
  for (;;) {
    uint64_t e = plan_phase(offsets, n_seq, s, PENG_FRONT_PAD, first, max_blocks);
    bool last = e == n_seq;
    if (last) {  // the tail pad must fit as well
      const uint64_t q0 = first ? 0 : PENG_FRONT_PAD + offsets[s] + 2 * s;
      if (n_blocks - (q0 >> 6) > max_blocks) last = false;
    }
    if (e == s && !last) return -1;
    Phase ph{src, offsets, s, e, ascii != 0, PENG_FRONT_PAD, first, last, q_total, sc.data(), sv.data()};
    std::vector<ValidExc> exc;
    bool exc_complete = false;
    ph.exc = &exc;
    ph.exc_complete = &exc_complete;
    uint64_t blk0, blk1;
    phase_blocks(ph, &blk0, &blk1);
    pack_phase(pool, ph, &has_carry, &carry);
    memcpy(codes32 + blk0 * 4, sc.data(), (blk1 - blk0) * 16);
    if (exc_complete) {  // what the device does with a phase that sends no validity words: implied ones, then the listed blocks
      for (uint64_t t = blk0 * 2; t < blk1 * 2; t++) valid32[t] = implied_valid_word(t, offsets, n_seq);
      for (const ValidExc& x : exc) {
        if (x.blk < blk0 || x.blk >= blk1) return -3;
        memcpy(valid32 + x.blk * 2, &x.v, 8);
      }
      if (n_exc) *n_exc += exc.size();
      if (memcmp(valid32 + blk0 * 2, sv.data(), (blk1 - blk0) * 8)) return -4;
    } else {
      memcpy(valid32 + blk0 * 2, sv.data(), (blk1 - blk0) * 8);
    }
    phases++;
    if (last) break;
    s = e;
    first = false;
  }
  return has_carry ? -2 : phases;
}

// Several callers on ONE pool at the same time (peng_motif --gpus N: one host thread per device, all of them ingest
// through the process-wide pool), each with two phases in flight the way ingest_host_hybrid runs them (pack_phase_begin of
// phase i + 1 before phase i is taken out of its staging buffer).  Every caller packs the whole set `rounds` times and
// compares its codes with want_codes32 (2 * n_words32 words).  Returns the number of callers that saw a difference or
// an error.
int sim_host_pack_concurrent(const uint8_t* src, const uint64_t* offsets, uint64_t n_seq, int n_callers, int n_threads, int rounds,
                             uint64_t max_blocks, uint64_t n_words32, const uint32_t* want_codes32) {
  using namespace peng_hostpack;
  Pool pool(n_threads);
  const uint64_t q_total = ((n_words32 * 32 + 63) / 64) * 64;
  const uint64_t n_blocks = q_total / 64;
  std::vector<int> bad((size_t)n_callers, 0);
  auto caller = [&](int who) {
    std::vector<uint32_t> codes32(2 * n_words32 + 8, 0);
    struct Slot {
      std::vector<uint64_t> sc, sv;
      Phase ph;
      PhaseJob* job = nullptr;
      uint64_t blk0 = 0, blk1 = 0, e = 0;
      bool last = false;
    } slots[2];
    for (auto& c : slots) { c.sc.resize(2 * max_blocks + 2); c.sv.resize(max_blocks + 1); }
    for (int r = 0; r < rounds && !bad[(size_t)who]; r++) {
      std::fill(codes32.begin(), codes32.end(), 0u);
      bool has_carry = false;
      Partial carry{};
      auto begin = [&](Slot& c, uint64_t s, bool first) -> bool {
        uint64_t e = plan_phase(offsets, n_seq, s, PENG_FRONT_PAD, first, max_blocks);
        bool last = e == n_seq;
        if (last) {
          const uint64_t q0 = first ? 0 : PENG_FRONT_PAD + offsets[s] + 2 * s;
          if (n_blocks - (q0 >> 6) > max_blocks) last = false;
        }
        if (e == s && !last) return false;
        c.ph = Phase{src, offsets, s, e, false, PENG_FRONT_PAD, first, last, q_total, c.sc.data(), c.sv.data()};
        c.e = e;
        c.last = last;
        phase_blocks(c.ph, &c.blk0, &c.blk1);
        c.job = pack_phase_begin(pool, c.ph);
        return true;
      };
      if (!begin(slots[0], 0, true)) { bad[(size_t)who] = 1; break; }
      for (int cur = 0;;) {
        Slot& c = slots[cur];
        pack_phase_end(pool, c.job, &has_carry, &carry);
        c.job = nullptr;
        const bool more = !c.last;
        if (more && !begin(slots[cur ^ 1], c.e, false)) { bad[(size_t)who] = 1; break; }
        memcpy(codes32.data() + c.blk0 * 4, c.sc.data(), (c.blk1 - c.blk0) * 16);  // "issue" of phase i beside the packing of i + 1
        if (!more) break;
        cur ^= 1;
      }
      for (auto& c : slots)
        if (c.job) {
          bool hc = false;
          Partial pc{};
          pack_phase_end(pool, c.job, &hc, &pc);
          c.job = nullptr;
        }
      if (has_carry || memcmp(codes32.data(), want_codes32, 2 * n_words32 * 4)) bad[(size_t)who] = 1;
    }
  };
  std::vector<std::thread> ts;
  for (int i = 0; i < n_callers; i++) ts.emplace_back(caller, i);
  for (auto& t : ts) t.join();
  int n_bad = 0;
  for (int b : bad) n_bad += b;
  return n_bad;
}
}
