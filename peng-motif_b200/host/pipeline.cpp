// pipeline.cpp — src/main.cpp + Peng::process (src/peng.cpp:315-435) on top of the C ABI.
//
// Device work per phase (all through include/peng_cuda.h):
//   ingest            peng_cuda_ingest_fasta                      SequenceSet::readFASTA
//   background model  peng_cuda_background_model                  BackgroundModel ctor
//   base patterns     peng_cuda_count + peng_cuda_stats           BasePattern ctor
//   seeds             peng_cuda_select_base_patterns              BasePattern::select_base_patterns
//   hill climb        peng_cuda_iupac_aggregate, one call per sweep (<= W x 10 candidates of one mother pattern)
//   PWMs              peng_cuda_iupac_count, one call for all motifs (4 W substitutions each)
//   EM                peng_cuda_em, all PWMs in one batch
// Host work: the bookkeeping between those calls and everything after EM (merging, redundancy filter, writers).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <iomanip>
#include <iostream>
#include <memory>
#include <set>
#include <thread>
#include <unordered_map>
#include <unordered_set>

#include "peng_host.h"

namespace peng_host {
namespace {

struct Device {  // owns one handle; any failing call ends the run with the library's message, like the reference's exit()
  peng_handle_t h = nullptr;
  explicit Device(int device) {
    if (peng_cuda_create(device, &h) != PENG_OK) {
      std::cerr << "Error: " << peng_cuda_last_error(nullptr) << std::endl;
      exit(1);
    }
  }
  ~Device() { peng_cuda_destroy(h); }
  void check(int rc, int exit_code = 1) const {
    if (rc == PENG_OK) return;
    std::cerr << peng_cuda_last_error(h) << std::endl;
    exit(exit_code);
  }
};

struct Stopwatch {
  bool on;
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  void lap(const char* what) {
    if (!on) return;
    const auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[timing] %-28s %10.3f ms\n", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};

void print_status(const std::string& header, bool leading_newline = true) {  // src/peng.cpp:315-320
  if (leading_newline) std::cout << std::endl;
  std::cout << "[STATUS] " << header << ":" << std::endl;
}

void print_table_header() {
  std::cout << std::setw(15) << "pattern" << "\t" << std::setw(15) << "observed" << "\t" << std::setw(15) << "enrichment"
            << "\t" << std::setw(15) << "zscore" << std::endl
            << std::endl;
}

void print_climb_line(const Motif& m, float score) {  // src/peng.cpp:456-463,489-496
  std::cout << "\t" << std::setw(15) << iupac_to_string(m.pattern, m.length) << "\t" << std::setw(10) << m.n_sites << "\t"
            << std::setw(5) << std::setprecision(2) << (float)m.n_sites / m.expected << "\t" << std::setw(10)
            << std::setprecision(6) << score << std::endl;
}

struct BaseTables {  // the columns of the 4^W tables the host reads for the selected seeds only
  std::vector<uint64_t> counts;
  std::vector<float> expected, zscore, logp;
};

}  // namespace

namespace {

// Peng::optimize_iupac_patterns (src/peng.cpp:437-541).  One sweep = every single-letter substitution of the
// current mother pattern by a similar IUPAC letter; the candidates of a sweep are independent of each other, so
// their degenerate sums come back from ONE device call and the host replays the reference's sequential
// better-than-best scan over them (which also decides what is printed).
void optimize_iupac_patterns(const Device& dev, const Options& opt, int W, const std::vector<uint64_t>& seeds,
                             const BaseTables& seed_cols, uint64_t n_sequences, std::vector<std::unique_ptr<Motif>>& out) {
  std::set<uint64_t> seen, best;
  const size_t pseudo_expected = (size_t)((float)n_sequences * opt.enrich_pseudocount_factor);
  const unsigned n_seq_u = (unsigned)n_sequences;

  // Degenerate sums are pure functions of the pattern, so they are cached per IUPAC id: `fetch` asks the device for the ids
  // it has not seen yet (one call, chunked only to bound the kernel's scratch), `aggregate` reads the cache.
  struct Agg { uint64_t sites; float bgp, expected; };
  std::unordered_map<uint64_t, Agg> cache;
  auto fetch = [&](const std::vector<uint64_t>& ids) {
    std::vector<uint64_t> need;
    {
      std::unordered_set<uint64_t> asked;
      for (uint64_t id : ids)
        if (!cache.count(id) && asked.insert(id).second) need.push_back(id);
    }
    const size_t max_call = W <= 10 ? 8192 : W == 12 ? 1024 : 64;  // BOTH: 4^W bits of scratch per pattern
    for (size_t at = 0; at < need.size(); at += max_call) {
      const int n = (int)std::min(max_call, need.size() - at);
      std::vector<uint64_t> sites((size_t)n);
      std::vector<float> bgp((size_t)n), expected((size_t)n);
      dev.check(peng_cuda_iupac_aggregate(dev.h, need.data() + at, n, PENG_F32REF, sites.data(), bgp.data(), expected.data()));
      for (int i = 0; i < n; i++) cache[need[at + (size_t)i]] = Agg{sites[(size_t)i], bgp[(size_t)i], expected[(size_t)i]};
    }
  };
  auto aggregate = [&](const std::vector<uint64_t>& ids, std::vector<Motif>& motifs) {
    fetch(ids);
    const int n = (int)ids.size();
    motifs.assign((size_t)n, Motif());
    for (int i = 0; i < n; i++) {
      const Agg& a = cache[ids[(size_t)i]];
      motifs[(size_t)i].pattern = ids[(size_t)i];
      motifs[(size_t)i].length = W;
      set_aggregated_attributes(motifs[(size_t)i], a.sites, a.bgp, a.expected);
    }
  };
  auto seed_score = [&](size_t si) -> float {  // BasePattern::getOptimizationScore (src/base_pattern.cpp:180-224)
    switch (opt.score) {
      case Score::LogPval: return seed_cols.logp[si];
      case Score::ExpCounts: return (seed_cols.expected[si] + (float)pseudo_expected) / (float)seed_cols.counts[si];
      default: return mutual_information_score((float)(unsigned)seed_cols.counts[si], seed_cols.expected[si], n_seq_u);
    }
  };
  auto sweep_candidates = [&](uint64_t mother, std::vector<uint64_t>& candidates) {
    for (int p = 0; p < W; p++) {
      const int c = iupac_letter_at(mother, p);
      const uint64_t masked = mother - (uint64_t)c * iupac_factor(p);
      for (int r : similar_letters(c)) candidates.push_back(masked + (uint64_t)r * iupac_factor(p));
    }
  };

  // Speculative pass: the climbs of ALL seeds advance in lockstep, one device call per round for the sweeps of every climb
  // that is still improving (the reference climbs seed after seed, <= W x 10 patterns per call and a host round trip
  // each).  A climb here ignores the `seen` set, which can only cut a climb short, so the sequential replay below walks a
  // prefix of what was computed here and finds every sum in the cache; nothing is printed in this pass.
  {
    struct Climb { uint64_t pattern; float score; bool active; };
    std::vector<Climb> climbs;
    std::vector<uint64_t> ids;
    for (size_t si = 0; si < seeds.size(); si++) {
      climbs.push_back(Climb{base_id_to_iupac_id(seeds[si], W), seed_score(si), true});
      ids.push_back(climbs.back().pattern);
    }
    fetch(ids);
    for (int round = 0; round < 64; round++) {
      ids.clear();
      for (const Climb& c : climbs)
        if (c.active) sweep_candidates(c.pattern, ids);
      if (ids.empty()) break;
      fetch(ids);
      for (Climb& c : climbs) {
        if (!c.active) continue;
        std::vector<uint64_t> cand;
        sweep_candidates(c.pattern, cand);
        std::vector<Motif> mutants;
        aggregate(cand, mutants);
        c.active = false;
        for (const Motif& mutant : mutants) {
          const float score = motif_score(mutant, opt.score, pseudo_expected, n_seq_u);
          if (score < c.score) {
            c.active = true;
            c.score = score;
            c.pattern = mutant.pattern;
          }
        }
      }
    }
  }

  for (size_t si = 0; si < seeds.size(); si++) {
    const uint64_t seed = seeds[si];
    std::vector<Motif> one;
    aggregate({base_id_to_iupac_id(seed, W)}, one);
    Motif best_mutant = one[0];
    float best_score = seed_score(si);  // from the BASE tables
    print_climb_line(best_mutant, best_score);

    bool found_better = true;
    while (found_better) {
      found_better = false;
      const uint64_t mother = best_mutant.pattern;
      std::set<uint64_t> current_seen;
      std::vector<uint64_t> candidates;
      sweep_candidates(mother, candidates);
      std::vector<Motif> mutants;
      aggregate(candidates, mutants);
      for (const Motif& mutant : mutants) {
        current_seen.insert(mutant.pattern);
        const float score = motif_score(mutant, opt.score, pseudo_expected, n_seq_u);
        if (score < best_score) {
          found_better = true;
          best_score = score;
          best_mutant = mutant;
          print_climb_line(best_mutant, best_score);
        }
      }
      if (seen.count(best_mutant.pattern) == 1) found_better = false;
      current_seen.erase(best_mutant.pattern);
      seen.insert(current_seen.begin(), current_seen.end());
    }

    if (best.count(best_mutant.pattern) == 0 && seen.count(best_mutant.pattern) == 0) {
      best.insert(best_mutant.pattern);
      seen.insert(best_mutant.pattern);
      std::cout << "optimization: " << base_to_string(seed, W) << " -> " << iupac_to_string(best_mutant.pattern, W) << std::endl
                << std::endl;
      out.push_back(std::make_unique<Motif>(best_mutant));
    } else {
      std::cout << "optimization: " << base_to_string(seed, W) << " removed" << '\t' << std::endl << std::endl;
    }
  }

  print_table_header();
  std::cout << std::fixed << std::setprecision(2);
  for (const auto& m : out)
    std::cout << std::setw(15) << iupac_to_string(m->pattern, W) << "\t" << std::setw(15) << m->n_sites << "\t" << std::setw(15)
              << ((float)m->n_sites / m->expected) << "\t" << std::setw(15) << m->zscore << std::endl;
}

// Peng::filter_iupac_patterns (src/peng.cpp:543-600)
void filter_iupac_patterns(int W, size_t minimum_retained, std::vector<std::unique_ptr<Motif>>& motifs) {
  std::vector<std::unique_ptr<Motif>> informative;
  for (auto& m : motifs) {
    int n_letters = 0;
    for (int p = 0; p < W; p++) n_letters += iupac_letter_at(m->pattern, p) == kLetterN;
    if (W - n_letters > 3) informative.push_back(std::move(m));
  }
  // the reference sorts raw pointers with std::sort; same algorithm, same comparator, same input order
  std::vector<Motif*> order;
  for (auto& m : informative) order.push_back(m.release());
  std::sort(order.begin(), order.end(), by_log_pvalue);
  float min_pvalue = -5.f;
  if (!order.empty()) min_pvalue = std::min(-5.f, order[0]->log_pvalue * 0.2f);
  motifs.clear();
  for (size_t i = 0; i < order.size(); i++) {
    if (order[i]->log_pvalue < min_pvalue || i < minimum_retained) motifs.emplace_back(order[i]);
    else delete order[i];
  }
}

}  // namespace

namespace {

// IUPACPattern::calculate_adv_pwm (src/iupac_pattern.cpp:505-536) for all motifs: the 4 W single-letter
// substitutions of every motif are counted in one device call.
void calculate_adv_pwms(const Device& dev, int W, int pseudo_counts, const float* v0, std::vector<std::unique_ptr<Motif>>& motifs) {
  std::vector<uint64_t> ids;
  for (const auto& m : motifs)
    for (int p = 0; p < W; p++) {
      const uint64_t masked = m->pattern - (uint64_t)iupac_letter_at(m->pattern, p) * iupac_factor(p);
      for (int i = 0; i < 4; i++) ids.push_back(masked + (uint64_t)i * iupac_factor(p));
    }
  std::vector<uint64_t> occ(ids.size());
  if (!ids.empty()) dev.check(peng_cuda_iupac_count(dev.h, ids.data(), (int)ids.size(), occ.data()));
  size_t at = 0;
  for (auto& m : motifs) {
    m->pwm.assign((size_t)W, PwmRow{0, 0, 0, 0});
    for (int p = 0; p < W; p++) {
      uint64_t total = 0, per_letter[4];
      for (int i = 0; i < 4; i++) {
        per_letter[i] = (uint64_t)((float)pseudo_counts * v0[i]) + occ[at++];
        total += per_letter[i];
      }
      for (int i = 0; i < 4; i++) m->pwm[(size_t)p][i] = (float)(1.0 * (double)per_letter[i] / (double)total);
    }
    make_comp_pwm(*m);
  }
}

// IUPACPattern::calculate_pwm (src/iupac_pattern.cpp:475-503).  The list of base patterns it loops over is never
// filled in the reference, so the matrix is pseudo-counts only — reproduced as is.
void calculate_default_pwms(int W, int pseudo_counts, const float* v0, std::vector<std::unique_ptr<Motif>>& motifs) {
  for (auto& m : motifs) {
    m->pwm.assign((size_t)W, PwmRow{});
    for (int p = 0; p < W; p++)
      for (int i = 0; i < 4; i++)
        m->pwm[(size_t)p][i] = (float)((double)((float)pseudo_counts * v0[i]) / (1.0 * (double)m->n_sites + (double)pseudo_counts));
    make_comp_pwm(*m);
  }
}

// Peng::merge_iupac_patterns (src/peng.cpp:238-313)
void merge_iupac_patterns(int W, const Options& opt, const float* v0, uint32_t max_sequence_length,
                          std::vector<std::unique_ptr<Motif>>& motifs) {
  bool found_better = true;
  while (found_better) {
    found_better = false;
    float best_score = -std::numeric_limits<float>::infinity();
    size_t best_i = 0, best_j = 0;
    int best_shift = 0;
    bool best_comp = false;
    for (size_t i = 0; i < motifs.size(); i++) {
      if (motifs[i]->log_pvalue > -5) continue;
      for (size_t j = i + 1; j < motifs.size(); j++) {
        if (motifs[j]->log_pvalue > -5) continue;
        const BestShift r = best_alignment(*motifs[i], *motifs[j], opt.strand, v0);
        if (r.score > best_score) {
          best_i = i;
          best_j = j;
          best_score = r.score;
          best_shift = r.shift;
          best_comp = r.comp;
        }
      }
    }
    if (motifs.empty()) break;  // the reference would index an empty vector here
    if (best_score > (float)W * opt.merge_bit_factor && (size_t)motifs[best_i]->length <= opt.max_merged_length &&
        (size_t)motifs[best_j]->length <= opt.max_merged_length) {
      const Motif& a = *motifs[best_i];
      const Motif& b = *motifs[best_j];
      Motif merged = a.length < b.length ? merge_motifs(b, a, best_comp, v0, best_shift) : merge_motifs(a, b, best_comp, v0, best_shift);
      if ((size_t)merged.length <= (size_t)max_sequence_length && (size_t)merged.length <= opt.max_merged_length) {
        std::cout << "merge: " << pwm_string(b.pwm) << " + " << pwm_string(a.pwm) << " -> " << pwm_string(merged.pwm) << std::endl;
        motifs.erase(motifs.begin() + (long)best_j);
        motifs.erase(motifs.begin() + (long)best_i);
        motifs.push_back(std::make_unique<Motif>(std::move(merged)));
        found_better = true;
      }
    }
  }
}

// Peng::filter_redundancy (src/peng.cpp:199-236): each kept motif removes at most the first later motif of the
// same length that is too similar to it (or to its reverse complement)
void filter_redundancy(float merge_bit_factor, const float* v0, std::vector<Motif*>& motifs) {
  std::sort(motifs.begin(), motifs.end(), by_log_pvalue);
  std::set<size_t, std::greater<size_t>> deselected;
  for (size_t i = 0; i < motifs.size(); i++) {
    if (deselected.count(i)) continue;
    for (size_t j = i + 1; j < motifs.size(); j++) {
      if (deselected.count(j) || motifs[i]->length != motifs[j]->length) continue;
      const int L = motifs[i]->length;
      const float s = similarity(motifs[i]->pwm, motifs[j]->pwm, v0, 0, 0, L);
      const float sc = similarity(motifs[i]->comp_pwm, motifs[j]->pwm, v0, 0, 0, L);
      const float threshold = merge_bit_factor * (float)L;
      if (s > threshold || sc > threshold) {
        deselected.insert(j);
        break;
      }
    }
  }
  for (size_t index : deselected) {
    delete motifs[index];
    motifs.erase(motifs.begin() + (long)index);
  }
}

}  // namespace

// ---- --gpus N: the data-parallel front of the run (ingest, background counts, k-mer counts) sharded over N devices ----
// One handle and one host thread per device (the ABI's rule).  Every device reads its shard of the FASTA file
// (peng_cuda_ingest_fasta_shard: contiguous runs of records, balanced by bases), counts it, and the per-device tables are
// combined by the fused exchange over peer memory (peng_cuda_count_exchange32: all handles live in this process, so their
// device pointers are exchanged directly) — or, for sets the 32-bit exchange does not cover, on the host.  Every device ends
// with the complete table; the table passes, the hill climb and EM then run on the first device as in a single-GPU run
// (4^W-sized work, "replicas only").  Integer sums: the result is bit-identical to the single-GPU run.
struct MultiGpu {
  std::vector<std::unique_ptr<Device>> extra;  // devices 1 .. N-1 (device 0 is the run's main handle)
  std::vector<peng_handle_t> hs;               // all N handles, rank order
  template <typename F>
  void parallel(F f) {  // f(rank) on one thread per handle; rank 0 on the caller
    std::vector<std::thread> th;
    for (size_t r = 1; r < hs.size(); r++) th.emplace_back([&, r] { f((int)r); });
    f(0);
    for (auto& t : th) t.join();
  }
};

static void multi_count(MultiGpu& mg, const Device& dev, int W, int strand, uint64_t total_positions, uint64_t max_shard_positions,
                        uint64_t* ltot) {
  const int N = (int)mg.hs.size();
  std::vector<int> rc((size_t)N, PENG_OK);
  auto first_error = [&]() {
    bool bad = false;
    for (int r = 0; r < N; r++)
      if (rc[(size_t)r] != PENG_OK) {
        std::cerr << "[device " << r << "] " << peng_cuda_last_error(mg.hs[(size_t)r]) << std::endl;
        bad = true;
      }
    if (bad) exit(1);
  };
  const bool exchange32 = W >= 8 && total_positions < ((uint64_t)1 << 32) - ((uint64_t)1 << 20) && max_shard_positions < ((uint64_t)1 << 30) - 4096;
  if (exchange32) {
    std::vector<unsigned char> blobs((size_t)N * PENG_PEER_BLOB_BYTES);
    for (int r = 0; r < N; r++) rc[(size_t)r] = peng_cuda_peer_export(mg.hs[(size_t)r], W, blobs.data() + (size_t)r * PENG_PEER_BLOB_BYTES);
    first_error();
    for (int r = 0; r < N; r++) rc[(size_t)r] = peng_cuda_peer_attach(mg.hs[(size_t)r], r, N, blobs.data());
    first_error();
    std::vector<uint64_t> lt((size_t)N, 0);
    // two passes: the exchange kernels spin on the device until every rank has arrived, so nothing that may block on the
    // host (first-use kernel loading, allocation) is left to do once the first of them is launched
    mg.parallel([&](int r) {
      rc[(size_t)r] = peng_cuda_count_local32(mg.hs[(size_t)r], W, strand);
      if (rc[(size_t)r] == PENG_OK) rc[(size_t)r] = peng_cuda_synchronize(mg.hs[(size_t)r]);
    });
    first_error();
    // launch everywhere first (ltot = NULL: nothing blocks), then fetch: a blocking read-back of one handle must not sit in
    // the driver while another handle of the same device still has its kernel to launch
    mg.parallel([&](int r) { rc[(size_t)r] = peng_cuda_count_exchange32(mg.hs[(size_t)r], nullptr); });
    first_error();
    mg.parallel([&](int r) { rc[(size_t)r] = peng_cuda_count_finish(mg.hs[(size_t)r], &lt[(size_t)r]); });
    first_error();
    for (int r = 0; r < N; r++) peng_cuda_peer_detach(mg.hs[(size_t)r]);
    *ltot = lt[0];
    return;
  }
  // general case (> 2^32 positions, or W < 8): canonical 64-bit tables summed on the host, loaded into the main handle
  const uint64_t n = (uint64_t)1 << (2 * W);
  std::vector<std::vector<uint64_t>> tab((size_t)N);
  std::vector<uint64_t> lt((size_t)N, 0);
  mg.parallel([&](int r) {
    tab[(size_t)r].resize((size_t)n);
    rc[(size_t)r] = peng_cuda_count(mg.hs[(size_t)r], W, strand, &lt[(size_t)r]);  // per-shard table, mirrored: the sum of mirrored tables is the mirrored sum
    if (rc[(size_t)r] == PENG_OK) rc[(size_t)r] = peng_cuda_get_counts(mg.hs[(size_t)r], tab[(size_t)r].data());
  });
  first_error();
  uint64_t total = 0;
  for (int r = 0; r < N; r++) total += lt[(size_t)r];
  for (int r = 1; r < N; r++)
    for (uint64_t i = 0; i < n; i++) tab[0][(size_t)i] += tab[(size_t)r][(size_t)i];
  dev.check(peng_cuda_set_counts(dev.h, W, strand, tab[0].data(), total));
  *ltot = total;
}

int run(const Options& opt) {
  Stopwatch watch{opt.timings};
  Device dev(opt.device);
  MultiGpu mg;
  mg.hs.push_back(dev.h);
  // PENG_GPUS_SAME_DEVICE=1 (tests on a one-GPU box): all N handles on --device; the whole multi-GPU protocol runs, the
  // exchange kernels meet through the same flag words and pointers as over NVLink
  const bool same_device = getenv("PENG_GPUS_SAME_DEVICE") != nullptr;
  if (opt.gpus > 1) {
    if (!same_device && opt.device + opt.gpus > peng_cuda_device_count()) {
      std::cerr << "Error: --gpus " << opt.gpus << " from device " << opt.device << " on, but only " << peng_cuda_device_count() << " CUDA devices are visible" << std::endl;
      exit(1);
    }
    mg.extra.resize((size_t)opt.gpus - 1);
    mg.hs.resize((size_t)opt.gpus);
    mg.parallel([&](int r) {
      if (r == 0) return;
      mg.extra[(size_t)r - 1] = std::make_unique<Device>(same_device ? opt.device : opt.device + r);
      mg.hs[(size_t)r] = mg.extra[(size_t)r - 1]->h;
    });
  }
  const int N = (int)mg.hs.size();
  watch.lap("CUDA context");

  // ---- Global::init (src/Global.cpp:60-79): the input set, and the background set if it is another file ----
  uint64_t n_seq = 0, total_bases = 0;
  uint32_t min_len = 0, max_len = 0;
  std::vector<uint64_t> shard_positions((size_t)N, 0);
  if (N == 1) {
    dev.check(peng_cuda_ingest_fasta(dev.h, opt.input_fasta.c_str(), &n_seq, &total_bases, &min_len, &max_len));
  } else {
    std::vector<uint64_t> ns((size_t)N, 0), tb((size_t)N, 0);
    std::vector<uint32_t> mn((size_t)N, 0), mx((size_t)N, 0);
    std::vector<int> rc((size_t)N, PENG_OK);
    mg.parallel([&](int r) {
      rc[(size_t)r] = peng_cuda_ingest_fasta_shard(mg.hs[(size_t)r], opt.input_fasta.c_str(), r, N, &ns[(size_t)r], &tb[(size_t)r], &mn[(size_t)r], &mx[(size_t)r]);
    });
    for (int r = 0; r < N; r++)
      if (rc[(size_t)r] != PENG_OK) {  // every shard applies the line rules to the whole file: one message is enough
        std::cerr << peng_cuda_last_error(mg.hs[(size_t)r]) << std::endl;
        exit(1);
      }
    min_len = 0xFFFFFFFFu;
    for (int r = 0; r < N; r++) {
      n_seq += ns[(size_t)r];
      total_bases += tb[(size_t)r];
      shard_positions[(size_t)r] = tb[(size_t)r] + 2 * ns[(size_t)r];
      if (ns[(size_t)r]) { min_len = std::min(min_len, mn[(size_t)r]); max_len = std::max(max_len, mx[(size_t)r]); }
    }
    if (!n_seq) min_len = 0;
  }
  watch.lap("ingest");

  // ---- background model (src/main.cpp:21-26) ----
  const int K = std::max(opt.bg_model_order, opt.max_opt_bg_model_order);
  if (K + 1 > 15) {  // BackgroundModel ctor, src/shared/BackgroundModel.cpp:13-20: floor(log(INT_MAX) / log(4)) = 15
    std::cerr << "Error: The maximum possible order is " << 15 << std::endl;
    exit(-1);
  }
  const std::vector<float> alpha((size_t)K + 1, 1.0f);
  const size_t v_len = (((size_t)1 << (2 * (K + 2))) - 4) / 3;
  std::vector<float> v32(v_len);
  std::vector<double> v64(v_len);
  {
    std::unique_ptr<Device> other;
    const Device* bg_dev = &dev;
    if (!opt.background_fasta.empty()) {
      other = std::make_unique<Device>(opt.device);
      uint64_t n = 0, t = 0;
      uint32_t a = 0, b = 0;
      other->check(peng_cuda_ingest_fasta(other->h, opt.background_fasta.c_str(), &n, &t, &a, &b));
      bg_dev = other.get();
    }
    if (N > 1 && opt.background_fasta.empty()) {
      // sharded input set: the (k+1)-mer counts of the shards add up (no k-mer crosses a sequence), calculateV runs once
      std::vector<std::vector<int64_t>> cnt((size_t)N, std::vector<int64_t>(v_len));
      std::vector<int> rc((size_t)N, PENG_OK);
      mg.parallel([&](int r) {
        std::vector<float> scratch(v_len);
        rc[(size_t)r] = peng_cuda_background_model(mg.hs[(size_t)r], K, alpha.data(), opt.interpolate_bg, PENG_F32REF, scratch.data(), cnt[(size_t)r].data());
      });
      for (int r = 0; r < N; r++) {
        if (rc[(size_t)r] != PENG_OK) { std::cerr << peng_cuda_last_error(mg.hs[(size_t)r]) << std::endl; exit(1); }
        if (r) for (size_t i = 0; i < v_len; i++) cnt[0][i] += cnt[(size_t)r][i];
      }
      dev.check(peng_cuda_background_from_counts(K, alpha.data(), opt.interpolate_bg, PENG_F32REF, cnt[0].data(), v32.data()));
      dev.check(peng_cuda_background_from_counts(K, alpha.data(), opt.interpolate_bg, PENG_F64, cnt[0].data(), v64.data()));
    } else {
      bg_dev->check(peng_cuda_background_model(bg_dev->h, K, alpha.data(), opt.interpolate_bg, PENG_F32REF, v32.data(), nullptr));
      bg_dev->check(peng_cuda_background_model(bg_dev->h, K, alpha.data(), opt.interpolate_bg, PENG_F64, v64.data(), nullptr));
    }
  }
  const float* v0 = v32.data();
  watch.lap("background model");

  // ---- Peng ctor (src/peng.cpp:22-42) ----
  const int k = opt.bg_model_order;
  const int max_k = std::max(k, opt.max_opt_bg_model_order);
  init_motif_tables(v0);

  // ---- Peng::process (src/peng.cpp:322-435) ----
  const int W = opt.pattern_length;
  {
    const double lim_iupac = std::log((double)SIZE_MAX) / std::log((double)kIupacLetters) - 1;
    const double lim_base = std::log((double)SIZE_MAX) / std::log(4.0) - 1;
    if ((double)(size_t)W > lim_iupac || (double)(size_t)W > lim_base) {
      std::cerr << "Warning: pattern length too long!" << std::endl;
      std::cerr << "max pattern length: " << std::max(lim_iupac, lim_base) << std::endl;
      exit(1);
    }
  }
  print_status("Processing kmers of length " + std::to_string(W), false);
  print_status("Finding overrepresented kmers (base patterns)", false);
  const int current_k = std::min(W - 1, k), current_k_max = std::min(W - 1, max_k);
  uint64_t ltot = 0;
  if (N == 1) {
    dev.check(peng_cuda_count(dev.h, W, opt.strand, &ltot));
  } else {
    multi_count(mg, dev, W, opt.strand, total_bases + 2 * n_seq, *std::max_element(shard_positions.begin(), shard_positions.end()), &ltot);
    mg.extra.clear();  // the other devices are done: everything that follows is 4^W-sized work on the main handle
  }
  watch.lap("count k-mers");
  dev.check(peng_cuda_stats(dev.h, K, v32.data(), v64.data(), current_k, current_k_max));
  watch.lap("background / z-score tables");

  const uint64_t n_patterns = (uint64_t)1 << (2 * W);
  std::vector<uint64_t> seeds(n_patterns > ((uint64_t)1 << 24) ? ((uint64_t)1 << 24) : n_patterns);
  uint64_t n_selected = 0;
  dev.check(peng_cuda_select_base_patterns(dev.h, opt.zscore_threshold, opt.count_threshold, opt.strand == PENG_STRAND_PLUS,
                                           opt.filter_neighbors, seeds.data(), seeds.size(), &n_selected));
  seeds.resize((size_t)n_selected);
  watch.lap("select base patterns");
  if (seeds.empty()) std::cout << "No overrepresented seed patterns found. Stopping." << std::endl;

  // BasePattern::print_patterns (src/base_pattern.cpp:517-532): only the rows of the selected ids leave the device
  BaseTables cols;
  cols.counts.resize(seeds.size());
  cols.expected.resize(seeds.size());
  cols.zscore.resize(seeds.size());
  cols.logp.resize(seeds.size());
  if (!seeds.empty()) {
    dev.check(peng_cuda_gather_counts(dev.h, seeds.data(), seeds.size(), cols.counts.data()));
    dev.check(peng_cuda_gather_table(dev.h, PENG_TABLE_EXPECTED, 0, PENG_F32REF, seeds.data(), seeds.size(), cols.expected.data()));
    dev.check(peng_cuda_gather_table(dev.h, PENG_TABLE_ZSCORE, 0, PENG_F32REF, seeds.data(), seeds.size(), cols.zscore.data()));
    dev.check(peng_cuda_gather_table(dev.h, PENG_TABLE_LOGP, 0, PENG_F32REF, seeds.data(), seeds.size(), cols.logp.data()));
  }
  print_table_header();
  std::cout << std::fixed << std::setprecision(2);
  for (size_t i = 0; i < seeds.size(); i++)
    std::cout << std::setw(15) << base_to_string(seeds[i], W) << "\t" << std::setw(15) << cols.counts[i] << "\t" << std::setw(15)
              << ((float)cols.counts[i] / cols.expected[i]) << "\t" << std::setw(15) << cols.zscore[i] << std::endl;

  print_status("Optimizing base patterns");
  std::cout << std::endl;
  if (seeds.size() > (size_t)opt.max_optimized_patterns) seeds.resize((size_t)opt.max_optimized_patterns);
  std::vector<std::unique_ptr<Motif>> unoptimized;
  optimize_iupac_patterns(dev, opt, W, seeds, cols, n_seq, unoptimized);
  watch.lap("optimize IUPAC patterns");
  std::cout << std::endl;

  print_status("Filtering degenerated IUPAC patterns");
  filter_iupac_patterns(W, opt.minimum_processed_motifs, unoptimized);
  for (const auto& m : unoptimized) std::cout << "selected iupac pattern: " << iupac_to_string(m->pattern, W) << std::endl;

  print_status("Calculating PWMs");
  if (opt.adv_pwm) calculate_adv_pwms(dev, W, opt.pseudo_counts, v0, unoptimized);
  else calculate_default_pwms(W, opt.pseudo_counts, v0, unoptimized);
  for (const auto& m : unoptimized) {
    const float avg_info = pwm_information(m->pwm) / (float)(size_t)W;
    std::cout << (opt.adv_pwm ? "adv pwm: " : "def pwm: ") << iupac_to_string(m->pattern, W) << " -> " << pwm_string(m->pwm)
              << "   [ avg. info: " << std::setprecision(2) << avg_info << " ]" << std::endl;
  }
  watch.lap("PWMs");

  print_status("Optimizing expectation-maximization / merging patterns");
  std::vector<Motif*> result;
  {
    const int background = max_k;
    std::cout << std::endl << "background order: " << background << std::endl;
    std::vector<std::unique_ptr<Motif>> optimized;
    if (opt.use_em) {
      const int P = (int)unoptimized.size();
      std::vector<float> in((size_t)P * W * 4), outp((size_t)P * W * 4);
      std::vector<int> iterations((size_t)P);
      for (int i = 0; i < P; i++)
        for (int p = 0; p < W; p++)
          for (int a = 0; a < 4; a++) in[((size_t)i * W + p) * 4 + a] = unoptimized[(size_t)i]->pwm[(size_t)p][a];
      if (P)
        dev.check(peng_cuda_em(dev.h, P, in.data(), PENG_F32REF, (double)opt.em_saturation, (double)opt.em_threshold,
                               opt.em_max_iterations, background, outp.data(), iterations.data()));
      for (int i = 0; i < P; i++) {
        Pwm pwm((size_t)W);
        for (int p = 0; p < W; p++)
          for (int a = 0; a < 4; a++) pwm[(size_t)p][a] = outp[((size_t)i * W + p) * 4 + a];
        // peng_cuda_em already applied the normalisation of IUPACPattern(ori, pwm); with_pwm would repeat it, so the
        // optimized motif is assembled here from the returned matrix as is
        auto m = std::make_unique<Motif>(*unoptimized[(size_t)i]);
        m->pwm = pwm;
        m->zscore = 0;
        make_comp_pwm(*m);
        const float avg_info = pwm_information(m->pwm) / (float)(size_t)W;
        std::cout << "em: " << iupac_to_string(m->pattern, W) << " -> " << pwm_string(m->pwm) << "   [ avg. info: " << std::setprecision(2)
                  << avg_info << " ]" << std::endl;
        optimized.push_back(std::move(m));
      }
      watch.lap("EM");
    } else {
      optimized = std::move(unoptimized);
    }
    if (opt.use_merging) {
      if (W >= 6) merge_iupac_patterns(W, opt, v0, max_len, optimized);
      else std::cerr << "Warning: Specified pattern length (" << W << ") is too low for merging!" << std::endl;
    }
    for (auto& m : optimized) {
      m->opt_bg_order = background;
      result.push_back(m.release());
    }
    watch.lap("merging");
  }

  // ---- src/main.cpp:64-75 ----
  filter_redundancy(opt.merge_bit_factor, v0, result);
  if (!opt.meme_path.empty()) write_meme(result, opt.meme_path, v0);
  if (!opt.json_path.empty()) write_json(result, opt.json_path, v0);
  for (Motif* m : result) delete m;
  watch.lap("filter + output");
  return 0;
}

}  // namespace peng_host
