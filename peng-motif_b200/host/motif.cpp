// motif.cpp — the small per-motif arithmetic that stays on the host (a few hundred W x 4 matrices per run).
//
// Float discipline: the reference is float32 with DOUBLE libm calls (its translation units include <cmath> only
// and call log / log2 / exp / sqrt unqualified, which resolves to the C double functions) — so a float operand
// is widened for the call, the surrounding expression continues in double and is narrowed when it is stored in
// a float.  Every cast below spells out one of those implicit conversions; -ffp-contract=off keeps the compiler
// from fusing what the reference does not fuse.
#include <algorithm>
#include <cmath>
#include <limits>

#include "peng_host.h"

namespace peng_host {

// ---- IUPAC alphabet ---------------------------------------------------------------------------------------
namespace {
enum { A = 0, C = 1, G = 2, T = 3, S = 4, W_ = 5, R = 6, Y = 7, M = 8, K = 9, N = 10 };
const char kLetters[kIupacLetters + 1] = "ACGTSWRYMK" "N";
// hill-climb neighbourhood of each letter, in the order the reference tries them (src/iupac_alphabet.cpp:47-136)
const std::vector<int> kSimilar[kIupacLetters] = {
    /* A */ {W_, R, M, N},
    /* C */ {S, Y, M, N},
    /* G */ {S, R, K, N},
    /* T */ {W_, Y, K, N},
    /* S */ {C, G, R, Y, M, K, N},
    /* W */ {A, T, R, Y, M, K, N},
    /* R */ {A, G, S, W_, M, K, N},
    /* Y */ {C, T, S, W_, M, K, N},
    /* M */ {A, C, S, W_, R, Y, N},
    /* K */ {G, T, S, W_, R, Y, N},
    /* N */ {A, C, G, T, S, W_, R, Y, M, K},
};
const std::vector<int> kRepresentatives[kIupacLetters] = {
    {A}, {C}, {G}, {T}, {C, G}, {A, T}, {A, G}, {C, T}, {A, C}, {G, T}, {A, C, G, T},
};
float g_log_bonferroni[kIupacLetters];
float g_profile[kIupacLetters][4];
}  // namespace

char iupac_char(int c) { return kLetters[c]; }
const std::vector<int>& similar_letters(int c) { return kSimilar[c]; }
const std::vector<int>& representative_letters(int c) { return kRepresentatives[c]; }

uint64_t iupac_factor(int p) {
  uint64_t f = 1;
  for (int i = 0; i < p; i++) f *= kIupacLetters;
  return f;
}
int iupac_letter_at(uint64_t id, int p) { return (int)((id / iupac_factor(p)) % kIupacLetters); }

std::string iupac_to_string(uint64_t id, int W) {
  std::string s;
  for (int p = 0; p < W; p++, id /= kIupacLetters) s += kLetters[id % kIupacLetters];
  return s;
}
uint64_t base_id_to_iupac_id(uint64_t base_id, int W) {  // BasePattern::baseId2IUPACId, src/base_pattern.cpp:170-178
  uint64_t out = 0;
  for (int p = 0; p < W; p++) out += ((base_id >> (2 * p)) & 3u) * iupac_factor(p);
  return out;
}
std::string base_to_string(uint64_t base_id, int W) {  // BasePattern::toString, src/base_pattern.cpp:109-117
  std::string s;
  for (int p = 0; p < W; p++) s += kLetters[(base_id >> (2 * p)) & 3u];
  return s;
}

// ---- per-run tables (IUPACPattern::init, src/iupac_pattern.cpp:193-240) --------------------------------
void init_motif_tables(const float* v0) {
  static const double bonferroni[kIupacLetters] = {8, 8, 8, 8, 16, 16, 16, 16, 24, 24, 6};
  for (int c = 0; c < kIupacLetters; c++) g_log_bonferroni[c] = (float)std::log(bonferroni[c]);
  const float mixin_factor = 0.2f, mixin_bias = 0.7f;  // MIXIN_FACTOR / MIXIN_BIAS, :25-26
  for (int c = 0; c < kIupacLetters; c++)
    for (int a = 0; a < 4; a++) {
      float x = 0;
      x += mixin_factor * v0[a];
      const auto& rep = kRepresentatives[c];
      if (std::find(rep.begin(), rep.end(), a) != rep.end()) x += mixin_bias;
      g_profile[c][a] = x;
    }
}

// ---- scores --------------------------------------------------------------------------------------------
namespace {
float entropy_nat(float p) {  // calculate_entropy, src/utils.h:25-27
  return (float)(-(double)p * std::log((double)p) - (double)(1 - p) * std::log((double)(1 - p)));
}
float mutual_information_fast(float observed, float expected, unsigned n_sequences, float prior) {  // src/utils.h:29-37
  const float p_obs = (float)(1 - std::exp((double)(-observed / (float)n_sequences)));
  const float p_exp = (float)(1 - std::exp((double)(-expected / (float)n_sequences)));
  const float q = prior;
  const float p = p_obs * q + p_exp * (1 - q);
  return -q * entropy_nat(p_obs) - (1 - q) * entropy_nat(p_exp) + entropy_nat(p);
}
}  // namespace

float mutual_information_score(float observed, float expected, unsigned n_sequences) {
  if (observed < expected) return 0;  // "certainly not interesting" (src/iupac_pattern.cpp:656-659)
  float score = 0;
  for (double prior : {0.5, 0.1, 0.01}) {
    const float q = (float)prior;
    score += mutual_information_fast(observed, expected, n_sequences, q) / entropy_nat(q);
  }
  return -score;
}

// IUPACPattern::aggregate_attributes_from_basepatterns after the three sums (src/iupac_pattern.cpp:443-470)
void set_aggregated_attributes(Motif& m, uint64_t n_sites, float bg_p, float expected) {
  m.bg_p = bg_p;
  m.expected = expected;
  m.zscore = (float)((double)((float)n_sites - expected) / std::sqrt((double)expected));
  m.n_sites = n_sites;
  m.local_sites.assign((size_t)m.length, n_sites);
  if (n_sites == 0) {
    m.log_pvalue = std::numeric_limits<float>::infinity();
    return;
  }
  const float mu = expected;
  const float frac = 1 - mu / (float)(n_sites + 1);
  float logp = 0;
  if ((float)n_sites > mu && n_sites > 5 && m.zscore > 2) {
    const double n = (double)n_sites;
    logp = (float)(n * std::log((double)(mu / (float)n_sites)) + n - (double)mu -
                   0.5 * std::log(6.283 * n * (double)frac * (double)frac));
  }
  for (int p = 0; p < m.length; p++) logp += g_log_bonferroni[iupac_letter_at(m.pattern, p)];
  m.log_pvalue = logp;
}

float motif_score(const Motif& m, Score s, size_t pseudo_expected, unsigned n_sequences) {
  switch (s) {
    case Score::LogPval: return m.log_pvalue;
    case Score::ExpCounts: return (m.expected + (float)pseudo_expected) / (float)m.n_sites;  // :648-650
    default: return mutual_information_score((float)m.n_sites, m.expected, n_sequences);
  }
}

// ---- PWM helpers ---------------------------------------------------------------------------------------
void normalize_pwm(Pwm& pwm) {
  for (auto& row : pwm) {
    float sum = 0;
    for (int a = 0; a < 4; a++) sum += row[a];
    for (int a = 0; a < 4; a++) row[a] /= sum;
  }
}

void make_comp_pwm(Motif& m) {
  const size_t L = m.pwm.size();
  m.comp_pwm.assign(L, PwmRow{});
  for (size_t p = 0; p < L; p++)
    for (int a = 0; a < 4; a++) m.comp_pwm[p][a] = m.pwm[L - 1 - p][3 - a];
}

namespace {
inline double xlog2x(float x) { return (double)x * std::log2((double)x); }

// Jensen-Shannon style distance of two columns, summed over l columns (calculate_d, :539-549)
float column_distance(const PwmRow* a, const PwmRow* b, int l, float epsilon) {
  float d = 0;
  for (int i = 0; i < l; i++)
    for (int x = 0; x < 4; x++) {
      const float mean = (a[i][x] + b[i][x] + 2 * epsilon) / 2;
      d = (float)((double)d + (xlog2x(a[i][x] + epsilon) + xlog2x(b[i][x] + epsilon) - (double)(2 * mean) * std::log2((double)mean)));
    }
  return d;
}
// same against the order-0 background (calculate_d_bg, :551-561)
float background_distance(const PwmRow* a, const float* bg, int l, float epsilon = 1E-4f) {
  float d = 0;
  for (int i = 0; i < l; i++)
    for (int x = 0; x < 4; x++) {
      const float mean = (a[i][x] + bg[x] + 2 * epsilon) / 2;
      d = (float)((double)d + (xlog2x(a[i][x] + epsilon) + xlog2x(bg[x] + epsilon) - (double)(2 * mean) * std::log2((double)mean)));
    }
  return d;
}
}  // namespace

std::string pwm_string(const Pwm& pwm) {
  std::string out;
  for (const PwmRow& row : pwm) {
    float best = std::numeric_limits<float>::infinity();
    int best_letter = 0;
    for (int c = 0; c < kIupacLetters; c++) {
      const PwmRow profile = {g_profile[c][0], g_profile[c][1], g_profile[c][2], g_profile[c][3]};
      const float d = column_distance(&row, &profile, 1, 1E-7f);
      if (d < best) {
        best = d;
        best_letter = c;
      }
    }
    out += kLetters[best_letter];
  }
  return out;
}

float pwm_information(const Pwm& pwm) {
  float total = 0;
  for (const PwmRow& row : pwm)
    for (int a = 0; a < 4; a++)
      if (row[a] != 0) total = (float)((double)total + xlog2x(row[a]));
  return (float)((double)total + (double)pwm.size() * std::log2(4.0));
}

void no_zero_pwm(Pwm& pwm, unsigned precision) {
  const float delta = (float)std::pow(10.0, -(double)(int)precision);
  const float epsilon = delta / (1 - 4 * delta);
  for (auto& row : pwm)
    for (int a = 0; a < 4; a++) row[a] += epsilon;
  normalize_pwm(pwm);
}

float similarity(const Pwm& a, const Pwm& b, const float* bg, int off_a, int off_b, int l) {
  const double half = 0.5 * (double)(background_distance(a.data() + off_a, bg, l) + background_distance(b.data() + off_b, bg, l));
  return (float)(half - (double)column_distance(a.data() + off_a, b.data() + off_b, l, 1E-4f));
}

static const int kMinMergeOverlap = 6;  // MIN_MERGE_OVERLAP, src/iupac_pattern.h:18

BestShift best_alignment(const Motif& p1, const Motif& p2, int strand, const float* bg) {
  const Motif* larger = &p1;
  const Motif* shorter = &p2;
  if (p1.length < p2.length) std::swap(larger, shorter);
  BestShift best{-std::numeric_limits<float>::infinity(), -255, false};
  const int n_orient = strand == PENG_STRAND_BOTH ? 2 : 1;
  for (int orient = 0; orient < n_orient; orient++) {
    const bool comp = orient == 1;
    for (int shift = kMinMergeOverlap - shorter->length; shift <= larger->length - kMinMergeOverlap; shift++) {
      const int off_s = -std::min(shift, 0), off_l = std::max(shift, 0);
      const int overlap = std::min(larger->length - off_l, shorter->length - off_s);
      float s;
      if (!comp) s = similarity(larger->pwm, shorter->pwm, bg, off_l, off_s, overlap);
      else if (larger->n_sites < shorter->n_sites) s = similarity(larger->comp_pwm, shorter->pwm, bg, off_l, off_s, overlap);
      else s = similarity(larger->pwm, shorter->comp_pwm, bg, off_l, off_s, overlap);
      if (s > best.score) best = BestShift{s, shift, comp};
    }
  }
  return best;
}

namespace {
// calculate_merged_pvalue (src/iupac_pattern.cpp:242-289); the orientation choice below is the reference's: the
// shorter motif's complement is used whenever the first condition fails, also for same-strand merges
float merged_log_pvalue(const Motif& longer, const Motif& shorter, bool comp, const float* bg, int shift) {
  const Pwm* lp = &longer.pwm;
  const Pwm* sp = &shorter.pwm;
  if (comp && longer.n_sites < shorter.n_sites) lp = &longer.comp_pwm;
  else sp = &shorter.comp_pwm;
  const int off_s = -std::min(shift, 0), off_l = std::max(shift, 0);
  const int overlap = std::min(longer.length - off_l, shorter.length - off_s);
  const Motif& weak = longer.log_pvalue < shorter.log_pvalue ? shorter : longer;    // the one that is discounted
  const Motif& strong = longer.log_pvalue < shorter.log_pvalue ? longer : shorter;
  const Pwm& wp = &weak == &shorter ? *sp : *lp;
  const int off_w = &weak == &shorter ? off_s : off_l;
  float d;
  if (off_w != 0) d = background_distance(wp.data(), bg, off_w);  // starts with its non-overlapping part
  else {
    const int start = off_w + overlap;
    d = background_distance(wp.data() + start, bg, weak.length - start);
  }
  const float divisor = background_distance(wp.data(), bg, weak.length);
  return strong.log_pvalue + d / divisor * weak.log_pvalue;
}
}  // namespace

Motif merge_motifs(const Motif& longer, const Motif& shorter, bool comp, const float* bg, int shift) {
  const int off_s = -std::min(shift, 0), off_l = std::max(shift, 0);
  const int overlap = std::min(longer.length - off_l, shorter.length - off_s);
  const Pwm* lp = &longer.pwm;
  const Pwm* sp = &shorter.pwm;
  if (comp && longer.n_sites < shorter.n_sites) lp = &longer.comp_pwm;
  else if (comp) sp = &shorter.comp_pwm;

  Motif m;
  m.length = longer.length + shorter.length - overlap;
  m.local_sites.assign((size_t)m.length, 0);
  for (int p = 0; p < shorter.length; p++) m.local_sites[(size_t)(std::max(shift, 0) + p)] += shorter.local_sites[(size_t)p];
  for (int p = 0; p < longer.length; p++) m.local_sites[(size_t)(-std::min(shift, 0) + p)] += longer.local_sites[(size_t)p];
  int sum_sites = 0;  // int in the reference (:103-107)
  for (int p = 0; p < m.length; p++) sum_sites += (int)m.local_sites[(size_t)p];
  m.n_sites = (uint64_t)sum_sites / (uint64_t)m.length;

  m.pwm.assign((size_t)m.length, PwmRow{0, 0, 0, 0});
  for (int p = 0; p < m.length; p++) {
    const int ps = p - std::max(0, shift);
    const int pl = p + std::min(shift, 0);
    const bool in_s = ps >= 0 && ps < shorter.length, in_l = pl >= 0 && pl < longer.length;
    if (in_s && in_l) {
      const uint64_t ns = shorter.local_sites[(size_t)ps], nl = longer.local_sites[(size_t)pl];
      for (int a = 0; a < 4; a++)
        m.pwm[(size_t)p][a] = ((float)ns * (*sp)[(size_t)ps][a] + (float)nl * (*lp)[(size_t)pl][a]) / (float)(ns + nl);
    } else if (in_l) {
      m.pwm[(size_t)p] = (*lp)[(size_t)pl];
    } else if (in_s) {
      m.pwm[(size_t)p] = (*sp)[(size_t)ps];
    }
  }
  normalize_pwm(m.pwm);
  make_comp_pwm(m);
  m.log_pvalue = merged_log_pvalue(longer, shorter, comp, bg, shift);
  m.bg_p = 0;
  m.merged = true;
  return m;
}

Motif with_pwm(const Motif& ori, const Pwm& pwm) {
  Motif m;
  m.pattern = ori.pattern;
  m.length = ori.length;
  m.local_sites = ori.local_sites;
  m.pwm = pwm;
  normalize_pwm(m.pwm);
  make_comp_pwm(m);
  m.n_sites = ori.n_sites;
  m.log_pvalue = ori.log_pvalue;
  m.bg_p = ori.bg_p;
  m.expected = ori.expected;
  m.merged = ori.merged;
  m.opt_bg_order = ori.opt_bg_order;
  return m;
}

bool by_log_pvalue(const Motif* a, const Motif* b) { return a->log_pvalue < b->log_pvalue; }

}  // namespace peng_host
