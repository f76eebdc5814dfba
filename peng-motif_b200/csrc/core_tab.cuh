// core_tab.cuh — per-element bodies of the 4^W-table kernels: background probabilities, expected counts /
// z-scores / log p-values, IUPAC expansion enumeration, EM terms.  PENG_HD, no warp intrinsics (see
// core_seq.cuh for why).  Reference citations are relative to the reference tree.
#pragma once
#include <math.h>
#include <stdint.h>

#include "core_seq.cuh"

namespace peng {

// ---- rounding-exact arithmetic ---------------------------------------------------------------------------
// PENG_F32REF results must equal what g++ -O3 emits for the reference on x86-64 (SSE, no FMA): every
// product / sum is a separately rounded IEEE operation.  nvcc would contract a*b+c into an FMA, so the
// bodies below spell the operations out.
template <typename R> struct Ar;
template <> struct Ar<float> {
  static PENG_HD float mul(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fmul_rn(a, b);
#else
    return a * b;
#endif
  }
  static PENG_HD float add(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fadd_rn(a, b);
#else
    return a + b;
#endif
  }
  static PENG_HD float sub(float a, float b) { return add(a, -b); }
  static PENG_HD float div(float a, float b) {
#if defined(__CUDA_ARCH__)
    return __fdiv_rn(a, b);
#else
    return a / b;
#endif
  }
  static PENG_HD float from_u64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __ull2float_rn(x);
#else
    return (float)x;
#endif
  }
};
template <> struct Ar<double> {
  static PENG_HD double mul(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    return a * b;
#endif
  }
  static PENG_HD double add(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __dadd_rn(a, b);
#else
    return a + b;
#endif
  }
  static PENG_HD double sub(double a, double b) { return add(a, -b); }
  static PENG_HD double div(double a, double b) {
#if defined(__CUDA_ARCH__)
    return __ddiv_rn(a, b);
#else
    return a / b;
#endif
  }
  static PENG_HD double from_u64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __ull2double_rn(x);
#else
    return (double)x;
#endif
  }
};

PENG_HD uint32_t v_offset(int k) { return ((1u << (2 * (k + 1))) - 4u) / 3u; }  // start of v[k] in v[0]|v[1]|..

// ---- (c) background probability of one W-mer -------------------------------------------------------------
// calculate_bg_probabilities + calculate_bg_probability (src/base_pattern.cpp:285-325): starting from 1,
// multiply left to right by v[0][x0], v[1][x0x1], ..., v[order][x0..x_order], then by one conditional of
// full order per further letter; conditionals are indexed first-letter-most-significant (get_bg_id,
// src/base_pattern.h:88-103).
template <typename R>
PENG_HD R bg_prob_of(const R* __restrict__ v, int W, int order, uint32_t id) {
  R p = (R)1;
  uint32_t y = 0;
  const uint32_t ymask = (1u << (2 * (order + 1))) - 1u;
  for (int j = 0; j < W; j++) {
    y = ((y << 2) | ((id >> (2 * j)) & 3u)) & ymask;
    const int k = j < order ? j : order;
    p = Ar<R>::mul(p, v[v_offset(k) + y]);
  }
  return p;
}

// aggregate_double_strand_background (src/base_pattern.cpp:268-283) fused: non-palindromes carry
// p(id) + p(revcomp id) at both ids.
template <typename R>
PENG_HD R bg_prob_entry(const R* __restrict__ v, int W, int order, uint32_t id, bool both) {
  const R p = bg_prob_of<R>(v, W, order, id);
  if (!both) return p;
  const uint32_t rc = revcomp_id(id, W);
  if (rc == id) return p;
  const R q = bg_prob_of<R>(v, W, order, rc);
  return id < rc ? Ar<R>::add(p, q) : Ar<R>::add(q, p);
}

// calculate_expected_counts / calculate_zscores / calculate_log_pvalues (src/base_pattern.cpp:231-265).
// The unqualified sqrt / log there are the double overloads, so the intermediate is double in both arms.
template <typename R>
PENG_HD void stats_of(uint64_t cnt, R bg, uint64_t ltot, R* expected, R* z, R* logp) {
  const R e = Ar<R>::mul(bg, Ar<R>::from_u64(ltot));
  *expected = e;
  const R diff = Ar<R>::sub(Ar<R>::from_u64(cnt), e);
  *z = (R)Ar<double>::div((double)diff, sqrt((double)e));
  if (cnt == 0) {
    *logp = (R)INFINITY;
  } else {
    const R mu = e;
    const R frac = (R)Ar<double>::sub(1.0, (double)Ar<R>::div(mu, Ar<R>::from_u64(cnt + 1)));
    if (Ar<R>::from_u64(cnt) > mu && cnt > 5) {
      const double n = (double)cnt;
      // n*log(mu/n) + n - mu - 0.5*log(6.283*n*frac*frac), evaluated left to right
      double t = Ar<double>::mul(n, log((double)Ar<R>::div(mu, Ar<R>::from_u64(cnt))));
      t = Ar<double>::add(t, n);
      t = Ar<double>::sub(t, (double)mu);
      double a = Ar<double>::mul(6.283, n);
      a = Ar<double>::mul(a, (double)frac);
      a = Ar<double>::mul(a, (double)frac);
      t = Ar<double>::sub(t, Ar<double>::mul(0.5, log(a)));
      *logp = (R)t;
    } else {
      *logp = (R)0;
    }
  }
}

// ---- (d) IUPAC expansion -----------------------------------------------------------------------------------
// An IUPAC pattern arrives as a 64-bit letter mask: nibble p = set of bases allowed at position p
// (representative_iupac_nucleotides, src/iupac_alphabet.cpp:138-180; every set has 1, 2 or 4 members).
struct IupacShape {
  uint32_t nbits[16];  // log2(|set_p|)
  uint32_t total_bits; // log2(number of expansions)
};

PENG_HD IupacShape iupac_shape(uint64_t mask, int W) {
  IupacShape s;
  s.total_bits = 0;
  for (int p = 0; p < 16; p++) {
    uint32_t n = p < W ? (uint32_t)popc32((uint32_t)(mask >> (4 * p)) & 15u) : 1u;
    s.nbits[p] = n == 4 ? 2u : (n == 2 ? 1u : 0u);
    s.total_bits += s.nbits[p];
  }
  return s;
}

// the d-th member (ascending base code) of the set in nibble m
PENG_HD uint32_t nth_member(uint32_t m, uint32_t d) {
  for (uint32_t c = 0; c < 4; c++) {
    if ((m >> c) & 1u) {
      if (d == 0) return c;
      d--;
    }
  }
  return 0;
}

// t-th expansion in the reference's list order for PLUS (basepatterns_from_iupac_single_stranded,
// src/iupac_pattern.cpp:371-408): a mixed-radix count with position 0 slowest and, per position, the
// letters visited as l0, l(n-1), ..., l1 (explicit stack: first alternative followed, others popped LIFO).
PENG_HD uint32_t iupac_expansion_dfs(uint64_t mask, int W, const IupacShape& s, uint64_t t) {
  uint32_t id = 0;
  uint32_t shift = s.total_bits;
  for (int p = 0; p < W; p++) {
    const uint32_t nb = s.nbits[p];
    shift -= nb;
    const uint32_t n = 1u << nb;
    uint32_t d = (uint32_t)(t >> shift) & (n - 1u);
    d = d == 0 ? 0u : n - d;  // digit d -> letter index: 0, n-1, ..., 1
    id |= nth_member((uint32_t)(mask >> (4 * p)) & 15u, d) << (2 * p);
  }
  return id;
}

// t-th expansion in ANY fixed order (position 0 fastest) — enough where the order is irrelevant
PENG_HD uint32_t iupac_expansion_any(uint64_t mask, int W, const IupacShape& s, uint64_t t) {
  uint32_t id = 0;
  for (int p = 0; p < W; p++) {
    const uint32_t nb = s.nbits[p];
    const uint32_t d = (uint32_t)t & ((1u << nb) - 1u);
    t >>= nb;
    id |= nth_member((uint32_t)(mask >> (4 * p)) & 15u, d) << (2 * p);
  }
  return id;
}

PENG_HD bool iupac_matches(uint64_t mask, int W, uint32_t id) {
  bool ok = true;
  for (int p = 0; p < W; p++) ok &= (bool)(((mask >> (4 * p)) >> ((id >> (2 * p)) & 3u)) & 1u);
  return ok;
}

// BOTH strands: the reference maps every expansion to min(id, revcomp), sorts and skips adjacent duplicates
// (src/iupac_pattern.cpp:364-367, 433-441), i.e. it sums over the DISTINCT canonical ids.  Without sorting:
// expansion e represents its canonical id unless its reverse complement is a different expansion of the
// same pattern with a smaller id.
PENG_HD bool iupac_is_representative(uint64_t mask, int W, uint32_t e) {
  const uint32_t rc = revcomp_id(e, W);
  return e <= rc || !iupac_matches(mask, W, rc);
}

// ---- (e) EM ------------------------------------------------------------------------------------------------
// One term of Peng::em_optimize_pwms (src/peng.cpp:121-127) in the reference's float32 arithmetic:
//   odds = (prod_p pwm[p][digit_p], left to right from 1.0f) / bgprob[id]      (calculate_prob_odds, :180-197)
//   term = count * s / (1 + s / odds)
PENG_HD float em_term_f32(const float* __restrict__ pwm, int W, uint32_t id, uint64_t cnt, float bg, float s) {
  float p = 1.0f;
  for (int q = 0; q < W; q++) p = Ar<float>::mul(p, pwm[q * 4 + ((id >> (2 * q)) & 3u)]);
  const float odds = Ar<float>::div(p, bg);
  return Ar<float>::div(Ar<float>::mul(Ar<float>::from_u64(cnt), s), Ar<float>::add(1.0f, Ar<float>::div(s, odds)));
}

// n-th id (ascending) whose digit at position p equals a: the order in which new_pwm[p][a] receives its terms
PENG_HD uint32_t em_chain_id(int p, uint32_t a, uint32_t n) {
  const uint32_t low = n & ((1u << (2 * p)) - 1u);
  const uint32_t high = n >> (2 * p);
  return (uint32_t)(((uint64_t)high << (2 * p + 2)) | ((uint64_t)a << (2 * p)) | low);
}

}  // namespace peng
