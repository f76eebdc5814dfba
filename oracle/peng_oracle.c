/* peng_oracle.c — TEST INFRASTRUCTURE ONLY.  See peng_oracle.h for scope and parity status.
 *
 * Compiled twice by oracle/Makefile: -DREAL=float -DSUF=_f32 -DPO_WITH_COMMON and
 * -DREAL=double -DSUF=_f64.  With REAL=float every expression below has the
 * operand types of the reference expression it restates, so that gcc emits the
 * same IEEE operations (x86-64 SSE, no contraction: -ffp-contract=off).  Note
 * that the reference calls the unqualified sqrt/log/exp with only <cmath>
 * included, which g++/libstdc++ resolve to the double overloads; plain C does
 * the same, so `sqrt(x)` below means sqrt((double)x) exactly as there.
 */
#include "peng_oracle.h"

#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#ifndef REAL
#error "compile with -DREAL=float|double -DSUF=_f32|_f64"
#endif
#define CAT_(a, b) a##b
#define CAT(a, b) CAT_(a, b)
#define FN(name) CAT(name, SUF)

static inline uint64_t pow4(int e) { return (uint64_t)1 << (2 * e); }

static uint64_t pow11(int e) {
  uint64_t r = 1;
  while (e-- > 0) r *= 11;
  return r;
}

static inline int iupac_letter(uint64_t id, int pos) { return (int)((id / pow11(pos)) % 11); }

#ifdef PO_WITH_COMMON
/* IUPACAlphabet::representative_iupac_nucleotides (src/iupac_alphabet.cpp:138-180) */
static const int kRepN[11] = {1, 1, 1, 1, 2, 2, 2, 2, 2, 2, 4};
static const int kRep[11][4] = {{0}, {1}, {2}, {3}, {1, 2}, {0, 3}, {0, 2}, {1, 3}, {0, 1}, {2, 3}, {0, 1, 2, 3}};

/* ------------------------------------------------------------------------- */
/* ids                                                                        */
/* ------------------------------------------------------------------------- */

/* BasePattern::getRevCompId (src/base_pattern.cpp:119-134): digit p of the result is the complement
 * (3 - c) of digit W-1-p. */
uint64_t po_revcomp(uint64_t id, int W) {
  uint64_t r = 0;
  for (int p = 0; p < W; p++) {
    uint64_t c = (id >> (2 * p)) & 3;
    r |= (3 - c) << (2 * (W - 1 - p));
  }
  return r;
}

/* BasePattern::get_bg_id (src/base_pattern.h:88-103): letters pattern_length-k-1 .. pattern_length-1 of a
 * PEnG id, re-encoded first-letter-most-significant (BaMM order). */
uint64_t po_bg_id(uint64_t pattern, int pattern_length, int k) {
  uint64_t y = 0;
  for (int i = pattern_length - k - 1; i < pattern_length; i++) {
    uint64_t c = (pattern >> (2 * i)) & 3;
    y += c * pow4(pattern_length - i - 1);
  }
  return y;
}

/* ------------------------------------------------------------------------- */
/* FASTA ingest (src/shared/SequenceSet.cpp:285-447)                           */
/* ------------------------------------------------------------------------- */
typedef struct {
  uint8_t* codes; /* NULL on the sizing pass */
  uint64_t* offsets;
  uint64_t n_seq, total;
  uint32_t min_len, max_len;
} fasta_sink;

static void sink_commit(fasta_sink* s, const char* seq, uint64_t L) {
  /* Alphabet::getCode (src/shared/Alphabet.cpp:33-41): ACGT / acgt -> 1..4, anything else -> 0 */
  if (s->codes) {
    for (uint64_t i = 0; i < L; i++) {
      uint8_t c = 0;
      switch (seq[i]) {
        case 'A': case 'a': c = 1; break;
        case 'C': case 'c': c = 2; break;
        case 'G': case 'g': c = 3; break;
        case 'T': case 't': c = 4; break;
        default: c = 0;
      }
      s->codes[s->total + i] = c;
    }
    s->offsets[s->n_seq] = s->total;
    s->offsets[s->n_seq + 1] = s->total + L;
  }
  s->n_seq++;
  s->total += L;
  if (L > s->max_len) s->max_len = (uint32_t)L;
  if (L < s->min_len) s->min_len = (uint32_t)L;
}

static int fasta_parse(const char* path, fasta_sink* s) {
  FILE* f = fopen(path, "rb");
  if (!f) return 1; /* "Cannot open FASTA file" -> exit(1) (:430-433) */
  fseek(f, 0, SEEK_END);
  long sz = ftell(f);
  fseek(f, 0, SEEK_SET);
  char* buf = (char*)malloc((size_t)sz + 1);
  if (sz > 0 && fread(buf, 1, (size_t)sz, f) != (size_t)sz) { fclose(f); free(buf); return 1; }
  fclose(f);

  char* seq = (char*)malloc((size_t)sz + 1);
  uint64_t seq_len = 0;
  int have_header = 0;
  int rc = 0;
  long pos = 0;
  s->n_seq = 0; s->total = 0; s->max_len = 0; s->min_len = 0x7fffffffu; /* int max (:296) */
  /* `while (getline(file, line).good())` (:304): a final line that is not terminated by '\n' sets eofbit
   * and is therefore dropped */
  while (pos < sz) {
    char* nl = (char*)memchr(buf + pos, '\n', (size_t)(sz - pos));
    if (!nl) break;
    long len = (long)(nl - (buf + pos));
    const char* line = buf + pos;
    pos += len + 1;
    if (len == 0) continue; /* blank lines skipped (:306) */
    if (line[0] == '>') {
      if (have_header) {
        if (seq_len) { sink_commit(s, seq, seq_len); seq_len = 0; } /* else: warning, entry ignored (:346) */
      }
      have_header = 1; /* ">" alone gets the running index as header (:351-354): non-empty either way */
    } else if (have_header) {
      if (memchr(line, ' ', (size_t)len)) { rc = 1; break; } /* space in sequence -> exit(1) (:361-364) */
      memcpy(seq + seq_len, line, (size_t)len);
      seq_len += (uint64_t)len;
    } else {
      rc = 1; /* sequence before any header -> exit(1) (:369-373) */
      break;
    }
  }
  if (!rc && have_header && seq_len) sink_commit(s, seq, seq_len); /* last record (:379-421) */
  free(seq);
  free(buf);
  return rc;
}

int po_fasta_scan(const char* path, uint64_t* n_seq, uint64_t* total_len, uint32_t* min_len, uint32_t* max_len) {
  fasta_sink s;
  memset(&s, 0, sizeof s);
  int rc = fasta_parse(path, &s);
  *n_seq = s.n_seq; *total_len = s.total; *min_len = s.min_len; *max_len = s.max_len;
  return rc;
}

int po_fasta_read(const char* path, uint8_t* codes, uint64_t* offsets) {
  fasta_sink s;
  memset(&s, 0, sizeof s);
  s.codes = codes;
  s.offsets = offsets;
  offsets[0] = 0;
  return fasta_parse(path, &s);
}

/* ------------------------------------------------------------------------- */
/* counting (src/base_pattern.cpp:331-441)                                     */
/* ------------------------------------------------------------------------- */
void po_count(const uint8_t* codes, const uint64_t* offsets, uint64_t n_seq, int W, int both, uint64_t* counts,
              uint64_t* ltot_out) {
  const uint64_t n = pow4(W), mask = n - 1;
  memset(counts, 0, n * sizeof(uint64_t));
  /* last_match_pos / j are `unsigned int` in the reference (:335-336); 64-bit here so that inputs beyond
   * 2^32 concatenated positions stay exact.  j = coordinate just right of the current W-mer in the
   * concatenation of all sequences, which starts with W padding positions. */
  uint64_t* last = (uint64_t*)calloc(n, sizeof(uint64_t));
  uint64_t j = (uint64_t)W, ltot = 0;
  for (uint64_t s = 0; s < n_seq; s++) {
    const uint8_t* seq = codes + offsets[s];
    const uint64_t L = offsets[s + 1] - offsets[s];
    uint64_t i = 0;
    while (i < L) {
      /* open a run: W consecutive letters starting at i (:350-352) */
      int p = 0;
      uint64_t id = 0;
      while (p < W && i < L && seq[i] > 0) {
        id += (uint64_t)(seq[i] - 1) << (2 * p);
        p++; i++; j++;
      }
      if (p < W) { i++; j++; continue; } /* failed on seq[i]: resume behind it (:353 + loop ++ at :347) */
      uint64_t idrev = po_revcomp(id, W);
      for (;;) { /* slide the window to the right (:359-379) */
        uint64_t key = both ? (id < idrev ? id : idrev) : id;
        if (last[key] + (uint64_t)W <= j) { /* previous COUNTED hit of this key ended >= W positions ago */
          counts[key]++;
          last[key] = j;
        }
        ltot++;
        if (i >= L || seq[i] == 0) break;
        uint64_t c = (uint64_t)(seq[i] - 1);
        id = (id >> 2) | (c << (2 * (W - 1)));
        idrev = ((idrev << 2) & mask) | (3 - c);
        i++; j++;
      }
      /* the run stopped at index i (end of sequence or a non-ACGT letter): the reference steps over i
       * (:380) AND the for-increment (:347) runs, so scanning resumes at i + 2 */
      i += 2; j += 2;
    }
    j += (uint64_t)W; /* (:382) */
  }
  free(last);
  if (both) { /* second copy of the pair count at the reverse-complement id (:387-392) */
    for (uint64_t id = 0; id < n; id++) {
      uint64_t rc = po_revcomp(id, W);
      if (id > rc) counts[id] = counts[rc];
    }
  }
  *ltot_out = ltot;
}

/* ------------------------------------------------------------------------- */
/* IUPAC expansion (src/iupac_pattern.cpp:331-408)                             */
/* ------------------------------------------------------------------------- */
uint64_t po_iupac_expansion_size(uint64_t iupac_id, int W) {
  uint64_t n = 1;
  for (int p = 0; p < W; p++) n *= (uint64_t)kRepN[iupac_letter(iupac_id, p)];
  return n;
}

static int cmp_u64(const void* a, const void* b) {
  uint64_t x = *(const uint64_t*)a, y = *(const uint64_t*)b;
  return x < y ? -1 : x > y;
}

uint64_t po_iupac_expand(uint64_t iupac_id, int W, int both, uint64_t* out, uint64_t cap) {
  /* The reference walks an explicit stack: the first alternative of every position is followed
   * immediately, the others are pushed in order and popped last-in-first-out.  The resulting order is a
   * mixed-radix count with position 0 slowest and per-position letter order l0, l(n-1), ..., l1. */
  int letters[32][4], nl[32], digit[32];
  uint64_t n = 1;
  for (int p = 0; p < W; p++) {
    int c = iupac_letter(iupac_id, p);
    nl[p] = kRepN[c];
    letters[p][0] = kRep[c][0];
    for (int d = 1; d < nl[p]; d++) letters[p][d] = kRep[c][nl[p] - d];
    digit[p] = 0;
    n *= (uint64_t)nl[p];
  }
  if (n > cap) return n;
  for (uint64_t t = 0; t < n; t++) {
    uint64_t id = 0;
    for (int p = 0; p < W; p++) id |= (uint64_t)letters[p][digit[p]] << (2 * p);
    if (both) {
      uint64_t rc = po_revcomp(id, W);
      out[t] = rc < id ? rc : id; /* (:364) */
    } else {
      out[t] = id;
    }
    for (int p = W - 1; p >= 0; p--) { /* last position varies fastest */
      if (++digit[p] < nl[p]) break;
      digit[p] = 0;
    }
  }
  if (both) qsort(out, n, sizeof(uint64_t), cmp_u64); /* std::sort of integers: order is unique (:367) */
  return n;
}

uint64_t po_count_combined(uint64_t iupac_id, int W, int both, const uint64_t* counts) {
  uint64_t n = po_iupac_expansion_size(iupac_id, W);
  uint64_t* ids = (uint64_t*)malloc(n * sizeof(uint64_t));
  po_iupac_expand(iupac_id, W, both, ids, n);
  uint64_t total = counts[ids[0]];
  for (uint64_t t = 1; t < n; t++) {
    if (!both || ids[t] != ids[t - 1]) total += counts[ids[t]]; /* adjacent duplicates skipped (:822-830) */
  }
  free(ids);
  return total;
}

int po_similar_letters(int c, int* out) {
  /* src/iupac_alphabet.cpp:47-136, letters A,C,G,T,S,W,R,Y,M,K,N = 0..10 */
  static const int n[11] = {4, 4, 4, 4, 7, 7, 7, 7, 7, 7, 10};
  static const int t[11][10] = {
      {5, 6, 8, 10},          {4, 7, 8, 10},          {4, 6, 9, 10},          {5, 7, 9, 10},
      {1, 2, 6, 7, 8, 9, 10}, {0, 3, 6, 7, 8, 9, 10}, {0, 2, 4, 5, 8, 9, 10}, {1, 3, 4, 5, 8, 9, 10},
      {0, 1, 4, 5, 6, 7, 10}, {2, 3, 4, 5, 6, 7, 10}, {0, 1, 2, 3, 4, 5, 6, 7, 8, 9}};
  for (int i = 0; i < n[c]; i++) out[i] = t[c][i];
  return n[c];
}
#endif /* PO_WITH_COMMON */

/* ------------------------------------------------------------------------- */
/* background model (src/shared/BackgroundModel.cpp:60-84, 490-530)            */
/* ------------------------------------------------------------------------- */
void FN(po_background_model)(const uint8_t* codes, const uint64_t* offsets, uint64_t n_seq, int K,
                             const REAL* alpha, int interpolate, REAL* v_out, int64_t* n_out) {
  int64_t* n[16];
  REAL* v[16];
  size_t off = 0;
  int64_t* n_store = (int64_t*)calloc((size_t)((pow4(K + 2) - 4) / 3), sizeof(int64_t));
  for (int k = 0; k <= K; k++) {
    n[k] = n_store + off;
    v[k] = v_out + off;
    off += (size_t)pow4(k + 1);
  }
  for (uint64_t s = 0; s < n_seq; s++) {
    const uint8_t* seq = codes + offsets[s];
    const int64_t L = (int64_t)(offsets[s + 1] - offsets[s]);
    for (int64_t i = 0; i < L; i++) {
      /* Sequence::kmer_[i] (src/shared/Sequence.cpp:28-33): BaMM id of the <=9-mer ending at i, every
       * non-ACGT letter inside the window contributing -4^10 */
      int64_t kmer = 0;
      for (int64_t jj = i < 8 ? i : 8; jj >= 0; jj--)
        kmer += seq[i - jj] > 0 ? (int64_t)(seq[i - jj] - 1) * (int64_t)pow4((int)jj) : -(int64_t)pow4(10);
      for (int k = 0; k <= K && k <= i; k++) {
        int64_t y = kmer % (int64_t)pow4(k + 1); /* C remainder keeps the sign of kmer (:75) */
        if (y >= 0) n[k][y]++;                   /* (:77-80) */
      }
    }
  }
  /* calculateV (:490-530) */
  int64_t base_counts = 0;
  for (int y = 0; y < 4; y++) base_counts += n[0][y];
  for (int y = 0; y < 4; y++)
    v[0][y] = ((REAL)n[0][y] + alpha[0] * 0.25f) / ((REAL)base_counts + alpha[0]);
  for (int k = 1; k <= K; k++) {
    const int64_t Yk1 = (int64_t)pow4(k + 1), Yk = (int64_t)pow4(k);
    for (int64_t y = 0; y < Yk1; y++) {
      int64_t y2 = y % Yk, yk = y / 4;
      if (interpolate)
        v[k][y] = ((REAL)n[k][y] + alpha[k] * v[k - 1][y2]) / ((REAL)n[k - 1][yk] + alpha[k]);
      else
        v[k][y] = ((REAL)n[k][y] + alpha[k] * 0.25f) / ((REAL)n[k - 1][yk] + alpha[k]);
    }
    REAL factor = 0.0f;
    for (int64_t y = 0; y < Yk1; y++) {
      factor += v[k][y];
      if ((y + 1) % 4 == 0) {
        for (int i = 0; i < 4; i++) v[k][y - i] /= factor;
        factor = 0.0f;
      }
    }
  }
  if (n_out) memcpy(n_out, n_store, off * sizeof(int64_t));
  free(n_store);
}

/* ------------------------------------------------------------------------- */
/* background probabilities of all W-mers (src/base_pattern.cpp:285-325)       */
/* ------------------------------------------------------------------------- */
void FN(po_bg_probabilities)(const REAL* v_all, int W, int order, REAL* out) {
  const REAL* v[16];
  size_t off = 0;
  for (int k = 0; k <= order; k++) { v[k] = v_all + off; off += (size_t)pow4(k + 1); }
  const uint64_t n = pow4(W);
  for (uint64_t id = 0; id < n; id++) {
    /* ramp-up over the first order+1 letters (:291-294), then one conditional of full order per further
     * letter (:312-316); strictly left-to-right float products starting from 1.0f */
    REAL p = 1.0;
    for (int kp = 0; kp <= order; kp++) p *= v[kp][po_bg_id(id & (pow4(order + 1) - 1), kp + 1, kp)];
    for (int len = order + 2; len <= W; len++) p = p * v[order][po_bg_id(id & (pow4(len) - 1), len, order)];
    out[id] = p;
  }
}

void FN(po_aggregate_double_strand)(REAL* bg, int W) {
  const uint64_t n = pow4(W);
  for (uint64_t id = 0; id < n; id++) { /* ascending, in place (:271-280) */
    uint64_t rc = po_revcomp(id, W);
    if (id < rc) bg[id] += bg[rc];
    else if (id > rc) bg[id] = bg[rc];
  }
}

/* ------------------------------------------------------------------------- */
/* expected counts, z-scores, log p-values (src/base_pattern.cpp:231-265)      */
/* ------------------------------------------------------------------------- */
void FN(po_stats)(const uint64_t* counts, const REAL* bgprob_k, uint64_t ltot, int W, REAL* expected, REAL* z,
                  REAL* logp) {
  const uint64_t n = pow4(W);
  for (uint64_t id = 0; id < n; id++) {
    expected[id] = bgprob_k[id] * (REAL)ltot; /* float * size_t (:263) */
    const uint64_t cnt = counts[id];
    if (cnt == 0) {
      logp[id] = (REAL)INFINITY;
    } else {
      REAL mu = expected[id];
      REAL frac = (REAL)(1.0 - mu / (REAL)(cnt + 1)); /* (:239) */
      if ((REAL)cnt > mu && cnt > 5) {
        logp[id] = (REAL)((double)cnt * log(mu / (REAL)cnt) + (double)cnt - mu -
                          0.5 * log(6.283 * (double)cnt * frac * frac)); /* (:242-243) */
      } else {
        logp[id] = 0;
      }
    }
    z[id] = (REAL)(((REAL)cnt - expected[id]) / sqrt(expected[id])); /* (:255-256) */
  }
}

/* utils.h:25-27, 29-37 */
static REAL entropy_nat(REAL p) { return (REAL)(-p * log(p) - (1 - p) * log(1 - p)); }

static REAL mutual_information_fast(REAL observed, REAL expected, unsigned int n_sequences, REAL prior) {
  REAL p_obs = (REAL)(1 - exp(-observed / (REAL)n_sequences));
  REAL p_exp = (REAL)(1 - exp(-expected / (REAL)n_sequences));
  REAL q = prior;
  REAL p = p_obs * q + p_exp * (1 - q);
  return -q * entropy_nat(p_obs) - (1 - q) * entropy_nat(p_exp) + entropy_nat(p);
}

static REAL mi_score(REAL observed, REAL expected, unsigned int n_sequences) {
  static const double priors[3] = {0.5, 0.1, 0.01};
  REAL score = 0;
  for (int i = 0; i < 3; i++) {
    REAL q = (REAL)priors[i];
    score += mutual_information_fast(observed, expected, n_sequences, q) / entropy_nat(q);
  }
  return -score;
}

REAL FN(po_base_mi_score)(uint64_t count, REAL expected, uint64_t n_sequences) {
  unsigned int observed = (unsigned int)count; /* (:186) */
  if ((REAL)observed < expected) return 0;
  return mi_score((REAL)observed, expected, (unsigned int)n_sequences);
}

/* ------------------------------------------------------------------------- */
/* IUPAC sums (src/iupac_pattern.cpp:410-473, 505-536)                         */
/* ------------------------------------------------------------------------- */
void FN(po_iupac_aggregate)(uint64_t iupac_id, int W, int both, const uint64_t* counts, const REAL* bgprob_k,
                            const REAL* expected, uint64_t n_sequences, uint64_t* n_sites_out, REAL* out5) {
  uint64_t n = po_iupac_expansion_size(iupac_id, W);
  uint64_t* ids = (uint64_t*)malloc(n * sizeof(uint64_t));
  po_iupac_expand(iupac_id, W, both, ids, n);
  uint64_t last = ids[0];
  REAL sum_bg = bgprob_k[last];
  uint64_t sum_counts = counts[last];
  REAL sum_exp = expected[last];
  for (uint64_t t = 1; t < n; t++) { /* sequential float sums in list order (:433-441) */
    uint64_t cur = ids[t];
    if (cur != last) {
      sum_bg += bgprob_k[cur];
      sum_counts += counts[cur];
      sum_exp += expected[cur];
    }
    last = cur;
  }
  free(ids);
  REAL zscore = (REAL)(((REAL)sum_counts - sum_exp) / sqrt(sum_exp)); /* (:446) */
  REAL logp;
  if (sum_counts == 0) {
    logp = (REAL)INFINITY;
  } else {
    REAL mu = sum_exp;
    REAL frac = 1 - mu / (REAL)(sum_counts + 1); /* float arithmetic here (:457) */
    logp = 0;
    if ((REAL)sum_counts > mu && sum_counts > 5 && zscore > 2)
      logp = (REAL)((double)sum_counts * log(mu / (REAL)sum_counts) + (double)sum_counts - mu -
                    0.5 * log(6.283 * (double)sum_counts * frac * frac));
    /* IUPACPattern::init log_bonferroni (:199-210) */
    static const double bonf[11] = {8, 8, 8, 8, 16, 16, 16, 16, 24, 24, 6};
    for (int p = 0; p < W; p++) logp += (REAL)log(bonf[iupac_letter(iupac_id, p)]);
  }
  *n_sites_out = sum_counts;
  out5[0] = sum_bg;
  out5[1] = sum_exp;
  out5[2] = zscore;
  out5[3] = logp;
  /* getMutualInformationScore (:652-669): observed is the float image of n_sites */
  REAL observed = (REAL)sum_counts;
  out5[4] = observed < sum_exp ? (REAL)0 : mi_score(observed, sum_exp, (unsigned int)n_sequences);
}

void FN(po_adv_pwm)(uint64_t iupac_id, int W, int both, const uint64_t* counts, int pseudo_counts,
                    const REAL* v0, REAL* pwm) {
  for (int p = 0; p < W; p++) {
    uint64_t f = pow11(p);
    uint64_t base = iupac_id - (uint64_t)iupac_letter(iupac_id, p) * f;
    uint64_t i_total[4], n_total = 0;
    for (int i = 0; i < 4; i++) {
      i_total[i] = (uint64_t)((REAL)pseudo_counts * v0[i]); /* int * float truncated to size_t (:524) */
      i_total[i] += po_count_combined(base + (uint64_t)i * f, W, both, counts);
      n_total += i_total[i];
    }
    for (int i = 0; i < 4; i++) pwm[p * 4 + i] = (REAL)(1.0 * (double)i_total[i] / (double)n_total); /* (:530) */
  }
}

/* ------------------------------------------------------------------------- */
/* EM (src/peng.cpp:48-197)                                                    */
/* ------------------------------------------------------------------------- */
void FN(po_prob_odds)(const REAL* pwm, int W, const REAL* bgprob, REAL* odds) {
  const uint64_t n = pow4(W);
  for (uint64_t id = 0; id < n; id++) {
    REAL p = 1.0; /* product over positions 0..W-1, left to right (:186, recursion depth = position) */
    for (int q = 0; q < W; q++) p = p * pwm[q * 4 + ((id >> (2 * q)) & 3)];
    odds[id] = p / bgprob[id]; /* (:195) */
  }
}

static void normalize_pwm(REAL* pwm, int W) { /* src/iupac_pattern.cpp:291-303 */
  for (int p = 0; p < W; p++) {
    REAL sum = 0;
    for (int a = 0; a < 4; a++) sum += pwm[p * 4 + a];
    for (int a = 0; a < 4; a++) pwm[p * 4 + a] /= sum;
  }
}

void FN(po_em_iteration)(const REAL* old_pwm, int W, const uint64_t* counts, const REAL* bgprob, REAL saturation,
                         REAL* new_pwm, REAL* odds) {
  const uint64_t n = pow4(W);
  for (int i = 0; i < 4 * W; i++) new_pwm[i] = 0.0;
  FN(po_prob_odds)(old_pwm, W, bgprob, odds);
  for (uint64_t id = 0; id < n; id++) { /* ascending ids, sequential accumulation (:121-127) */
    REAL w = (REAL)counts[id] * saturation / (1 + saturation / odds[id]);
    for (int p = 0; p < W; p++) new_pwm[p * 4 + ((id >> (2 * p)) & 3)] += w;
  }
  normalize_pwm(new_pwm, W);
}

int FN(po_em)(const REAL* pwm_in, int W, const uint64_t* counts, const REAL* bgprob, REAL saturation,
              REAL min_threshold, int max_iterations, REAL* pwm_out) {
  const uint64_t n = pow4(W);
  REAL* odds = (REAL*)malloc(n * sizeof(REAL));
  REAL a[64], b[64];
  REAL *old_pwm = a, *new_pwm = b;
  memcpy(old_pwm, pwm_in, sizeof(REAL) * 4 * (size_t)W);
  REAL change = (REAL)W;
  int it = 0;
  while (!(change <= min_threshold || it >= max_iterations)) { /* (:104-107) */
    it++;
    FN(po_em_iteration)(old_pwm, W, counts, bgprob, saturation, new_pwm, odds);
    change = 0;
    for (int i = 0; i < 4 * W; i++) change += (REAL)fabs(new_pwm[i] - old_pwm[i]); /* std::abs(float) (:135) */
    REAL* t = old_pwm; old_pwm = new_pwm; new_pwm = t;
  }
  memcpy(pwm_out, old_pwm, sizeof(REAL) * 4 * (size_t)W);
  normalize_pwm(pwm_out, W); /* IUPACPattern(ori, pwm) normalises once more (src/iupac_pattern.cpp:61) */
  free(odds);
  return it;
}
