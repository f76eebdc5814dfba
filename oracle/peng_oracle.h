/* peng_oracle.h — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement ("port") of the PEnG-motif hot path, written from the
 * reference's behaviour; each function cites the reference file:line it follows
 * (paths relative to /root/reference).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product (peng-motif_b200/) never does.
 *
 * PARITY STATUS: pinned.  tests/test_oracle_vs_golden.py checks every function
 * against (a) the encoding identities the reference's own unit tests hold
 * (test/test_base_pattern.cpp:38-131) and (b) tables dumped from the UNMODIFIED
 * reference built by oracle/Makefile (`make ref`, harness oracle/ref_dump.cpp),
 * committed under tests/golden/ together with the generating script
 * tests/golden/make_golden.py.
 *
 * Numeric functions exist twice: suffix _f32 follows the reference's float32
 * arithmetic operation by operation (bit-exact against Oracle A), suffix _f64
 * is the same code with every float widened to double (matches Oracle B, the
 * `#define float double` build).  Counters are 64-bit clean (the reference's
 * are 32-bit, src/base_pattern.cpp:335-336).
 *
 * Encodings (src/base_pattern.h:24-29, src/shared/Alphabet.cpp:33-41,
 * src/iupac_pattern.h:26-30): sequence codes 0 = other, 1..4 = ACGT; base id =
 * sum c_p 4^p (position 0 least significant, ACGT = 0..3); IUPAC id = sum c_p 11^p
 * with A,C,G,T,S,W,R,Y,M,K,N = 0..10.
 */
#ifndef PENG_ORACLE_H_
#define PENG_ORACLE_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* ---- ingest: SequenceSet::readFASTA (src/shared/SequenceSet.cpp:285-447) ---- */
/* returns 0 ok, 1 = reference would exit(1) (space in sequence / sequence before header / cannot open) */
int po_fasta_scan(const char* path, uint64_t* n_seq, uint64_t* total_len, uint32_t* min_len, uint32_t* max_len);
int po_fasta_read(const char* path, uint8_t* codes, uint64_t* offsets /* n_seq+1 */);

/* ---- id helpers (src/base_pattern.cpp:119-144, src/base_pattern.h:88-107) ---- */
uint64_t po_revcomp(uint64_t id, int W);
uint64_t po_bg_id(uint64_t pattern, int pattern_length, int k);

/* ---- counting: BasePattern::count_patterns / _single_strand (src/base_pattern.cpp:331-441) ---- */
void po_count(const uint8_t* codes, const uint64_t* offsets, uint64_t n_seq, int W, int both_strands,
              uint64_t* counts /* 4^W, zeroed by callee */, uint64_t* ltot);

/* ---- IUPAC expansion (src/iupac_pattern.cpp:331-408) ---- */
uint64_t po_iupac_expansion_size(uint64_t iupac_id, int W);
/* writes the list exactly as the reference builds it (BOTH: canonicalised + sorted, duplicates kept) */
uint64_t po_iupac_expand(uint64_t iupac_id, int W, int both_strands, uint64_t* out, uint64_t cap);
/* src/iupac_pattern.cpp:806-833 */
uint64_t po_count_combined(uint64_t iupac_id, int W, int both_strands, const uint64_t* counts);
/* Peng::optimize_iupac_patterns neighbourhood (src/iupac_alphabet.cpp:47-136): writes <=10 letters, returns n */
int po_similar_letters(int c, int* out);

#define PO_DECL(SUF, REAL)                                                                                      \
  /* BackgroundModel ctor count loop + calculateV (src/shared/BackgroundModel.cpp:60-84, 490-530,               \
     src/shared/Sequence.cpp:28-33); v_out = v[0] (4) | v[1] (16) | ... | v[K] (4^(K+1)) */                     \
  void po_background_model##SUF(const uint8_t* codes, const uint64_t* offsets, uint64_t n_seq, int K,           \
                                const REAL* alpha, int interpolate, REAL* v_out, int64_t* n_out);               \
  /* calculate_bg_probabilities (src/base_pattern.cpp:285-325) for one order */                                 \
  void po_bg_probabilities##SUF(const REAL* v, int W, int order, REAL* out);                                    \
  /* aggregate_double_strand_background (src/base_pattern.cpp:268-283), in place */                             \
  void po_aggregate_double_strand##SUF(REAL* bgprob, int W);                                                    \
  /* calculate_expected_counts / _zscores / _log_pvalues (src/base_pattern.cpp:231-265) */                      \
  void po_stats##SUF(const uint64_t* counts, const REAL* bgprob_k, uint64_t ltot, int W, REAL* expected,        \
                     REAL* z, REAL* logp);                                                                      \
  /* select_base_patterns (src/base_pattern.cpp:443-515), std::sort tie order included */                       \
  uint64_t po_select##SUF(const REAL* z, const uint64_t* counts, int W, REAL z_threshold,                       \
                          uint64_t count_threshold, int single_stranded, int filter_neighbors, uint64_t* out,  \
                          uint64_t cap);                                                                        \
  /* BasePattern::getMutualInformationScore (src/base_pattern.cpp:184-200) */                                   \
  REAL po_base_mi_score##SUF(uint64_t count, REAL expected, uint64_t n_sequences);                              \
  /* IUPACPattern::aggregate_attributes_from_basepatterns (src/iupac_pattern.cpp:410-473)                       \
     out[0..5] = n_sites, bg_p, expected, zscore, log_pvalue, MI score (src/iupac_pattern.cpp:652-669) */       \
  void po_iupac_aggregate##SUF(uint64_t iupac_id, int W, int both_strands, const uint64_t* counts,              \
                               const REAL* bgprob_k, const REAL* expected, uint64_t n_sequences,                \
                               uint64_t* n_sites, REAL* out5);                                                  \
  /* IUPACPattern::calculate_adv_pwm (src/iupac_pattern.cpp:505-536); pwm = W x 4 */                            \
  void po_adv_pwm##SUF(uint64_t iupac_id, int W, int both_strands, const uint64_t* counts, int pseudo_counts,  \
                       const REAL* v0, REAL* pwm);                                                              \
  /* Peng::calculate_prob_odds (src/peng.cpp:180-197) */                                                        \
  void po_prob_odds##SUF(const REAL* pwm, int W, const REAL* bgprob, REAL* odds);                               \
  /* Peng::em_optimize_pwms for ONE pwm (src/peng.cpp:84-146) incl. the final renormalisation of              \
     IUPACPattern(ori, pwm) (src/iupac_pattern.cpp:44-72); returns iterations run */                            \
  int po_em##SUF(const REAL* pwm_in, int W, const uint64_t* counts, const REAL* bgprob, REAL saturation,        \
                 REAL min_threshold, int max_iterations, REAL* pwm_out);                                        \
  /* one accumulation pass only (src/peng.cpp:111-129): new_pwm (normalised) from old_pwm */                    \
  void po_em_iteration##SUF(const REAL* old_pwm, int W, const uint64_t* counts, const REAL* bgprob,             \
                            REAL saturation, REAL* new_pwm, REAL* scratch_odds);

PO_DECL(_f32, float)
PO_DECL(_f64, double)

#ifdef __cplusplus
}
#endif
#endif
