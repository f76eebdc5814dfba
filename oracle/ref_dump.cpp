// ref_dump — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A small harness of OUR OWN that is compiled against the UNMODIFIED reference
// sources where they lie under /root/reference (see oracle/Makefile) and dumps
// the hot-path tables of the reference itself as raw little-endian files, so
// that the CPU restatement (oracle/peng_oracle.c) and the CUDA path can be
// pinned against outputs of the real reference.
//
// It follows the call order of Peng::process (/root/reference/src/peng.cpp:322-435)
// but stops after each stage to write what the reference computed:
//   meta.txt            n_seq, min/max length, W, strand, ltot, n_patterns, k, max_k
//   v.bin               BackgroundModel::getV()[0..K]   (src/shared/BackgroundModel.cpp:490-530)
//   counts.u64          BasePattern::pattern_counter    (src/base_pattern.cpp:331-441)
//   bgprob_k<k>.bin     BasePattern::pattern_bg_probabilities[k] (src/base_pattern.cpp:268-325)
//   expected.bin, z.bin, logp.bin                       (src/base_pattern.cpp:231-265)
//   seeds.u64           select_base_patterns output     (src/base_pattern.cpp:443-515)
//   iupac.txt           every IUPAC pattern the hill-climb evaluated with its sums
//   adv_pwm.bin / em_pwm.bin   PWMs before / after Peng::em_optimize_pwms (src/peng.cpp:48-178)
//
// REAL is `float` for the as-is build and `double` for the FP64 variant
// (oracle/Makefile builds the latter with a `#define float double` prelude).
// Compiled with -fno-access-control so private members can be read.

#include "Global.h"
#include "base_pattern.h"
#include "iupac_alphabet.h"
#include "iupac_pattern.h"
#include "peng.h"
#include "utils.h"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

typedef float REAL;  // becomes double under the FP64 prelude

static void write_raw(const std::string& path, const void* p, size_t bytes) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) { perror(path.c_str()); exit(2); }
  if (bytes) fwrite(p, 1, bytes, f);
  fclose(f);
}

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int main(int argc, char** argv) {
  if (argc < 5) {
    fprintf(stderr,
            "usage: ref_dump FASTA W BOTH|PLUS OUTDIR [--no-em] [--tables-only] [--counts-only]\n"
            "               [--iupac-file F] [--max-seeds N] [-t Z] [--count-threshold C]\n"
            "               [--em-threshold T] [--em-iterations N]   (bench: T = 0 makes every PWM run all N iterations)\n");
    return 2;
  }
  std::string fasta = argv[1];
  int W = atoi(argv[2]);
  Strand strand = strcmp(argv[3], "PLUS") == 0 ? Strand::PLUS_STRAND : Strand::BOTH_STRANDS;
  std::string out = argv[4];
  bool use_em = true, tables_only = false, counts_only = false;
  std::string iupac_file;
  size_t max_seeds = 50;
  REAL zthr = 10;
  size_t cthr = 3;
  REAL em_thr = 0.08;
  int em_its = 10;
  for (int i = 5; i < argc; i++) {
    if (!strcmp(argv[i], "--no-em")) use_em = false;
    else if (!strcmp(argv[i], "--tables-only")) tables_only = true;
    else if (!strcmp(argv[i], "--counts-only")) counts_only = true;
    else if (!strcmp(argv[i], "--iupac-file") && i + 1 < argc) iupac_file = argv[++i];
    else if (!strcmp(argv[i], "--max-seeds") && i + 1 < argc) max_seeds = atol(argv[++i]);
    else if (!strcmp(argv[i], "-t") && i + 1 < argc) zthr = atof(argv[++i]);
    else if (!strcmp(argv[i], "--count-threshold") && i + 1 < argc) cthr = atol(argv[++i]);
    else if (!strcmp(argv[i], "--em-threshold") && i + 1 < argc) em_thr = atof(argv[++i]);
    else if (!strcmp(argv[i], "--em-iterations") && i + 1 < argc) em_its = atoi(argv[++i]);
  }

  Alphabet::init("STANDARD");
  double t0 = now_s();
  SequenceSet* seqs = new SequenceSet(fasta, true);
  double t_read = now_s() - t0;
  int K = 2;
  t0 = now_s();
  BackgroundModel* bg = new BackgroundModel(*seqs, K, Global::bgModelAlpha, Global::interpolateBG);
  double t_bg = now_s() - t0;

  // Peng ctor does IUPACAlphabet::init + IUPACPattern::init (src/peng.cpp:22-41)
  Peng peng(strand, K, K, seqs, bg);

  int k = std::min(W - 1, K), max_k = std::min(W - 1, K);
  t0 = now_s();
  BasePattern* bp = new BasePattern(W, strand, k, max_k, seqs, bg);
  double t_bp = now_s() - t0;
  size_t n = bp->getNumberPatterns();
  // the hot loop alone: run count_patterns once more on the finished object (private -> -fno-access-control);
  // it adds to pattern_counter, so take a copy of the table first and put it back afterwards
  double t_count = 0;
  {
    std::vector<size_t> keep(bp->getPatternCounter(), bp->getPatternCounter() + n);
    size_t keep_ltot = bp->ltot;
    t0 = now_s();
    if (strand == Strand::BOTH_STRANDS) bp->count_patterns(seqs); else bp->count_patterns_single_strand(seqs);
    t_count = now_s() - t0;
    std::copy(keep.begin(), keep.end(), bp->getPatternCounter());
    bp->ltot = keep_ltot;
  }

  {
    FILE* f = fopen((out + "/meta.txt").c_str(), "w");
    fprintf(f, "n_seq %zu\nmin_len %u\nmax_len %u\nW %d\nstrand %s\nltot %zu\nn_patterns %zu\nk %d\nmax_k %d\nreal_bytes %zu\n",
            seqs->getN(), seqs->getMinL(), seqs->getMaxL(), W, argv[3], bp->getLtot(), n, k, bp->max_k,
            sizeof(REAL));
    fprintf(f, "t_read_fasta_s %.6f\nt_background_model_s %.6f\nt_base_pattern_s %.6f\nt_count_patterns_s %.6f\n", t_read,
            t_bg, t_bp, t_count);
    fclose(f);
  }
  {
    std::vector<REAL> v;
    size_t sz = 4;
    for (int o = 0; o <= K; o++, sz *= 4)
      for (size_t y = 0; y < sz; y++) v.push_back(bg->getV()[o][y]);
    write_raw(out + "/v.bin", v.data(), v.size() * sizeof(REAL));
  }
  write_raw(out + "/counts.u64", bp->getPatternCounter(), n * sizeof(size_t));
  if (counts_only) return 0;
  for (int o = 0; o <= bp->max_k; o++)
    write_raw(out + "/bgprob_k" + std::to_string(o) + ".bin", bp->getBackgroundProb(o), n * sizeof(REAL));
  write_raw(out + "/expected.bin", bp->getExpectedCounts(), n * sizeof(REAL));
  write_raw(out + "/z.bin", bp->pattern_zscore, n * sizeof(REAL));
  write_raw(out + "/logp.bin", bp->pattern_logp, n * sizeof(REAL));

  t0 = now_s();
  std::vector<size_t> seeds = bp->select_base_patterns(zthr, cthr, strand == Strand::PLUS_STRAND, true);
  double t_sel = now_s() - t0;
  write_raw(out + "/seeds.u64", seeds.data(), seeds.size() * sizeof(size_t));

  // optional: aggregate an explicit list of IUPAC ids (one decimal id per line)
  if (!iupac_file.empty()) {
    FILE* fi = fopen(iupac_file.c_str(), "r");
    FILE* fo = fopen((out + "/iupac_query.txt").c_str(), "w");
    unsigned long long id;
    while (fi && fscanf(fi, "%llu", &id) == 1) {
      IUPACPattern p((size_t)id, W);
      p.aggregate_attributes_from_basepatterns(bp);
      unsigned long cc = p.count_combined_occurences(bp, (size_t)id);
      fprintf(fo, "%llu %zu %.17g %.17g %.17g %.17g %.17g %lu\n", id, p.get_sites(), (double)p.get_bg_p(),
              (double)p.getExpectedCounts(), (double)p.getZscore(), (double)p.getLogPval(),
              (double)p.getMutualInformationScore(seqs->getN()), cc);
    }
    if (fi) fclose(fi);
    fclose(fo);
  }
  if (tables_only) return 0;

  if (seeds.size() > max_seeds) seeds.resize(max_seeds);
  std::vector<IUPACPattern*> unopt;
  t0 = now_s();
  peng.optimize_iupac_patterns(OPTIMIZATION_SCORE::MutualInformation, bp, seeds, unopt, 0.005);
  double t_opt = now_s() - t0;
  peng.filter_iupac_patterns(W, 0, unopt);
  {
    FILE* f = fopen((out + "/iupac.txt").c_str(), "w");
    for (auto p : unopt)
      fprintf(f, "%zu %s %zu %.17g %.17g %.17g %.17g\n", p->get_pattern(),
              IUPACPattern::toString(p->get_pattern(), W).c_str(), p->get_sites(), (double)p->get_bg_p(),
              (double)p->getExpectedCounts(), (double)p->getZscore(), (double)p->getLogPval());
    fclose(f);
  }
  std::vector<REAL> adv;
  for (auto p : unopt) {
    p->calculate_adv_pwm(bp, 10, bp->getPatternCounter(), bg->getV()[0]);
    for (int q = 0; q < W; q++)
      for (int a = 0; a < 4; a++) adv.push_back(p->get_pwm()[q][a]);
  }
  write_raw(out + "/adv_pwm.bin", adv.data(), adv.size() * sizeof(REAL));

  double t_em = 0;
  if (use_em) {
    std::vector<IUPACPattern*> opt;
    t0 = now_s();
    peng.em_optimize_pwms(unopt, bp, 1e4, em_thr, em_its, bp->getBackgroundProb(bp->max_k), opt);
    t_em = now_s() - t0;
    std::vector<REAL> em;
    for (auto p : opt)
      for (int q = 0; q < W; q++)
        for (int a = 0; a < 4; a++) em.push_back(p->get_pwm()[q][a]);
    write_raw(out + "/em_pwm.bin", em.data(), em.size() * sizeof(REAL));
  }
  {
    FILE* f = fopen((out + "/meta.txt").c_str(), "a");
    fprintf(f, "t_select_s %.6f\nt_optimize_iupac_s %.6f\nt_em_s %.6f\nn_em_pwms %zu\nem_threshold %g\nem_max_iterations %d\n", t_sel,
            t_opt, t_em, unopt.size(), (double)em_thr, em_its);
    fclose(f);
  }
  return 0;
}
