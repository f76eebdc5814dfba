// peng_host.h — host driver of the B200 hot path: the `peng_motif` command line on top of libpeng_cuda.so.
//
// What stays on the host is what the reference runs once per *motif* (a few hundred small objects), not
// once per *base* or per *k-mer id*: the hill-climb bookkeeping of Peng::optimize_iupac_patterns
// (src/peng.cpp:437-541), the IUPAC filter (:543-600), PWM similarity / merging (src/peng.cpp:199-313,
// src/iupac_pattern.cpp:75-172,539-615) and the MEME / JSON writers (src/peng.cpp:602-727).  Everything that
// touches the sequence stream or a 4^W table goes through the C ABI of include/peng_cuda.h and runs on the GPU.
//
// The observable contract — command line options, console report, MEME / JSON files — is kept byte for byte;
// the float32 arithmetic of every printed number follows the reference operation by operation (each function
// cites the lines it mirrors).  There is no CPU fallback: without a CUDA device the driver stops with the
// library's error message.
#pragma once
#include <array>
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/peng_cuda.h"

namespace peng_host {

// ---- command line (src/Global.cpp:18-58 defaults, :81-306 parser) ---------------------------------------
enum class Score { LogPval = 0, ExpCounts = 1, MutualInformation = 2 };  // OPTIMIZATION_SCORE, src/Global.h:17-21

struct Options {
  std::string input_fasta;
  std::string background_fasta;  // empty: the input set is its own background
  std::string meme_path, json_path;
  Score score = Score::MutualInformation;
  float enrich_pseudocount_factor = 0.005f;
  int pattern_length = 10;
  int strand = PENG_STRAND_BOTH;
  bool use_em = true;
  float em_saturation = 1E4f;
  float em_threshold = 0.08f;
  int em_max_iterations = 10;
  bool use_merging = true;
  int pseudo_counts = 10;
  bool adv_pwm = true;
  float zscore_threshold = 10.f;
  size_t count_threshold = 3;
  float merge_bit_factor = 0.4f;
  size_t max_merged_length = 14;
  bool interpolate_bg = true;
  int bg_model_order = 2;
  int max_opt_bg_model_order = 2;
  int verbosity = 2;
  int threads = 1;
  bool filter_neighbors = true;
  unsigned minimum_processed_motifs = 0;
  int max_optimized_patterns = 50;
  // additions of this driver (not options of the reference; all default to the reference's behaviour)
  int device = 0;          // --device N
  bool timings = false;    // --timings: per-phase wall clock on stderr
  int gpus = 1;            // --gpus N: ingest + counting sharded over the devices [device, device + N)
};

// Parses argv like Global::readArguments; prints help / version and exits where the reference does.
Options parse_command_line(int argc, char** argv);
void print_help();
extern const char* const kVersion;  // "1.0.0", src/Global.h:23

// ---- IUPAC alphabet (src/iupac_alphabet.cpp, src/iupac_pattern.h:26-30) --------------------------------
constexpr int kIupacLetters = 11;  // A C G T S W R Y M K N
constexpr int kLetterN = 10;
char iupac_char(int c);
const std::vector<int>& similar_letters(int c);         // hill-climb neighbourhood, src/iupac_alphabet.cpp:47-136
const std::vector<int>& representative_letters(int c);  // src/iupac_alphabet.cpp:138-180
uint64_t iupac_factor(int p);                           // 11^p
int iupac_letter_at(uint64_t id, int p);
std::string iupac_to_string(uint64_t id, int W);
uint64_t base_id_to_iupac_id(uint64_t base_id, int W);
std::string base_to_string(uint64_t base_id, int W);

// ---- motif = the reference's IUPACPattern object -----------------------------------------------------
typedef std::array<float, 4> PwmRow;
typedef std::vector<PwmRow> Pwm;

struct Motif {
  uint64_t pattern = 0;  // IUPAC id (meaningless for merged motifs, like in the reference)
  int length = 0;
  float log_pvalue = 0, zscore = 0, bg_p = 0, expected = 0;
  int opt_bg_order = 0;
  uint64_t n_sites = 0;
  std::vector<uint64_t> local_sites;
  Pwm pwm, comp_pwm;  // empty until calculated
  bool merged = false;
};

// Static per-run tables of IUPACPattern::init (src/iupac_pattern.cpp:193-240): log Bonferroni terms and the
// 11 IUPAC letter profiles mixed with the order-0 background.
void init_motif_tables(const float* v0);

// attributes from the device sums (src/iupac_pattern.cpp:443-470)
void set_aggregated_attributes(Motif& m, uint64_t n_sites, float bg_p, float expected);
float mutual_information_score(float observed, float expected, unsigned n_sequences);  // src/utils.h:29-37 + :652-669
float motif_score(const Motif& m, Score s, size_t pseudo_expected, unsigned n_sequences);

void normalize_pwm(Pwm& pwm);                       // src/iupac_pattern.cpp:291-303
void make_comp_pwm(Motif& m);                       // :617-632
std::string pwm_string(const Pwm& pwm);             // get_pattern_string, :694-712
float pwm_information(const Pwm& pwm);              // calculate_pwm_info, src/utils.h:52-63
void no_zero_pwm(Pwm& pwm, unsigned precision);     // src/utils.h:39-50
float similarity(const Pwm& a, const Pwm& b, const float* bg, int off_a, int off_b, int l);  // calculate_s, :563-566
struct BestShift { float score; int shift; bool comp; };
BestShift best_alignment(const Motif& a, const Motif& b, int strand, const float* bg);       // calculate_S, :568-614
Motif merge_motifs(const Motif& longer, const Motif& shorter, bool comp, const float* bg, int shift);  // :75-172
Motif with_pwm(const Motif& ori, const Pwm& pwm);   // IUPACPattern(ori, pwm), :44-72
bool by_log_pvalue(const Motif* a, const Motif* b);  // sort_IUPAC_patterns, :847-849

// ---- pipeline (src/main.cpp, Peng::process) -------------------------------------------------------------
int run(const Options& opt);

// writers (src/peng.cpp:602-727); sort the list by log p-value first, like the reference
void write_meme(std::vector<Motif*>& motifs, const std::string& path, const float* v0);
void write_json(std::vector<Motif*>& motifs, const std::string& path, const float* v0);

}  // namespace peng_host
