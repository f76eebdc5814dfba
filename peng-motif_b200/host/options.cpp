// options.cpp — the peng_motif command line (Global::readArguments / printHelp, src/Global.cpp:81-374).
// Same flags, same defaults, same exit codes: 0 for -h / -version / --version, -1 without arguments, 4 for a flag
// that lacks its value or has an unknown one; unknown flags are reported on stderr and skipped.
#include <sys/time.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <functional>
#include <iostream>
#include <string>

#include "peng_host.h"

namespace peng_host {

const char* const kVersion = "1.0.0";

namespace {

// stderr line in the layout of the reference's LOG() macro (src/log.h:41-54,116-130): "- HH:MM:SS.mmm LEVEL: text"
void log_line(const char* level, const std::string& text) {
  char clock[16];
  time_t t;
  time(&t);
  tm r = {};
  strftime(clock, sizeof clock, "%X", localtime_r(&t, &r));
  timeval tv;
  gettimeofday(&tv, nullptr);
  fprintf(stderr, "- %s.%03ld %s: %s\n\n", clock, (long)tv.tv_usec / 1000, level, text.c_str());
  fflush(stderr);
}

struct Flag {
  const char* name;
  bool takes_value;
  const char* missing_name;  // what the "No expression following ..." message names (the reference's typos included)
  std::function<bool(Options&, const char*)> apply;  // false: unknown value
};

}  // namespace

void print_help() {
  // one entry per option, in the reference's order
  static const char* const entries[][2] = {
      {"-o, <OUTPUT_FILE>", "best IUPAC motives will be written in OUTPUT_FILE\n           in minimal MEME format"},
      {"-j, <OUTPUT_FILE>", "best IUPAC motives will be written in OUTPUT_FILE\n           in JSON format"},
      {"--background-sequences, <FASTA_FILE>",
       "file with fasta sequences to be used for the\n           background model calculation"},
      {"-t, <ZSCORE_THRESHOLD>", "lower zscore threshold for basic patterns"},
      {"-w, <PATTERN_LENGTH>", "length of patterns to be searched"},
      {"--bg-model-order, <BG_MODEL_ORDER>", "order of the background model"},
      {"--count-threshold, <COUNT_THRESHOLD>", "lower threshold for counts of basic patterns"},
      {"--strand, <PLUS|BOTH>", "select the strands to work on"},
      {"--optimization_score, <ENRICHMENT|LOGPVAL|MUTUAL_INFO>", "select the iupac optimization score"},
      {"--enrich_pseudocount_factor, <PSEUDO_COUNTS>",
       "add (enrich_pseudocount_factor x #seqs) pseudocounts\n           in the EXPCOUNTS optimization"},
      {"-b, <BIT_FACTOR_THRESHOLD>", "bit factor threshold for merging IUPAC patterns"},
      {"--no-em", "shuts off the em optimization "},
      {"-a, <EM_SATURATION_THRESHOLD>", "saturation factor for em optimization "},
      {"--em-threshold, <EM_THRESHOLD>", "threshold for finishing the em optimization "},
      {"--em-max-iterations, <EM_MAX_ITERATIONS>", "max number of em optimization iterations"},
      {"--no-merging", "shuts off the merging "},
      {"--max_merged_length", "define the maximum length of motifs after merging"},
      {"--use-default-pwm", "use the default calculation of the pwm"},
      {"--pseudo-counts, <PSEUDO_COUNTS>", "number of pseudo-counts for optimization"},
      {"--threads, <NUMBER_THREADS>", "number of threads to be used for parallelization"},
      {"--no-neighbor-filtering", "do not filter similar base patterns before running the optimization"},
      {"--minimum-processed-patterns <NUMBER_PATTERNS>",
       "minimum number of iupac patterns that are selected for em optimization"},
      {"--version", "print the version number"},
      {"-h", "print this help "},
      {"--max-optimized-patterns", "maximum number of iupac patterns that are selected for pattern optimization"},
  };
  const char* rule = "\n=================================================================\n";
  fputs(rule, stdout);
  fputs("\n Usage: peng_motif SEQFILE [options] \n\n", stdout);
  fputs("\t SEQFILE: file with sequences in FASTA format. \n", stdout);
  for (const auto& e : entries) printf("\n      %s\n           %s\n", e[0], e[1]);
  fputs(rule, stdout);
}

Options parse_command_line(int argc, char** argv) {
  Options o;
  if (argc > 1 && !strcmp(argv[1], "-h")) {
    print_help();
    exit(0);
  }
  if (argc > 1 && !strcmp(argv[1], "-version")) {
    std::cout << "peng_motif version " << kVersion << std::endl;
    exit(0);
  }
  if (argc < 2) {
    fprintf(stderr, "Error: Arguments are missing! \n");
    print_help();
    exit(-1);
  }
  o.input_fasta = argv[1];

  auto as_int = [](int Options::*field) {
    return [field](Options& opt, const char* v) { opt.*field = std::stoi(v); return true; };
  };
  auto as_float = [](float Options::*field) {
    return [field](Options& opt, const char* v) { opt.*field = std::stof(v); return true; };
  };
  auto set_flag = [](bool Options::*field, bool value) {
    return [field, value](Options& opt, const char*) { opt.*field = value; return true; };
  };
  const Flag flags[] = {
      {"-w", true, "-w", [](Options& opt, const char* v) {
         opt.pattern_length = std::stoi(v);
         if (opt.pattern_length % 2 == 1) {
           log_line("ERROR", "Due to optimizations the pattern length has to be a multiple of 2");
           exit(4);
         }
         return true;
       }},
      {"--background-sequences", true, "--background-sequences",
       [](Options& opt, const char* v) { opt.background_fasta = v; return true; }},
      {"--optimization_score", true, "--optimization_score", [](Options& opt, const char* v) {
         if (!strcmp(v, "LOGPVAL")) opt.score = Score::LogPval;
         else if (!strcmp(v, "ENRICHMENT")) opt.score = Score::ExpCounts;
         else if (!strcmp(v, "MUTUAL_INFO")) opt.score = Score::MutualInformation;
         else return false;
         return true;
       }},
      {"--enrich_pseudocount_factor", true, "--enrich_pseudocount_factor", as_float(&Options::enrich_pseudocount_factor)},
      {"-v", true, "-v", as_int(&Options::verbosity)},
      {"-o", true, "-o", [](Options& opt, const char* v) { opt.meme_path = v; return true; }},
      {"-j", true, "-j", [](Options& opt, const char* v) { opt.json_path = v; return true; }},
      {"-t", true, "-t", as_float(&Options::zscore_threshold)},
      {"--count-threshold", true, "--count-threshold",
       [](Options& opt, const char* v) { opt.count_threshold = (size_t)std::stoi(v); return true; }},
      {"-b", true, "-b", as_float(&Options::merge_bit_factor)},
      {"--use-default-pwm", false, "", set_flag(&Options::adv_pwm, false)},
      {"--pseudo-counts", true, "--pseudo-counts", as_int(&Options::pseudo_counts)},
      {"--threads", true, "-t", as_int(&Options::threads)},
      {"--no-em", false, "", set_flag(&Options::use_em, false)},
      {"-a", true, "-a", as_float(&Options::em_saturation)},
      {"--em-threshold", true, "--em-threshold", as_float(&Options::em_threshold)},
      {"--em-max-iterations", true, "--em-max-iterations", as_int(&Options::em_max_iterations)},
      {"--no-merging", false, "", set_flag(&Options::use_merging, false)},
      {"--max_merged_length", true, "--max_merged_length",
       [](Options& opt, const char* v) { opt.max_merged_length = (size_t)std::stoi(v); return true; }},
      {"--strand", true, "--strand", [](Options& opt, const char* v) {
         if (!strcmp(v, "BOTH")) opt.strand = PENG_STRAND_BOTH;
         else if (!strcmp(v, "PLUS")) opt.strand = PENG_STRAND_PLUS;
         else return false;
         return true;
       }},
      {"--bg-model-order", true, "--bg-model-order", as_int(&Options::bg_model_order)},
      {"--no-neighbor-filtering", false, "", set_flag(&Options::filter_neighbors, false)},
      {"--minimum-processed-patterns", true, "--minimum-processed-patterns",
       [](Options& opt, const char* v) { opt.minimum_processed_motifs = (unsigned)std::stoi(v); return true; }},
      {"--max-optimized-patterns", true, "--max-optimized-patterns", as_int(&Options::max_optimized_patterns)},
      // this driver only
      {"--device", true, "--device", as_int(&Options::device)},
      {"--timings", false, "", set_flag(&Options::timings, true)},
      {"--gpus", true, "--gpus", as_int(&Options::gpus)},
  };

  for (int i = 2; i < argc; i++) {
    const char* a = argv[i];
    if (!strcmp(a, "--version")) {
      std::cout << "peng_motif " << kVersion << std::endl;
      exit(0);
    }
    if (!strcmp(a, "-h")) {
      print_help();
      exit(0);
    }
    const Flag* hit = nullptr;
    for (const Flag& f : flags)
      if (!strcmp(a, f.name)) hit = &f;
    if (!hit) {
      log_line("WARNING", std::string("Ignoring unknown option ") + a);
      continue;
    }
    const char* value = nullptr;
    if (hit->takes_value) {
      if (++i >= argc) {
        print_help();
        log_line("ERROR", std::string("No expression following ") + hit->missing_name);
        exit(4);
      }
      value = argv[i];
    }
    if (!hit->apply(o, value)) {
      print_help();
      log_line("ERROR", std::string("Unknown expression following ") + hit->name);
      exit(4);
    }
  }
  return o;
}

}  // namespace peng_host
